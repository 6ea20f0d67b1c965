"""RAD augmentation kernels (csrc/rad_augs.cu) against the HBM roofline, with the reference's own expression timed
beside them: torch eager on the same GPU for the tensor ops (flip / rotation / grayscale as written in
data_augs.py), the oracle's per-image numpy loop on the host for the array ops (bounded sample).
Batch x 9 x 100 x 100 (three stacked frames).  usage: python tools/rad_augs_bench.py [batch]"""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import data_augs_oracle as DO  # noqa: E402  (cpu baseline leg only)
import s2p_b200.data_augs as R  # noqa: E402
from s2p_b200 import _lib  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
peak = json.load(open(pk)).get("hbm_gbs", 6552.0) if os.path.exists(pk) else 6552.0


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def eager_flip(images, mask):        # data_augs.py:131-145 on the device
    m = mask[:, None, None, None].type(images.dtype)
    return m * images.flip([3]) + (1 - m) * images


def eager_rot(images, turns):        # data_augs.py:164-186 on the device
    rots = [images, images.rot90(1, [2, 3]), images.rot90(2, [2, 3]), images.rot90(3, [2, 3])]
    return sum((turns == k)[:, None, None, None].type(images.dtype) * rots[k] for k in range(4))


res = {}
rng = np.random.RandomState(0)
for dtype, eb in ((torch.uint8, 1), (torch.float32, 4)):
    x = (torch.randint(0, 256, (B, 9, 100, 100), device="cuda", dtype=torch.uint8) if eb == 1
         else torch.rand(B, 9, 100, 100, device="cuda"))
    draws = {k: torch.from_numpy(rng.randint(lo, hi, B).astype(np.int32)).cuda() for k, (lo, hi) in
             dict(crop=(0, 17), cut=(10, 30), tr=(0, 9), flag=(0, 2), turns=(0, 4)).items()}
    lib, st = _lib.load(), torch.cuda.current_stream().cuda_stream

    def run(op, ho, wo, p0, p1):
        out = torch.empty(B, 9, ho, wo, dtype=dtype, device="cuda")
        return lambda: _lib.check(lib.s2p_rad_augment(x.data_ptr(), out.data_ptr(), eb, B, 9, 100, 100, ho, wo, op, p0.data_ptr(),
                                                      0 if p1 is None else p1.data_ptr(), 0, st), "rad")
    cases = [("crop84", _lib.RAD_CROP, 84, 84, draws["crop"], draws["crop"]), ("translate108", _lib.RAD_TRANSLATE, 108, 108, draws["tr"], draws["tr"]),
             ("cutout", _lib.RAD_CUTOUT, 100, 100, draws["cut"], draws["cut"]), ("flip", _lib.RAD_FLIP, 100, 100, draws["flag"], None),
             ("rot90", _lib.RAD_ROT90, 100, 100, draws["turns"], None)]
    for name, op, ho, wo, p0, p1 in cases:
        ms = timed(run(op, ho, wo, p0, p1))
        read = 84 * 84 if op == _lib.RAD_CROP else 100 * 100
        byts = B * 9 * (read + ho * wo) * eb
        res[f"{name}_{'u8' if eb == 1 else 'f32'}"] = dict(us=round(ms * 1e3, 1), gbs=round(byts / ms / 1e6), frac=round(byts / ms / 1e6 / peak, 3))
    if eb == 4:
        m = draws["flag"]
        out = torch.empty_like(x)
        ms = timed(lambda: _lib.check(lib.s2p_rad_grayscale(x.data_ptr(), out.data_ptr(), B, 3, 100, 100, m.data_ptr(), st), "gray"))
        byts = B * 9 * 10000 * 8
        res["grayscale_f32"] = dict(us=round(ms * 1e3, 1), gbs=round(byts / ms / 1e6), frac=round(byts / ms / 1e6 / peak, 3))
        res["eager_flip_f32_us"] = round(timed(lambda: eager_flip(x, draws["flag"])) * 1e3, 1)
        res["eager_rotation_f32_us"] = round(timed(lambda: eager_rot(x, draws["turns"])) * 1e3, 1)

# host loop of the reference for the array ops: bounded sample of 256 stacks
xs = rng.randint(0, 256, size=(256, 9, 100, 100)).astype(np.uint8)
t0 = time.perf_counter(); DO.random_crop(xs, 84); t1 = time.perf_counter(); DO.random_translate(xs, 108); t2 = time.perf_counter()
res["cpu_numpy_loop_us_per_1024_u8"] = dict(crop84=round((t1 - t0) * 4e6), translate108=round((t2 - t1) * 4e6), sample="256 stacks, 1 core")
res["batch"], res["peak_gbs"] = B, peak
print(json.dumps(res))
