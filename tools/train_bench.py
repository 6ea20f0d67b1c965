"""Times the GAN training step (BASELINE configs[1]/[4]) on one GPU: this repo's S2PTrainer and,
for context, the fp32 oracle through eager PyTorch (cuDNN/cuBLAS).
usage: python tools/train_bench.py [--batch 16] [--steps 5] [--profile] [--no-oracle]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch

from s2p_b200 import _lib

if os.environ.get("S2P_TOOL_LIB"):          # same-box A/B against another build of the library (tools only)
    import ctypes
    _lib.LIB_PATH = os.path.abspath(os.environ["S2P_TOOL_LIB"])
    _probe = ctypes.CDLL(_lib.LIB_PATH)
    for _n in [n for n in _lib.SIGNATURES if not hasattr(_probe, n)]:
        del _lib.SIGNATURES[_n]
import s2p_b200  # noqa: E402
from s2p_b200.training import S2PTrainer

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--env", type=str, default="walker")
ap.add_argument("--profile", action="store_true")
ap.add_argument("--no-oracle", action="store_true")
ap.add_argument("--graph", action="store_true")
ap.add_argument("--no-vgg", action="store_true")
a = ap.parse_args()
B = a.batch
torch.manual_seed(1234)
opt = s2p_b200.make_options(env_type=a.env, gpu_ids="0", init_type="scaled")
nG, nD = s2p_b200.define_G(opt), s2p_b200.define_D(opt)
nV = None if a.no_vgg else s2p_b200.VGG19FeaturesB200().cuda()
tr = S2PTrainer(nG, nD, vgg=nV)
g = torch.Generator().manual_seed(0)
prev = (torch.rand(B, 3, 100, 100, generator=g) * 2 - 1).cuda()
real = (torch.rand(B, 3, 100, 100, generator=g) * 2 - 1).cuda()
state = torch.randn(B, opt.state_dim, generator=g).cuda()


def timed(fn, steps, warm):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


l0 = _lib.launch_count()
ms = timed(lambda: tr.train_step(prev, state, real), a.steps, a.warmup)
launches = (_lib.launch_count() - l0) / (a.steps + a.warmup)
print(f"s2p_b200 train step: batch={B} {ms:.2f} ms/step  {B / ms * 1e3:.1f} frames/s  ({launches:.0f} own kernel launches/step)")
if a.graph:
    tr.capture(prev, state, real)
    l0 = _lib.launch_count()
    ms_g = timed(lambda: tr.train_step_graphed(prev, state, real), a.steps, a.warmup)
    print(f"s2p_b200 train step (CUDA graph replay): batch={B} {ms_g:.2f} ms/step  {B / ms_g * 1e3:.1f} frames/s")
if a.profile:
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True) as prof:
        tr.train_step(prev, state, real)
        torch.cuda.synchronize()
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=70))
    # the torch glue between this repo's kernels, by op and input shape (what is left to fuse away)
    rows = [e for e in prof.key_averages(group_by_input_shape=True) if e.key.startswith("aten::") and e.self_device_time_total > 0]
    rows.sort(key=lambda e: -e.self_device_time_total)
    print("aten ops by self CUDA time (us), calls, input shapes")
    for e in rows[:45]:
        print(f"{e.key:28s} {e.self_device_time_total:9.0f} {e.count:4d}  {str(e.input_shapes)[:150]}")
    print("aten total self CUDA us:", sum(e.self_device_time_total for e in rows))
if not a.no_oracle:
    import s2p_oracle as O
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    oG = O.define_G(O.make_options(env_type=a.env, init_type="scaled")).cuda().train()
    oD = O.define_D(O.make_options(init_type="scaled")).cuda()
    optG = torch.optim.Adam(oG.parameters(), **O.ADAM)
    optD = torch.optim.Adam(oD.parameters(), **O.ADAM)
    oV = None if a.no_vgg else O.VGG19Features().cuda()
    ms_o = timed(lambda: O.training_step(oG, oD, optG, optD, prev, state, real, vgg=oV), max(2, a.steps // 2), 2)
    print(f"oracle eager PyTorch (TF32 cuDNN): {ms_o:.2f} ms/step  {B / ms_o * 1e3:.1f} frames/s")

    def amp_step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            return O.training_step(oG, oD, optG, optD, prev, state, real, vgg=oV)
    try:
        ms_a = timed(amp_step, max(2, a.steps // 2), 2)
        print(f"oracle eager PyTorch (bf16 autocast): {ms_a:.2f} ms/step  {B / ms_a * 1e3:.1f} frames/s")
    except Exception as e:  # noqa: BLE001
        print("bf16 autocast oracle failed:", repr(e)[:200])
