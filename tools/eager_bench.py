"""The practical bar: the SAME oracle generator run by eager PyTorch on the GPU (cuDNN / cuBLAS,
TF32 and bf16 autocast, optionally channels_last), frames/s at a 128-frame batch.  usage: python tools/eager_bench.py [batch]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch

import s2p_oracle as O

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
torch.manual_seed(1234)
net = O.define_G(O.make_options(env_type="cheetah", init_type="scaled")).cuda().eval()
g = torch.Generator().manual_seed(0)
frames = torch.randint(0, 256, (B, 100, 100, 3), dtype=torch.uint8, generator=g).cuda()
states = torch.randn(B, 17, generator=g).cuda()


def run():
    with torch.no_grad():
        return O.to_uint8_hwc(net(O.from_uint8_hwc(frames), states))


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
ms = timed(run)
print(f"eager oracle fp32 (no TF32): {ms:.1f} ms / {B} frames = {B / ms * 1e3:.0f} frames/s")
torch.backends.cudnn.allow_tf32 = True
torch.backends.cuda.matmul.allow_tf32 = True
ms = timed(run)
print(f"eager oracle TF32 cuDNN: {ms:.1f} ms / {B} frames = {B / ms * 1e3:.0f} frames/s")


def amp():
    with torch.autocast("cuda", dtype=torch.bfloat16):
        return run()


ms = timed(amp)
print(f"eager oracle bf16 autocast: {ms:.1f} ms / {B} frames = {B / ms * 1e3:.0f} frames/s")
net = net.to(memory_format=torch.channels_last)
ms = timed(amp)
print(f"eager oracle bf16 autocast, channels_last: {ms:.1f} ms / {B} frames = {B / ms * 1e3:.0f} frames/s")
