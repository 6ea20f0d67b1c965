"""One generator forward inside a cudaProfilerStart/Stop range, for ncu:

    ncu --profile-from-start off ... python tools/profile_step.py [--batch 128] [--warm 2]

Warm-up forwards run outside the range, so `ncu --profile-from-start off` sees exactly
the kernels of ONE forward of `--batch` dataset rows (uint8 HWC in -> uint8 HWC out).
Prints the launch count of that forward (s2p_launch_count delta).
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import s2p_b200
from s2p_b200 import _lib

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=128)
ap.add_argument("--warm", type=int, default=2)
ap.add_argument("--env", type=str, default="cheetah")
ap.add_argument("--fuse", type=int, default=1)
ap.add_argument("--precision", type=str, default="high")
a = ap.parse_args()

torch.manual_seed(1234)
net = s2p_b200.define_G(s2p_b200.make_options(env_type=a.env, gpu_ids="0", init_type="scaled", fuse_mat=bool(a.fuse), precision=a.precision)).eval()
net.pack()
g = torch.Generator().manual_seed(0)
frames = torch.randint(0, 256, (a.batch, 100, 100, 3), dtype=torch.uint8, generator=g).cuda()
states = torch.randn(a.batch, net.opt.state_dim, generator=g).cuda()
for _ in range(a.warm):
    net.forward_u8(frames, states)
torch.cuda.synchronize()
l0 = _lib.launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()
e0.record()
out = net.forward_u8(frames, states)
e1.record()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"profile_step: batch={a.batch} launches={_lib.launch_count() - l0} ms={e0.elapsed_time(e1):.3f} "
      f"checksum={int(out.sum().item())}")
