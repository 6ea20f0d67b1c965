"""Sustained forward throughput against a chosen build of the library (same-box A/B of kernel changes):
python tools/ab_forward.py --lib tools/_ab/lib_old.so --rows 1024 --reps 12  ->  ms per `rows` dataset rows (median)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from s2p_b200 import _lib

ap = argparse.ArgumentParser()
ap.add_argument("--lib", type=str, default=None)
ap.add_argument("--rows", type=int, default=1024)
ap.add_argument("--reps", type=int, default=12)
ap.add_argument("--precision", type=str, default="high")
ap.add_argument("--mb", type=int, default=0, help="rows per launch set (0: the default of options.py)")
a = ap.parse_args()
if a.lib:
    import ctypes
    _lib.LIB_PATH = os.path.abspath(a.lib)
    probe = ctypes.CDLL(_lib.LIB_PATH)          # an older build may lack entry points added since: drop their signatures
    for name in [n for n in _lib.SIGNATURES if not hasattr(probe, n)]:
        del _lib.SIGNATURES[name]
import s2p_b200  # noqa: E402

torch.manual_seed(1234)
kw = {"max_microbatch": a.mb} if a.mb else {}
net = s2p_b200.define_G(s2p_b200.make_options(env_type="cheetah", gpu_ids="0", init_type="scaled", precision=a.precision, **kw)).eval()
net.pack()
g = torch.Generator().manual_seed(0)
frames = torch.randint(0, 256, (a.rows, 100, 100, 3), dtype=torch.uint8, generator=g).cuda()
states = torch.randn(a.rows, net.opt.state_dim, generator=g).cuda()
for _ in range(3):
    out = net.forward_u8(frames, states)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.reps + 1)]
ev[0].record()
for i in range(a.reps):
    out = net.forward_u8(frames, states)
    ev[i + 1].record()
torch.cuda.synchronize()
ms = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(a.reps))
print(f"ab_forward lib={a.lib or 'product'} mb={net.max_microbatch} rows={a.rows} ms median={ms[len(ms) // 2]:.3f} min={ms[0]:.3f} "
      f"frames/s={a.rows / ms[len(ms) // 2] * 1e3:.0f} checksum={int(out.sum().item())}")
