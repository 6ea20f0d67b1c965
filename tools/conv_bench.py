"""Times individual conv layers of the generator (CUDA events, L2 flushed between runs).
usage: python tools/conv_bench.py [batch] [path of another build of the library, for a same-box A/B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from s2p_b200 import _lib

if len(sys.argv) > 2:
    import ctypes
    _lib.LIB_PATH = os.path.abspath(sys.argv[2])
    probe = ctypes.CDLL(_lib.LIB_PATH)
    for name in [n for n in _lib.SIGNATURES if not hasattr(probe, n)]:
        del _lib.SIGNATURES[name]
from s2p_b200 import ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
# name, H, Cin, Cout, k, act, residual, stats
LAYERS = [
    ("hs1 64->768 @100 relu", 100, 64, 768, 3, 1, False, False),
    ("hs2 128->384 @50 relu", 50, 128, 384, 3, 1, False, False),
    ("b3.conv0 256->128 @100 +stats", 100, 256, 128, 3, 0, False, True),
    ("b3.conv1 128->128 @100 +res+stats", 100, 128, 128, 3, 0, True, True),
    ("b4.conv0 128->64 @100 +stats", 100, 128, 64, 3, 0, False, True),
    ("b4.conv1 64->64 @100 +res lrelu", 100, 64, 64, 3, 2, True, False),
    ("b3.convs 1x1 256->128 @100", 100, 256, 128, 1, 0, False, False),
    ("b2.conv0 512->256 @50 +stats", 50, 512, 256, 3, 0, False, True),
    ("b0.conv0 512->512 @25 +stats", 25, 512, 512, 3, 0, False, True),
]
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
g = torch.Generator().manual_seed(0)
for name, hw, cin, cout, k, act, res, stats in LAYERS:
    x = torch.randn(B, hw, hw, cin, generator=g).cuda().to(torch.bfloat16)
    w = ops.pack_conv_weight((torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).cuda())
    bias = torch.randn(cout, generator=g).cuda()
    r = torch.randn(B, hw, hw, cout, generator=g).cuda().to(torch.bfloat16) if res else None
    st = torch.empty(B, ops.conv_stats_rows(hw, hw, cout, k, 1), cout, 2, device="cuda") if stats else None
    y = torch.empty(B, hw, hw, cout, dtype=torch.bfloat16, device="cuda")
    ts = []
    for i in range(6):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.conv_igemm(x, w, bias, cout, k, 1, act, r, y, st)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts[1:])[len(ts[1:]) // 2]
    fl = 2.0 * B * hw * hw * cout * cin * k * k
    print(f"{name:36s} {ms * 1e3:8.1f} us  {fl / ms / 1e9:7.1f} TF/s")
