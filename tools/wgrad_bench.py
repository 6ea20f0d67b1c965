"""Times s2p_conv_wgrad on the generator's layer shapes (CUDA events). usage: python tools/wgrad_bench.py [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from s2p_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
LAYERS = [("gb 128->512 @100", 100, 128, 512, 3), ("gb 128->256 @100", 100, 128, 256, 3), ("conv0 256->128 @100", 100, 256, 128, 3),
          ("hs1 64->768 @100", 100, 64, 768, 3), ("gb 128->1024 @50", 50, 128, 1024, 3), ("conv0 512->256 @50", 50, 512, 256, 3),
          ("conv 512->512 @25", 25, 512, 512, 3), ("conv1 64->64 @100", 100, 64, 64, 3), ("1x1 256->128 @100", 100, 256, 128, 1)]
g = torch.Generator().manual_seed(0)
tot = 0.0
for name, hw, cin, cout, k in LAYERS:
    x = torch.randn(B, hw, hw, cin, generator=g).cuda().to(torch.bfloat16)
    dy = torch.randn(B, hw, hw, cout, generator=g).cuda().to(torch.bfloat16)
    ts = []
    for i in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.conv_wgrad(x, dy, k, 1)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts[1:])[len(ts[1:]) // 2]
    tot += ms
    print(f"{name:24s} {ms * 1e3:8.1f} us  {2.0 * B * hw * hw * cout * cin * k * k / ms / 1e9:7.1f} TF/s")
print(f"total {tot:.2f} ms")
