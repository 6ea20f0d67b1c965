"""Bring-up probe of s2p_conv_wgrad against torch autograd (not a test)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from s2p_b200 import ops

torch.backends.cudnn.allow_tf32 = False
CASES = [  # b, h, w, cin, cout, k, stride
    (1, 16, 16, 64, 128, 1, 1), (1, 16, 16, 64, 128, 3, 1), (2, 25, 25, 128, 128, 3, 1), (2, 100, 100, 64, 64, 3, 1),
    (2, 50, 50, 128, 256, 3, 1), (3, 100, 100, 64, 128, 3, 2), (2, 25, 25, 512, 512, 3, 1), (2, 50, 50, 256, 128, 1, 1),
]
g = torch.Generator().manual_seed(0)
for b, h, w, cin, cout, k, st in CASES:
    pad = k // 2
    x = torch.randn(b, h, w, cin, generator=g).cuda().to(torch.bfloat16)
    ho, wo = (h + 2 * pad - k) // st + 1, (w + 2 * pad - k) // st + 1
    dy = torch.randn(b, ho, wo, cout, generator=g).cuda().to(torch.bfloat16)
    dw = ops.conv_wgrad(x, dy, k, st)
    torch.cuda.synchronize()
    wt = torch.zeros(cout, cin, k, k, device="cuda", requires_grad=True)
    y = F.conv2d(x.float().permute(0, 3, 1, 2), wt, None, st, pad)
    y.backward(dy.float().permute(0, 3, 1, 2))
    ref = wt.grad
    err = (dw - ref).abs().max().item()
    print(f"b{b} {h}x{w} {cin}->{cout} k{k} s{st}: max_abs_err={err:.4e} ref_absmax={ref.abs().max().item():.3f} "
          f"{'OK' if err <= 1e-3 * ref.abs().max().item() + 1e-3 else 'BAD'}", flush=True)
