"""diagnostic: per-slice gradient error of VGG19FeaturesB200 vs the fp32 oracle, with torch bf16 autocast as yardstick"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import torch
import torch.nn.functional as F
import s2p_b200, s2p_oracle as O
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
ora = O.VGG19Features().cuda()
net = s2p_b200.VGG19FeaturesB200().cuda()
g = torch.Generator().manual_seed(6)
fake = (torch.rand(3, 3, 100, 100, generator=g) * 2 - 1).cuda()
for k in range(5):
    fa, fb, fc = (fake.clone().requires_grad_() for _ in range(3))
    net(fa)[k].float().pow(2).mean().backward()
    ora(fb)[k].pow(2).mean().backward()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        lc = ora(fc)[k].float().pow(2).mean()
    lc.backward()
    rel = lambda a, b: ((a - b).norm() / b.norm()).item()
    cos = lambda a, b: F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item()
    print(f"slice {k}: ours rel {rel(fa.grad, fb.grad):.3e} cos {cos(fa.grad, fb.grad):.5f} | autocast rel {rel(fc.grad, fb.grad):.3e} cos {cos(fc.grad, fb.grad):.5f}")
