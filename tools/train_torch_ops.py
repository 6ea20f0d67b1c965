"""Which torch (aten) ops launch kernels inside one training step, with input shapes and the python line that called
them (torch.profiler).  usage: python tools/train_torch_ops.py [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

import s2p_b200
from s2p_b200.training import S2PTrainer

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
torch.manual_seed(1234)
opt = s2p_b200.make_options(env_type="walker", gpu_ids="0", init_type="scaled")
tr = S2PTrainer(s2p_b200.define_G(opt), s2p_b200.define_D(opt), vgg=s2p_b200.VGG19FeaturesB200().cuda())
g = torch.Generator().manual_seed(0)
prev = (torch.rand(B, 3, 100, 100, generator=g) * 2 - 1).cuda()
real = (torch.rand(B, 3, 100, 100, generator=g) * 2 - 1).cuda()
state = torch.randn(B, opt.state_dim, generator=g).cuda()
for _ in range(2):
    tr.train_step(prev, state, real)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=True) as prof:
    tr.train_step(prev, state, real)
    torch.cuda.synchronize()
rows = []
for e in prof.events():
    t = getattr(e, "self_device_time_total", 0) or getattr(e, "self_cuda_time_total", 0)
    if t <= 0 or not e.name.startswith("aten::"):
        continue
    where = next((s for s in (e.stack or []) if "/s2p_b200/" in s or "/repo/" in s), "?")
    rows.append((t, e.name, str(e.input_shapes)[:90], where.split("/repo/")[-1][:70]))
agg = {}
for t, n, sh, w in rows:
    k = (n, sh, w)
    a = agg.setdefault(k, [0, 0.0])
    a[0] += 1
    a[1] += t
tot = sum(v[1] for v in agg.values())
print(f"aten ops with device time: {tot:.0f} us")
for (n, sh, w), (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"{t:9.0f} us x{c:3d}  {n:28s} {sh:90s} {w}")
