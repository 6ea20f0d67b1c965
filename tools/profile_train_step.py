"""One eager G + D training step inside a cudaProfilerStart/Stop range, for ncu.  usage: python tools/profile_train_step.py [batch]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

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
torch.cuda.profiler.start()
tr.train_step(prev, state, real)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profile_train_step: done")
