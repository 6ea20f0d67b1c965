"""Per-parameter gradient comparison of forward_train vs the oracle's autograd (diagnostic)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch, torch.nn.functional as F
import common, s2p_b200
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
ora = common.build_oracle("cheetah", "scaled").cuda().train()
net = s2p_b200.define_G(s2p_b200.make_options(env_type="cheetah", gpu_ids="0"))
net.load_state_dict(ora.state_dict()); net.train()
image, states = common.make_inputs("cheetah", 2, 1)
g = torch.Generator().manual_seed(7)
target = (torch.rand(2, 3, 100, 100, generator=g) * 2 - 1).cuda()
image, st = image.cuda(), states[:, 0].cuda()
F.l1_loss(net(image, st), target).backward()
F.l1_loss(ora(image, st), target).backward()
refs = dict(ora.named_parameters())
for name, prm in net.named_parameters():
    a, b = prm.grad.double().flatten(), refs[name].grad.double().flatten()
    rel = ((a - b).norm() / (b.norm() + 1e-30)).item()
    cos = (torch.dot(a, b) / (a.norm() * b.norm() + 1e-30)).item()
    print(f"{name:40s} n={a.numel():8d} |ref|={b.norm().item():.3e} |got|={a.norm().item():.3e} rel={rel:.3e} cos={cos:.5f}")
