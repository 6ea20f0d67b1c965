"""Per-parameter gradient comparison of forward_train vs the oracle's autograd, with the bf16-autocast yardstick:
ours vs fp32 oracle next to torch-autocast(bf16) oracle vs fp32 oracle, per tensor (diagnostic for
tests/test_training_gpu.py::test_generator_gradients_vs_oracle)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch, torch.nn.functional as F
import common, s2p_b200
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
env = sys.argv[1] if len(sys.argv) > 1 else "cheetah"
init = sys.argv[2] if len(sys.argv) > 2 else "scaled"
ora = common.build_oracle(env, init).cuda().train()
ora_ac = common.build_oracle(env, init).cuda().train()
net = s2p_b200.define_G(s2p_b200.make_options(env_type=env, gpu_ids="0"))
net.load_state_dict(ora.state_dict()); net.train()
image, states = common.make_inputs(env, 2, 1)
g = torch.Generator().manual_seed(7)
target = (torch.rand(2, 3, 100, 100, generator=g) * 2 - 1).cuda()
image, st = image.cuda(), states[:, 0].cuda()
out = net(image, st)
F.l1_loss(out, target).backward()
ref = ora(image, st)
F.l1_loss(ref, target).backward()
with torch.autocast("cuda", dtype=torch.bfloat16):
    out_ac = ora_ac(image, st)
F.l1_loss(out_ac.float(), target).backward()
print(f"forward psnr: ours {common.psnr01(out.detach().cpu(), ref.detach().cpu()):.2f} dB, autocast {common.psnr01(out_ac.detach().float().cpu(), ref.detach().cpu()):.2f} dB")
refs, acs = dict(ora.named_parameters()), dict(ora_ac.named_parameters())
gmax = max(r.grad.norm().item() for r in refs.values())
worst_ratio = (0, "")
for name, prm in net.named_parameters():
    a, b, c = prm.grad.double().flatten(), refs[name].grad.double().flatten(), acs[name].grad.double().flatten()
    rel = ((a - b).norm() / (b.norm() + 1e-30)).item()
    rel_ac = ((c - b).norm() / (b.norm() + 1e-30)).item()
    cos = (torch.dot(a, b) / (a.norm() * b.norm() + 1e-30)).item()
    ratio = rel / max(rel_ac, 1e-12)
    flag = " <<<" if ratio > 1.5 and b.norm().item() > 1e-6 * gmax else ""
    if b.norm().item() > 1e-6 * gmax and ratio > worst_ratio[0]:
        worst_ratio = (ratio, name)
    print(f"{name:36s} n={a.numel():8d} |ref|={b.norm().item():.2e} ours={rel:.3e} autocast={rel_ac:.3e} ratio={ratio:6.2f} cos={cos:.5f}{flag}")
print("worst ours/autocast ratio", worst_ratio)
