"""GPU: VGG19 perceptual loss of the training step (ORACLE_SPEC v1.2) -- max-pool kernels against torch,
feature maps / loss / input gradient against the fp32 oracle, and the full G + D step with the VGG term.
Tolerances: bf16 activations through 13 conv layers -> relative L2 error per feature map <= 3e-2, loss within
2 %; input gradients: no worse than 1.1x (smooth loss, per slice) / 1.25x (L1 feature loss, sign gradients) the
deviation of torch's own bf16 autocast of the SAME oracle from fp32, and within the absolute per-slice bounds
measured on B200 (see the comments in the test)."""
import pytest
import torch
import torch.nn.functional as F

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    import s2p_b200
    return s2p_b200


@pytest.mark.parametrize("h,w,c", [(100, 100, 64), (25, 25, 256), (12, 13, 8)])
def test_maxpool_kernels_vs_torch(pkg, h, w, c):
    from s2p_b200 import autograd as A
    g = torch.Generator().manual_seed(h + c)
    x = torch.randn(3, c, h, w, generator=g).cuda()
    x = torch.relu(x).to(torch.bfloat16)                       # ties at 0, as after a ReLU
    xa = x.permute(0, 2, 3, 1).contiguous().requires_grad_()
    xb = x.float().requires_grad_()
    ya, yb = A.maxpool2(xa), F.max_pool2d(xb, 2, 2)
    assert torch.equal(ya.permute(0, 3, 1, 2).float(), yb)
    dy = torch.randn(yb.shape, generator=g).cuda().to(torch.bfloat16)
    ya.backward(dy.permute(0, 2, 3, 1).contiguous())
    yb.backward(dy.float())
    assert torch.equal(xa.grad.permute(0, 3, 1, 2).float(), xb.grad)     # same tie rule, exact routing


def test_vgg_features_loss_and_gradient_vs_oracle(pkg):
    import s2p_oracle as O
    from s2p_b200.networks.vgg import vgg_loss
    ora = O.VGG19Features().cuda()
    net = pkg.VGG19FeaturesB200().cuda()
    net.load_state_dict(ora.state_dict())
    g = torch.Generator().manual_seed(6)
    fake = (torch.rand(3, 3, 100, 100, generator=g) * 2 - 1).cuda()
    real = (torch.rand(3, 3, 100, 100, generator=g) * 2 - 1).cuda()
    fa, fb = fake.clone().requires_grad_(), fake.clone().requires_grad_()
    for a, b in zip(net(fa), ora(fb)):
        a = a.permute(0, 3, 1, 2).float()
        assert a.shape == b.shape
        rel = ((a - b).norm() / b.norm()).item()
        assert rel < 3e-2, rel
    la, lb = vgg_loss(net, fa, real), O.vgg_loss(ora, fb, real)
    print(f"VGG loss {la.item():.5f} vs oracle {lb.item():.5f}")
    assert abs(la.item() - lb.item()) <= 2e-2 * abs(lb.item())
    la.backward()
    lb.backward()
    rel = ((fa.grad - fb.grad).norm() / fb.grad.norm()).item()
    cos = F.cosine_similarity(fa.grad.flatten(), fb.grad.flatten(), dim=0).item()
    print(f"VGG L1-loss input-gradient rel L2 {rel:.3e} cosine {cos:.5f}")
    # The L1 feature loss back-propagates sign(a - b): any bf16 rounding of a feature flips signs where |a - b| is
    # small, so this gradient is discontinuous in the activations.  Yardstick: the SAME fp32 oracle under torch's own
    # bf16 autocast (cuDNN) deviates from fp32 by a comparable amount; we must not be worse than 1.25x that.
    fc = fake.clone().requires_grad_()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        lc = O.vgg_loss(ora, fc, real)
    lc.backward()
    rel_amp = ((fc.grad - fb.grad).norm() / fb.grad.norm()).item()
    print(f"torch bf16-autocast oracle vs fp32 oracle: rel L2 {rel_amp:.3e}")
    assert cos > 0.9 and rel <= 1.25 * rel_amp + 0.02
    assert all(p.grad is None for p in net.parameters())          # frozen extractor: no wgrad launched
    # smooth loss per slice: the backward path itself (dgrad convs, ReLU masks, max-pool routing).  ReLU masks and
    # pooling argmaxes flip under bf16 rounding, so the deviation from fp32 grows with depth (2e-3 after one conv,
    # 0.24 after thirteen) -- for torch's own bf16 autocast (cuDNN) exactly as for this path; the bar is that
    # yardstick, slice by slice, plus the absolute bound measured on B200.
    bounds = (5e-3, 5e-2, 0.1, 0.2, 0.3)
    for k in range(5):
        ga, gb_, gc = (fake.clone().requires_grad_() for _ in range(3))
        net(ga)[k].float().pow(2).mean().backward()
        ora(gb_)[k].pow(2).mean().backward()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            lk = ora(gc)[k].float().pow(2).mean()
        lk.backward()
        r_own = ((ga.grad - gb_.grad).norm() / gb_.grad.norm()).item()
        r_amp = ((gc.grad - gb_.grad).norm() / gb_.grad.norm()).item()
        print(f"slice {k}: input-gradient rel L2 {r_own:.3e} (torch bf16 autocast: {r_amp:.3e})")
        assert r_own <= 1.1 * r_amp + 1e-3 and r_own <= bounds[k]


def test_training_step_with_vgg_vs_oracle(pkg):
    import s2p_oracle as O
    from s2p_b200.training import S2PTrainer
    oG = common.build_oracle("walker", "scaled").cuda().train()
    torch.manual_seed(5)
    oD = O.define_D(O.make_options(init_type="scaled")).cuda()
    oV = O.VGG19Features().cuda()
    nG = pkg.define_G(pkg.make_options(env_type="walker", gpu_ids="0"))
    nD = pkg.define_D(pkg.make_options(gpu_ids="0"))
    nV = pkg.VGG19FeaturesB200().cuda()
    nG.load_state_dict(oG.state_dict())
    nD.load_state_dict(oD.state_dict())
    image, states = common.make_inputs("walker", 4, 1)
    g = torch.Generator().manual_seed(9)
    real = (torch.rand(4, 3, 100, 100, generator=g) * 2 - 1).cuda()
    image, st = image.cuda(), states[:, 0].cuda()
    optG = torch.optim.Adam(oG.parameters(), **O.ADAM)
    optD = torch.optim.Adam(oD.parameters(), **O.ADAM)
    g_ref, d_ref, _ = O.training_step(oG, oD, optG, optD, image, st, real, vgg=oV)
    tr = S2PTrainer(nG, nD, O.ADAM["lr"], O.ADAM["betas"], vgg=nV)
    g_out, d_out = tr.train_step(image, st, real)
    assert set(g_out) == set(g_ref) == {"GAN", "GAN_Feat", "L1", "VGG"}
    for k in g_ref:
        a, b = float(g_out[k].detach()), float(g_ref[k].detach())
        print(f"G loss {k}: {a:.5f} vs oracle {b:.5f}")
        assert abs(a - b) <= 2e-2 * abs(b) + 2e-3
    assert abs(float(d_out.detach()) - float(d_ref.detach())) <= 2e-2 * abs(float(d_ref.detach())) + 2e-3


def test_l1_mean_kernels_vs_torch(pkg):
    """fused mean|a-b| on NHWC bf16 (channel-sliced and NCHW-permuted views included) and its sign backward"""
    from s2p_b200 import autograd as A
    g = torch.Generator().manual_seed(11)
    for shape, sl in (((3, 25, 25, 256), 256), ((2, 50, 50, 128), 64), ((4, 6, 6, 512), 512)):
        a = torch.randn(shape, generator=g).cuda().to(torch.bfloat16)
        b = torch.randn(shape, generator=g).cuda().to(torch.bfloat16)
        b[0, 0, 0, :8] = a[0, 0, 0, :8]                               # exact ties -> zero gradient
        a.requires_grad_()
        av, bv = a[..., :sl].permute(0, 3, 1, 2), b[..., :sl].permute(0, 3, 1, 2)     # the discriminator's NCHW views
        out = A.l1_mean(av, bv)
        a2 = a.detach().clone().requires_grad_()
        ref = F.l1_loss(a2[..., :sl].float(), b[..., :sl].float())
        assert abs(out.item() - ref.item()) <= 1e-6 * abs(ref.item()) + 1e-7
        (out * 3.0).backward()
        (ref * 3.0).backward()
        assert torch.equal(a.grad, a2.grad)
        assert torch.equal(A.l1_mean(av, bv), out.detach())           # bit-stable
