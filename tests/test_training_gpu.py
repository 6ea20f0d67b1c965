"""GPU: the differentiable training path (s2p_b200.autograd + S2PGeneratorB200.forward_train)
against torch autograd of the frozen fp32 oracle on the same weights and inputs.
Tolerances: activations and their gradients are bf16 (8 mantissa bits) through ~40 layers, so
parameter gradients are compared by relative L2 error and cosine similarity, per tensor."""
import pytest
import torch
import torch.nn.functional as F

import common

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    import s2p_b200
    return s2p_b200


def nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def test_conv_autograd(pkg):
    from s2p_b200 import autograd as A
    g = torch.Generator().manual_seed(1)
    for (b, h, cin, cout, k, stride, act, res) in [(2, 25, 64, 128, 3, 1, 2, True), (2, 50, 128, 64, 3, 2, 0, False),
                                                   (2, 25, 128, 128, 1, 1, 1, False)]:
        x = torch.randn(b, cin, h, h, generator=g).cuda()
        w = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).cuda().requires_grad_()
        bias = (torch.randn(cout, generator=g) * 0.1).cuda().requires_grad_()
        ho = (h + 2 * (k // 2) - k) // stride + 1
        r = torch.randn(b, cout, ho, ho, generator=g).cuda() if res else None
        dy = torch.randn(b, cout, ho, ho, generator=g).cuda()
        xb = nhwc(x).to(torch.bfloat16).requires_grad_()
        rb = nhwc(r).to(torch.bfloat16).requires_grad_() if res else None
        y = A.conv(xb, w, bias, rb, stride, act)
        y.backward(nhwc(dy).to(torch.bfloat16))
        xr = xb.detach().float().permute(0, 3, 1, 2).requires_grad_()
        wr = w.detach().to(torch.bfloat16).float().requires_grad_()
        br = bias.detach().clone().requires_grad_()
        rr = rb.detach().float().permute(0, 3, 1, 2).requires_grad_() if res else None
        yr = F.conv2d(xr, wr, br, stride, k // 2)
        if res:
            yr = yr + rr
        yr = F.relu(yr) if act == 1 else F.leaky_relu(yr, 0.2) if act == 2 else yr
        yr.backward(nhwc(dy).to(torch.bfloat16).float().permute(0, 3, 1, 2))

        def rel(a, b_):
            return ((a.float() - b_).norm() / (b_.norm() + 1e-12)).item()
        assert rel(y, nhwc(yr)) < 5e-3
        assert rel(xb.grad, nhwc(xr.grad)) < 1e-2          # bf16-rounded dy/weights, bf16 store
        assert rel(w.grad, wr.grad) < 1e-2
        assert rel(bias.grad, br.grad) < 1e-2
        if res:
            assert rel(rb.grad, nhwc(rr.grad)) < 1e-2


@pytest.mark.parametrize("c,up,hx,mod", [(64, 1, 50, True), (128, 2, 25, True), (256, 1, 25, False), (512, 2, 12, True)])
def test_mat_norm_autograd(pkg, c, up, hx, mod):
    from s2p_b200 import autograd as A, ops
    g = torch.Generator().manual_seed(c + up)
    b, h = 2, hx * up
    x = (torch.randn(b, hx, hx, c, generator=g) * 2 + 0.5).cuda().to(torch.bfloat16).requires_grad_()
    gs = (torch.randn(b, c, generator=g) * 0.3).cuda().requires_grad_() if mod else None
    bs = (torch.randn(b, c, generator=g) * 0.3).cuda().requires_grad_() if mod else None
    gb = (torch.randn(b, h, h, 2 * c, generator=g) * 0.3).cuda().to(torch.bfloat16).requires_grad_() if mod else None
    dy = torch.randn(b, h, h, c, generator=g).cuda().to(torch.bfloat16)
    xf = x.detach().float()
    mr = torch.stack([xf.mean(dim=(1, 2)), 1 / torch.sqrt(xf.var(dim=(1, 2), unbiased=False) + 1e-5)], dim=-1).contiguous()
    y = A.mat_norm(x, mr, gs, bs, gb, up, 2)
    y.backward(dy)
    # torch restatement (fp32) of oracle MAT.forward + leaky_relu + nearest upsample
    xr = x.detach().float().requires_grad_()
    xh = F.instance_norm(xr.permute(0, 3, 1, 2), eps=1e-5)
    if up == 2:
        xh = F.interpolate(xh, scale_factor=2, mode="nearest")
    xh = xh.permute(0, 2, 3, 1)
    if mod:
        gsr, bsr = gs.detach().clone().requires_grad_(), bs.detach().clone().requires_grad_()
        gbr = gb.detach().float().requires_grad_()
        pre = xh * (1 + gsr[:, None, None] + gbr[..., :c]) + bsr[:, None, None] + gbr[..., c:]
    else:
        pre = xh
    yr = F.leaky_relu(pre, 0.2)
    yr.backward(dy.float())

    def rel(a, b_):
        return ((a.float() - b_).norm() / (b_.norm() + 1e-12)).item()
    assert rel(y, yr) < 5e-3
    assert rel(x.grad, xr.grad) < 2e-2
    if mod:
        assert rel(gb.grad, gbr.grad) < 1e-2
        assert rel(gs.grad, gsr.grad) < 1e-2 and rel(bs.grad, bsr.grad) < 1e-2


def test_generator_gradients_vs_oracle(pkg):
    """L1 loss to a random target: every parameter gradient of the B200 training path against
    the oracle's autograd (fp32, cuDNN) on the same weights."""
    ora = common.build_oracle("cheetah", "scaled").cuda().train()
    net = pkg.define_G(pkg.make_options(env_type="cheetah", gpu_ids="0"))
    net.load_state_dict(ora.state_dict())
    net.train()
    image, states = common.make_inputs("cheetah", 2, 1)
    g = torch.Generator().manual_seed(7)
    target = (torch.rand(2, 3, 100, 100, generator=g) * 2 - 1).cuda()
    image, st = image.cuda(), states[:, 0].cuda()
    out = net(image, st)
    assert out.requires_grad
    F.l1_loss(out, target).backward()
    ref = ora(image, st)
    F.l1_loss(ref, target).backward()
    p = common.psnr01(out.detach().cpu(), ref.detach().cpu())
    print(f"training forward psnr {p:.1f} dB")
    assert p >= 45.0
    refs = dict(ora.named_parameters())
    worst = (0.0, "")
    cos_min = (1.0, "")
    for name, prm in net.named_parameters():
        assert prm.grad is not None, f"no gradient for {name}"
        a, b = prm.grad.double().flatten(), refs[name].grad.double().flatten()
        relerr = ((a - b).norm() / (b.norm() + 1e-30)).item()
        cos = (torch.dot(a, b) / (a.norm() * b.norm() + 1e-30)).item()
        if relerr > worst[0]:
            worst = (relerr, name)
        if cos < cos_min[0]:
            cos_min = (cos, name)
    print(f"worst relative L2 gradient error {worst[0]:.3e} ({worst[1]}); lowest cosine {cos_min[0]:.5f} ({cos_min[1]})")
    assert worst[0] < 0.15 and cos_min[0] > 0.99
