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
    """L1 loss to a random target: every parameter gradient of the B200 training path against the oracle's autograd
    (fp32, cuDNN) on the same weights -- with a YARDSTICK instead of free tolerances: the same oracle run under
    torch's bf16 autocast (library kernels, bf16 activations, fp32 parameters and gradients).  Per tensor, this
    path's deviation from fp32 must not exceed 1.5x autocast's own deviation from fp32.  (Measured: 0.85-1.03x on all
    160 non-scalar tensors, tools/grad_probe.py; the 10-20 % error at the bottom of the trunk is what bf16 activations
    through ~40 random-weight layers cost any implementation: autocast shows 24 %.)"""
    ora = common.build_oracle("cheetah", "scaled").cuda().train()
    ora_ac = common.build_oracle("cheetah", "scaled").cuda().train()
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
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out_ac = ora_ac(image, st)
    F.l1_loss(out_ac.float(), target).backward()
    p = common.psnr01(out.detach().cpu(), ref.detach().cpu())
    p_ac = common.psnr01(out_ac.detach().float().cpu(), ref.detach().cpu())
    print(f"training forward psnr {p:.2f} dB (torch bf16 autocast of the oracle: {p_ac:.2f} dB)")
    assert p >= 47.0 and p >= p_ac
    refs, acs = dict(ora.named_parameters()), dict(ora_ac.named_parameters())
    gmax = max(r.grad.norm().item() for r in refs.values())
    scalars, worst = [], (0.0, "", 0.0, 0.0)
    for name, prm in net.named_parameters():
        assert prm.grad is not None, f"no gradient for {name}"
        a, b, c = prm.grad.double().flatten(), refs[name].grad.double().flatten(), acs[name].grad.double().flatten()
        if b.norm().item() < 1e-6 * gmax:
            # structurally zero gradients (a bias feeding an instance norm): ours must vanish too
            assert a.norm().item() < 1e-3 * gmax, name
            continue
        if a.numel() == 1:
            scalars.append((name, abs(a.item() - b.item()), abs(c.item() - b.item()), abs(b.item())))
            continue
        rel = ((a - b).norm() / b.norm()).item()
        rel_ac = ((c - b).norm() / b.norm()).item()
        if rel / max(rel_ac, 1e-12) > worst[0]:
            worst = (rel / max(rel_ac, 1e-12), name, rel, rel_ac)
        assert rel <= 1.5 * rel_ac + 1e-3, f"{name}: relative L2 error {rel:.3e} vs autocast yardstick {rel_ac:.3e}"
    print(f"worst ours / autocast ratio over the non-scalar tensors: {worst[0]:.2f} ({worst[1]}: {worst[2]:.3e} vs {worst[3]:.3e})")
    # blend scalars: each gradient is a sum of ~1e6 cancelling terms (|g| ~ 1e-4 of the largest tensor gradient), so a
    # single scalar's relative error is luck for either implementation; the yardstick is autocast's worst ABSOLUTE
    # error over the 26 scalars
    assert len(scalars) == 26
    yard = max(e_ac for _, _, e_ac, _ in scalars)
    bad = [(n, e, e_ac) for n, e, e_ac, _ in scalars if e > 1.5 * yard]
    print(f"blend scalars: worst abs error ours {max(e for _, e, _, _ in scalars):.3e}, autocast {yard:.3e}")
    assert not bad, bad


def test_discriminator_vs_oracle(pkg):
    import s2p_oracle as O
    torch.manual_seed(5)
    ora = O.define_D(O.make_options(init_type="scaled")).cuda()
    net = pkg.define_D(pkg.make_options(gpu_ids="0"))
    net.load_state_dict(ora.state_dict())
    g = torch.Generator().manual_seed(3)
    x = (torch.rand(4, 6, 100, 100, generator=g) * 2 - 1).cuda()
    xa = x.clone().requires_grad_()
    xb = x.clone().requires_grad_()
    out, ref = net(xa), ora(xb)
    loss = sum(f.float().pow(2).mean() for o in out for f in o)
    lref = sum(f.pow(2).mean() for o in ref for f in o)
    loss.backward()
    lref.backward()
    for o, r in zip(out, ref):
        for a, b in zip(o, r):
            assert tuple(a.shape) == tuple(b.shape)
            rel = ((a.float() - b).norm() / b.norm()).item()
            assert rel < 3e-2, rel
    assert abs(loss.item() - lref.item()) < 2e-2 * abs(lref.item())
    rel = ((xa.grad - xb.grad).norm() / xb.grad.norm()).item()
    print(f"D input-gradient relative error {rel:.3e}")
    assert rel < 0.1
    refs = dict(ora.named_parameters())
    for name, prm in net.named_parameters():
        a, b = prm.grad.double().flatten(), refs[name].grad.double().flatten()
        if b.norm().item() < 1e-7:
            continue
        rel = ((a - b).norm() / b.norm()).item()
        assert rel < 0.1, (name, rel)


def test_adam_kernel_matches_torch(pkg):
    from s2p_b200.training import FusedAdam
    g = torch.Generator().manual_seed(2)
    p0 = torch.randn(1000, 37, generator=g).cuda()
    a = torch.nn.Parameter(p0.clone())
    b = torch.nn.Parameter(p0.clone())
    mine, ref = FusedAdam([a], 2e-4, (0.0, 0.9)), torch.optim.Adam([b], lr=2e-4, betas=(0.0, 0.9))
    for i in range(3):
        gr = torch.randn(1000, 37, generator=g).cuda()
        a.grad, b.grad = gr.clone(), gr.clone()
        mine.step()
        ref.step()
    assert (a - b).abs().max().item() < 1e-6


def test_training_step_vs_oracle(pkg):
    """one G step + one D step: loss terms against the oracle's training_step on the same weights."""
    import s2p_oracle as O
    from s2p_b200.training import S2PTrainer
    oG = common.build_oracle("walker", "scaled").cuda().train()
    torch.manual_seed(5)
    oD = O.define_D(O.make_options(init_type="scaled")).cuda()
    nG = pkg.define_G(pkg.make_options(env_type="walker", gpu_ids="0"))
    nD = pkg.define_D(pkg.make_options(gpu_ids="0"))
    nG.load_state_dict(oG.state_dict())
    nD.load_state_dict(oD.state_dict())
    image, states = common.make_inputs("walker", 4, 1)
    g = torch.Generator().manual_seed(9)
    real = (torch.rand(4, 3, 100, 100, generator=g) * 2 - 1).cuda()
    image, st = image.cuda(), states[:, 0].cuda()
    optG = torch.optim.Adam(oG.parameters(), **O.ADAM)
    optD = torch.optim.Adam(oD.parameters(), **O.ADAM)
    g_ref, d_ref, _ = O.training_step(oG, oD, optG, optD, image, st, real)
    tr = S2PTrainer(nG, nD, O.ADAM["lr"], O.ADAM["betas"])
    g_out, d_out = tr.train_step(image, st, real)
    for k in g_ref:
        a, b = float(g_out[k]), float(g_ref[k])
        print(f"G loss {k}: {a:.5f} vs oracle {b:.5f}")
        assert abs(a - b) <= 2e-2 * abs(b) + 2e-3
    print(f"D loss: {float(d_out):.5f} vs oracle {float(d_ref):.5f}")
    assert abs(float(d_out) - float(d_ref)) <= 2e-2 * abs(float(d_ref)) + 2e-3
    # parameters moved, and by Adam-sized steps (|delta| <= lr / (1 - beta1) ...)
    moved = max((p - q).abs().max().item() for p, q in zip(nG.parameters(), oG.state_dict().values()))
    assert 0 < moved < 1e-2


def test_act_bwd_bias_grad_kernel(pkg):
    from s2p_b200 import ops
    g = torch.Generator().manual_seed(4)
    for c, act in ((64, 2), (128, 1), (512, 0), (72, 2)):
        dy = torch.randn(3, 37, 41, c, generator=g).cuda().to(torch.bfloat16)
        y = torch.randn(3, 37, 41, c, generator=g).cuda().to(torch.bfloat16)
        out, db = ops.act_bwd_bias_grad(dy, y, act, True)
        ref = dy.float() if act == 0 else torch.where(y.float() > 0, dy.float(), dy.float() * (0.2 if act == 2 else 0.0))
        ref = ref.to(torch.bfloat16)
        assert torch.equal(out, ref)
        assert torch.allclose(db, ref.float().sum(dim=(0, 1, 2)), rtol=1e-4, atol=1e-2)
        assert torch.equal(ops.act_bwd_bias_grad(dy, y, act, True)[1], db)     # bit-stable


def test_graphed_step_equals_eager(pkg):
    """the CUDA-graph replay of the G + D step is the same computation as the eager step."""
    from s2p_b200.training import S2PTrainer
    image, states = common.make_inputs("cheetah", 2, 1)
    g = torch.Generator().manual_seed(9)
    real = (torch.rand(2, 3, 100, 100, generator=g) * 2 - 1).cuda()
    image, st = image.cuda(), states[:, 0].cuda()
    trainers = []
    for _ in range(2):
        torch.manual_seed(77)
        nG = pkg.define_G(pkg.make_options(env_type="cheetah", gpu_ids="0", init_type="scaled"))
        nD = pkg.define_D(pkg.make_options(gpu_ids="0"))
        trainers.append(S2PTrainer(nG, nD))
    eager, graphed = trainers
    graphed.capture(image, st, real, warmup=2)              # warm-up steps are rolled back: still at step 0
    assert graphed.optG.step_count == 0
    for _ in range(2):
        eager.train_step(image, st, real)
        graphed.train_step_graphed(image, st, real)
    g_e, d_e = eager.train_step(image, st, real)          # step 3
    g_g, d_g = graphed.train_step_graphed(image, st, real)  # step 3
    g_g, d_g = {k: v.clone() for k, v in g_g.items()}, d_g.clone()
    torch.cuda.synchronize()
    for k in g_e:
        assert abs(float(g_e[k]) - float(g_g[k])) <= 1e-4 * abs(float(g_e[k])) + 1e-5, k
    assert abs(float(d_e) - float(d_g)) <= 1e-4 * abs(float(d_e)) + 1e-5
    assert graphed.optG.step_count == 3


def test_linear_splitk_vs_torch_and_batch_invariant(pkg):
    """K >= 256 runs split-K: same values as torch (fp32 reorder only), and a row's result never depends on the batch"""
    from s2p_b200 import ops
    g = torch.Generator().manual_seed(8)
    for (b, k, n) in ((16, 9600, 512), (16, 512, 512), (70, 357, 512), (3, 200, 9600)):
        x = torch.randn(b, k, generator=g).cuda()
        w = (torch.randn(n, k, generator=g) / k ** 0.5).cuda()
        bias = torch.randn(n, generator=g).cuda()
        y = ops.linear(x, w, bias, 2)
        ref = F.leaky_relu(x.double() @ w.double().t() + bias.double(), 0.2).float()
        assert (y - ref).abs().max().item() < 2e-5 * max(1.0, ref.abs().max().item())
        assert torch.equal(ops.linear(x[:1].contiguous(), w, bias, 2), y[:1])      # batch-invariant bits
        assert torch.equal(ops.linear(x, w, bias, 2), y)                           # run-to-run


def test_adam_multi_tensor_matches_torch(pkg):
    from s2p_b200.training import FusedAdam
    g = torch.Generator().manual_seed(12)
    shapes = [(64,), (512, 128, 3, 3), (1,), (9600, 512), (8193,)] + [(33, 7)] * 70      # > 64 tensors: two launches
    a = [torch.nn.Parameter(torch.randn(s, generator=g).cuda()) for s in shapes]
    b = [torch.nn.Parameter(p.detach().clone()) for p in a]
    mine, ref = FusedAdam(a, 2e-4, (0.0, 0.9)), torch.optim.Adam(b, lr=2e-4, betas=(0.0, 0.9))
    for _ in range(3):
        for p, q in zip(a, b):
            gr = torch.randn(p.shape, generator=g).cuda()
            p.grad, q.grad = gr.clone(), gr.clone()
        a[5].grad = None                                        # a parameter without a gradient is skipped
        b[5].grad = None
        mine.step()
        ref.step()
    for p, q in zip(a, b):
        assert (p - q).abs().max().item() < 1e-6


def test_multi_scale_fn_vs_torch(pkg):
    """blend scaling of many parameter tensors into concatenated operands, forward and backward"""
    from s2p_b200 import autograd as A
    g = torch.Generator().manual_seed(13)
    dev = torch.device("cuda")
    srcs = [torch.randn(s, generator=g).to(dev).requires_grad_() for s in ((64, 128, 3, 3), (32, 128, 3, 3), (64,), (32,), (10, 7))]
    scales = torch.rand(3, generator=g).to(dev).requires_grad_()
    shapes = [(96, 128, 3, 3), (96,), (10, 7)]
    where = [(0, 0, 0), (0, 64 * 128 * 9, 1), (1, 0, 0), (1, 64, 1), (2, 0, 2)]
    onehot = torch.zeros(3, 5, device=dev)
    onehot[torch.tensor([k for _, _, k in where]), torch.arange(5)] = 1.0
    outs = A.multi_scale(scales, (shapes, where, onehot), srcs)
    srcs2 = [t.detach().clone().requires_grad_() for t in srcs]
    scales2 = scales.detach().clone().requires_grad_()
    ref = [torch.cat([srcs2[0] * scales2[0], srcs2[1] * scales2[1]]), torch.cat([srcs2[2] * scales2[0], srcs2[3] * scales2[1]]),
           srcs2[4] * scales2[2]]
    gs = [torch.randn(o.shape, generator=g).to(dev) for o in ref]
    for o, r in zip(outs, ref):
        assert torch.equal(o, r)
    torch.autograd.backward(list(outs), gs)
    torch.autograd.backward(ref, gs)
    for t, t2 in zip(srcs, srcs2):
        assert torch.equal(t.grad, t2.grad)
    assert torch.allclose(scales.grad, scales2.grad, rtol=1e-4, atol=1e-4)


@pytest.mark.gpu
def test_pack_conv_weights_multi_matches_single_tensor_packs():
    """one launch for many tensors == the per-tensor pack kernels, bit for bit (incl. > 48 tensors, odd shapes, no-dgrad)"""
    from s2p_b200 import ops
    g = torch.Generator().manual_seed(5)
    shapes = [(64, 64, 3, 3), (128, 64, 1, 1), (8, 64, 4, 4), (512, 128, 3, 3), (64, 320, 1, 1)] + [(64, 64, 3, 3)] * 50
    ws = [torch.randn(*s, generator=g).cuda() for s in shapes]
    want = [i % 3 != 1 for i in range(len(ws))]
    out = ops.pack_conv_weights_multi(ws, want)
    assert len(out) == len(ws)
    for w, d, (f, gq) in zip(ws, want, out):
        assert f.data_ptr() % 256 == 0
        assert torch.equal(f, ops.pack_conv_weight(w))
        if d:
            assert gq.data_ptr() % 256 == 0 and torch.equal(gq, ops.pack_conv_weight_dgrad(w))
        else:
            assert gq is None
    assert ops.pack_conv_weights_multi([], []) == []


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3, 3, 100, 100), (2, 6, 50, 50), (1, 1, 7, 5), (2, 8, 9, 9)])
def test_pad_channels_fwd_bwd(shape):
    """NCHW fp32 -> NHWC bf16 (64 zero-padded channels) and its backward, bit-exact vs the torch expression it replaces"""
    from s2p_b200 import autograd as A
    g = torch.Generator().manual_seed(11)
    x = torch.randn(*shape, generator=g).cuda().requires_grad_(True)
    y = A.pad_channels(x)
    b, c, h, w = shape
    ref = torch.zeros(b, h, w, 64, dtype=torch.bfloat16, device="cuda")
    ref[..., :c] = x.detach().permute(0, 2, 3, 1)
    assert y.shape == ref.shape and torch.equal(y, ref)
    dy = torch.randn(b, h, w, 64, generator=g).to(torch.bfloat16).cuda()
    (dx,) = torch.autograd.grad(y, x, dy)
    assert torch.equal(dx, dy[..., :c].float().permute(0, 3, 1, 2))
    with pytest.raises(RuntimeError):
        A.pad_channels(torch.zeros(1, 9, 4, 4, device="cuda"))


@pytest.mark.gpu
@pytest.mark.parametrize("b,ho,wo,c,hout,wout", [(2, 51, 51, 64, 101, 101), (1, 3, 4, 8, 6, 9), (3, 13, 13, 128, 25, 25)])
def test_zero_insert2(b, ho, wo, c, hout, wout):
    from s2p_b200 import ops
    g = torch.Generator().manual_seed(3)
    dy = torch.randn(b, ho, wo, c, generator=g).to(torch.bfloat16).cuda()
    ref = torch.zeros(b, hout, wout, c, dtype=torch.bfloat16, device="cuda")
    ref[:, 0:2 * ho:2, 0:2 * wo:2] = dy
    assert torch.equal(ops.zero_insert2(dy, hout, wout), ref)
    wide = torch.randn(b, ho, wo, c + 8, generator=g).to(torch.bfloat16).cuda()
    ref[:, 0:2 * ho:2, 0:2 * wo:2] = wide[..., :c]
    assert torch.equal(ops.zero_insert2(wide[..., :c], hout, wout), ref)      # channel-sliced (pitched) source


@pytest.mark.gpu
def test_tanh_out_fwd_bwd():
    """generator output head: tanh + NHWC(64) -> NCHW(3), forward and backward vs the torch expression it replaces"""
    from s2p_b200 import autograd as A
    g = torch.Generator().manual_seed(13)
    rgb = (torch.randn(3, 20, 12, 64, generator=g) * 2).to(torch.bfloat16).cuda().requires_grad_(True)
    y = A.tanh_out(rgb, 3)
    ref_in = rgb.detach().clone().requires_grad_(True)
    ref = torch.tanh(ref_in[..., :3].float()).permute(0, 3, 1, 2).contiguous()
    assert y.shape == ref.shape and (y - ref).abs().max().item() < 2e-6
    dy = torch.randn(3, 3, 20, 12, generator=g).cuda()
    (dx,) = torch.autograd.grad(y, rgb, dy)
    (dref,) = torch.autograd.grad(ref, ref_in, dy)
    assert dx.shape == dref.shape and dx.dtype == torch.bfloat16
    assert torch.equal(dx[..., 3:], torch.zeros_like(dx[..., 3:]))
    assert (dx.float() - dref.float()).abs().max().item() <= 2 ** -7 * dref.float().abs().max().item()


@pytest.mark.gpu
def test_training_fits_a_fixed_batch_like_the_oracle(pkg):
    """Functional check over MANY steps (per-op parity cannot catch a sign or a stale-pack error that only shows after
    the optimiser has moved the weights): 30 G + D steps on one fixed batch.  The L1 term must fall, stay finite, and
    end within a few percent of where the eager fp32 oracle ends from the same weights on the same batch."""
    import s2p_oracle as O
    from s2p_b200.training import S2PTrainer
    # the ORACLE must be reproducible for a point-by-point comparison: cuDNN's default backward kernels are not
    # (atomics), and 30 GAN steps on one batch amplify that to +-2 % at the end of the oracle's own curve
    det = torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark
    torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = True, False
    oG = common.build_oracle("walker", "scaled").cuda().train()
    torch.manual_seed(6)
    oD = O.define_D(O.make_options(init_type="scaled")).cuda()
    nG = pkg.define_G(pkg.make_options(env_type="walker", gpu_ids="0"))
    nD = pkg.define_D(pkg.make_options(gpu_ids="0"))
    nG.load_state_dict(oG.state_dict())
    nD.load_state_dict(oD.state_dict())
    image, states = common.make_inputs("walker", 4, 1)
    g = torch.Generator().manual_seed(10)
    real = (torch.rand(4, 3, 100, 100, generator=g) * 0.6 - 0.3).cuda()
    image, st = image.cuda(), states[:, 0].cuda()
    optG = torch.optim.Adam(oG.parameters(), **O.ADAM)
    optD = torch.optim.Adam(oD.parameters(), **O.ADAM)
    tr = S2PTrainer(nG, nD, O.ADAM["lr"], O.ADAM["betas"])
    # yardstick for the points ALONG the curve: the same oracle trained under torch's bf16 autocast
    aG = common.build_oracle("walker", "scaled").cuda().train()
    torch.manual_seed(6)
    aD = O.define_D(O.make_options(init_type="scaled")).cuda()
    aoptG = torch.optim.Adam(aG.parameters(), **O.ADAM)
    aoptD = torch.optim.Adam(aD.parameters(), **O.ADAM)
    ours, ref, ac = [], [], []
    for _ in range(30):
        g_out, d_out = tr.train_step(image, st, real)
        g_ref, d_ref, _ = O.training_step(oG, oD, optG, optD, image, st, real)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            g_ac, _, _ = O.training_step(aG, aD, aoptG, aoptD, image, st, real)
        ours.append(float(g_out["L1"].detach()))
        ref.append(float(g_ref["L1"].detach()))
        ac.append(float(g_ac["L1"].detach()))
        assert all(torch.isfinite(v).all() for v in g_out.values()) and torch.isfinite(d_out).all()
    torch.backends.cudnn.deterministic, torch.backends.cudnn.benchmark = det
    rel = [abs(a - b) / b for a, b in zip(ours, ref)]
    rel_ac = [abs(a - b) / b for a, b in zip(ac, ref)]
    mean, mean_ac = sum(rel) / len(rel), sum(rel_ac) / len(rel_ac)
    print(f"L1 ours {ours[0]:.4f} -> {ours[-1]:.4f}; oracle {ref[0]:.4f} -> {ref[-1]:.4f}; deviation along the curve: mean "
          f"{100 * mean:.2f} % / max {100 * max(rel):.2f} % (bf16-autocast oracle: mean {100 * mean_ac:.2f} % / max "
          f"{100 * max(rel_ac):.2f} %, ends at {ac[-1]:.4f})")
    assert ours[-1] < 0.8 * ours[0] and ref[-1] < 0.8 * ref[0]
    assert abs(ours[-1] - ref[-1]) <= 0.05 * ref[-1]                  # the curve ends within 5 % of the fp32 oracle's
    # along the curve: 30 GAN steps on one batch are chaotic (the oracle's own curve moves by several percent per point
    # with cuDNN's algorithm choice), so single points are reported and the MEAN deviation is held to the autocast yardstick
    assert mean <= max(0.025, 1.5 * mean_ac)          # measured 1.90 % vs 0.92 % (both sides deterministic)


@pytest.mark.gpu
def test_inference_after_training_uses_the_updated_weights(pkg):
    """the inference schedule caches packed bf16 weights; a fused-Adam step, a replayed graph step and a plain in-place
    torch update must each invalidate that cache (the first two move the weights through raw pointers)"""
    from s2p_b200.training import S2PTrainer
    torch.manual_seed(3)
    opt = pkg.make_options(env_type="walker", gpu_ids="0", init_type="scaled")
    nG, nD = pkg.define_G(opt), pkg.define_D(opt)
    image, states = common.make_inputs("walker", 2, 1)
    g = torch.Generator().manual_seed(2)
    real = (torch.rand(2, 3, 100, 100, generator=g) * 2 - 1).cuda()
    image, st = image.cuda(), states[:, 0].cuda()

    def fresh_eval():
        ref = pkg.define_G(opt)
        ref.load_state_dict(nG.state_dict())
        return ref.eval()(image, st)

    def own_eval():
        nG.eval()
        with torch.no_grad():
            out = nG(image, st)
        nG.train()
        return out

    before = own_eval()                                  # builds the cache
    tr = S2PTrainer(nG, nD, lr=1e-2)                     # a step large enough to change the output visibly
    tr.train_step(image, st, real)
    after = own_eval()
    assert not torch.equal(after, before)
    assert torch.equal(after, fresh_eval())
    tr.capture(image, st, real)
    tr.train_step_graphed(image, st, real)
    assert torch.equal(own_eval(), fresh_eval())
    with torch.no_grad():
        nG.head.weight.mul_(1.5)
    assert torch.equal(own_eval(), fresh_eval())
