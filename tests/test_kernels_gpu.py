"""GPU: every kernel of libs2p_sm100.so, called through the C ABI, against a plain
torch fp32 restatement of the same op on the same (bf16-rounded) inputs."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from s2p_b200 import ops as o
    return o


def nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


CONV_CASES = [
    # b, h, w, cin, cout, k, stride, act, bias, residual, stats
    (1, 25, 25, 64, 64, 1, 1, 0, False, False, False),
    (1, 25, 25, 64, 64, 3, 1, 0, False, False, False),
    (2, 25, 25, 128, 128, 3, 1, 2, True, False, False),
    (2, 50, 50, 64, 256, 3, 1, 1, True, False, True),
    (2, 25, 25, 256, 192, 3, 1, 0, True, True, True),
    (3, 100, 100, 128, 64, 3, 1, 0, True, True, True),
    (2, 20, 36, 64, 128, 3, 1, 0, True, False, True),      # ragged: partial TMA boxes
    (1, 7, 9, 64, 64, 3, 1, 0, True, False, True),         # image smaller than a tile
    (4, 25, 25, 512, 1024, 3, 1, 0, True, False, False),   # 4 n-tiles, 72 k-blocks
    (1, 100, 100, 64, 128, 3, 2, 0, True, False, True),    # stride 2 (encoder)
    (2, 50, 50, 128, 256, 3, 2, 0, True, False, True),
    (2, 50, 50, 512, 256, 1, 1, 0, False, False, False),   # 1x1 shortcut
    (160, 25, 25, 64, 64, 3, 1, 0, True, False, True),     # > 148 tiles per CTA wave: persistent loop
]


@pytest.mark.parametrize("case", CONV_CASES, ids=lambda c: "-".join(map(str, c)))
def test_conv_igemm(ops, case):
    b, h, w, cin, cout, k, stride, act, bias, residual, stats = case
    g = torch.Generator().manual_seed(cin * 7 + cout)
    x = torch.randn(b, cin, h, w, generator=g).cuda()
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).cuda()
    bs = torch.randn(cout, generator=g).cuda() if bias else None
    xb = nhwc(x).to(torch.bfloat16)
    wp = ops.pack_conv_weight(wt)
    pad = k // 2
    ho, wo = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    res = torch.randn(b, ho, wo, cout, generator=g).cuda().to(torch.bfloat16) if residual else None
    st = torch.full((b, ops.conv_stats_rows(ho, wo), cout, 2), float("nan"), device="cuda") if stats else None
    y = ops.conv_igemm(xb, wp, bs, cout, k, stride, act, res, None, st)
    ref = F.conv2d(xb.float().permute(0, 3, 1, 2),
                   wp.float().view(cout, k, k, cin).permute(0, 3, 1, 2), bs, stride, pad)
    ref = nhwc(ref)
    if residual:
        ref = ref + res.float()
    ref = F.relu(ref) if act == 1 else F.leaky_relu(ref, 0.2) if act == 2 else ref
    # fp32 accumulation of exact bf16 products: only the final bf16 rounding differs (rel 2^-8)
    tol = 2.0 ** -8 * ref.abs().max().item() + 1e-3
    assert torch.isfinite(y.float()).all()
    assert (y.float() - ref).abs().max().item() <= tol
    if stats:
        yf = y.float()
        mr = ops.in_stats_finalize(st, ho * wo)
        assert torch.allclose(mr[..., 0], yf.mean(dim=(1, 2)), atol=1e-4)
        rstd = 1.0 / torch.sqrt(yf.var(dim=(1, 2), unbiased=False) + 1e-5)
        assert torch.allclose(mr[..., 1], rstd, rtol=1e-4, atol=1e-5)


def test_conv_channel_sliced_views(ops):
    """input = channel slice of a wider tensor, output = slice of a wider tensor."""
    g = torch.Generator().manual_seed(3)
    wide = torch.randn(2, 25, 25, 256, generator=g).cuda().to(torch.bfloat16)
    x = wide[..., 128:192]
    wt = (torch.randn(128, 64, 3, 3, generator=g) / 24).cuda()
    out_wide = torch.zeros(2, 25, 25, 384, dtype=torch.bfloat16, device="cuda")
    y = ops.conv_igemm(x, ops.pack_conv_weight(wt), None, 128, 3, 1, 0, None, out_wide[..., 256:384])
    ref = nhwc(F.conv2d(x.float().permute(0, 3, 1, 2), wt.to(torch.bfloat16).float(), None, 1, 1))
    assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-3
    assert out_wide[..., :256].abs().max().item() == 0


@pytest.mark.parametrize("b,h,cin,cin2,cout,act,stats", [(2, 25, 512, 512, 256, 0, True), (3, 50, 128, 256, 128, 0, True),
                                                      (2, 100, 64, 128, 64, 2, False), (1, 20, 64, 64, 192, 0, True)])
def test_conv_second_operand(ops, b, h, cin, cin2, cout, act, stats):
    """3x3 conv with a 1x1 conv over a second tensor accumulated into the same tile (the res-block's learned
    shortcut riding along with conv_1) == the two convolutions summed in fp32."""
    g = torch.Generator().manual_seed(cin + cin2 + cout)
    xb = torch.randn(b, h, h, cin, generator=g).cuda().to(torch.bfloat16)
    x2 = torch.randn(b, h, h, 2 * cin2, generator=g).cuda().to(torch.bfloat16)[..., cin2:]      # channel-sliced view
    w3 = (torch.randn(cout, cin, 3, 3, generator=g) / (9 * cin) ** 0.5).cuda()
    w1 = (torch.randn(cout, cin2, 1, 1, generator=g) / cin2 ** 0.5).cuda()
    bs = torch.randn(cout, generator=g).cuda()
    wp = torch.cat([ops.pack_conv_weight(w3), ops.pack_conv_weight(w1)], dim=1).contiguous()
    st = torch.full((b, ops.conv_stats_rows(h, h), cout, 2), float("nan"), device="cuda") if stats else None
    y = ops.conv_igemm(xb, wp, bs, cout, 3, 1, act, None, None, st, x2=x2)
    ref = F.conv2d(xb.float().permute(0, 3, 1, 2), w3.to(torch.bfloat16).float(), bs, 1, 1) \
        + F.conv2d(x2.float().permute(0, 3, 1, 2), w1.to(torch.bfloat16).float(), None, 1, 0)
    ref = nhwc(ref)
    ref = F.leaky_relu(ref, 0.2) if act == 2 else ref
    assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-3
    if stats:
        mr = ops.in_stats_finalize(st, h * h)
        assert torch.allclose(mr[..., 0], y.float().mean(dim=(1, 2)), atol=1e-4)
    with pytest.raises(RuntimeError):      # second operand of the wrong spatial size
        ops.conv_igemm(xb, wp, bs, cout, 3, 1, act, None, None, None, x2=x2[:, 1:])


def test_conv_bit_stable(ops):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(4, 50, 50, 128, generator=g).cuda().to(torch.bfloat16)
    wp = ops.pack_conv_weight((torch.randn(256, 128, 3, 3, generator=g) / 34).cuda())
    rows = ops.conv_stats_rows(50, 50)
    outs = []
    for _ in range(3):
        st = torch.empty(4, rows, 256, 2, device="cuda")
        y = ops.conv_igemm(x, wp, None, 256, 3, 1, 0, None, None, st)
        outs.append((y.clone(), st.clone()))
    for y, st in outs[1:]:
        assert torch.equal(y, outs[0][0]) and torch.equal(st, outs[0][1])


def test_conv_rejects_bad_shapes(ops):
    x = torch.zeros(1, 25, 25, 48, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError):
        ops.conv_igemm(x, torch.zeros(64, 9 * 48, dtype=torch.bfloat16, device="cuda"), None, 64, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.conv_igemm(torch.zeros(1, 25, 25, 64, dtype=torch.bfloat16),
                       torch.zeros(64, 9 * 64, dtype=torch.bfloat16), None, 64, 3)


@pytest.mark.parametrize("c,up,hx", [(64, 1, 100), (128, 2, 50), (256, 2, 25), (512, 1, 25), (512, 2, 25)])
def test_mat_norm(ops, c, up, hx):
    g = torch.Generator().manual_seed(c + up)
    b = 3
    x = (torch.randn(b, hx, hx, c, generator=g) * 2 + 0.5).cuda().to(torch.bfloat16)
    h = hx * up
    xf = x.float()
    mean = xf.mean(dim=(1, 2))
    rstd = 1 / torch.sqrt(xf.var(dim=(1, 2), unbiased=False) + 1e-5)
    mr = torch.stack([mean, rstd], dim=-1).contiguous()
    gsbs = (torch.randn(b, 2 * c + 64, generator=g) * 0.3).cuda()
    gs, bs = gsbs[:, :c], gsbs[:, c:2 * c]
    gb = (torch.randn(b, h, h, 2 * c, generator=g) * 0.3).cuda().to(torch.bfloat16)
    xu = xf.repeat_interleave(up, dim=1).repeat_interleave(up, dim=2)
    xh = (xu - mean[:, None, None]) * rstd[:, None, None]
    full = xh * (1 + gs[:, None, None] + gb[..., :c].float()) + bs[:, None, None] + gb[..., c:].float()
    for act, fn in ((2, lambda t: F.leaky_relu(t, 0.2)), (0, lambda t: t)):
        y = ops.mat_norm(x, mr, gs, bs, gb, up, act)
        ref = fn(full)
        assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-3
    y = ops.mat_norm(x, mr, None, None, None, up, 2)
    ref = F.leaky_relu(xh, 0.2)
    assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-3


@pytest.mark.parametrize("c,up,hx,b", [(64, 1, 100, 2), (128, 2, 50, 2), (256, 2, 25, 3), (512, 1, 25, 3),
                                       (128, 1, 20, 2), (32, 1, 9, 1)])
def test_conv_mat_fused(ops, c, up, hx, b):
    """gamma/beta conv with the MAT norm in its epilogue == conv_igemm -> mat_norm restated in
    torch fp32 (oracle MAT.forward arithmetic on bf16-rounded operands)."""
    g = torch.Generator().manual_seed(c * 3 + up)
    nh = 128
    h = hx * up
    x = (torch.randn(b, hx, hx, c, generator=g) * 2 + 0.5).cuda().to(torch.bfloat16)
    hid = torch.relu(torch.randn(b, h, h, 2 * nh, generator=g)).cuda().to(torch.bfloat16)[..., nh:]   # channel slice
    wg = (torch.randn(c, nh, 3, 3, generator=g) / (9 * nh) ** 0.5).cuda()
    wb = (torch.randn(c, nh, 3, 3, generator=g) / (9 * nh) ** 0.5).cuda()
    bias = (torch.randn(2 * c, generator=g) * 0.2).cuda()
    gsbs = (torch.randn(b, 2 * c + 8, generator=g) * 0.3).cuda()
    gs, bs = gsbs[:, :c], gsbs[:, c:2 * c]
    # statistics partials as a producing conv would emit them: one row holding the full sums
    xf = x.float()
    st = torch.zeros(b, 3, c, 2, device="cuda")
    st[:, 1, :, 0] = xf.sum(dim=(1, 2))
    st[:, 1, :, 1] = (xf * xf).sum(dim=(1, 2))
    tab = ops.mat_prepare(st, hx * hx, gs, bs, bias)
    mean = xf.mean(dim=(1, 2))
    rstd = 1 / torch.sqrt(xf.var(dim=(1, 2), unbiased=False) + 1e-5)
    assert torch.allclose(tab[..., 0], rstd, rtol=1e-4, atol=1e-6)
    assert torch.allclose(tab[..., 1], -mean * rstd, rtol=1e-4, atol=1e-4)
    assert torch.allclose(tab[..., 2], 1 + gs + bias[:c], atol=1e-6)
    assert torch.allclose(tab[..., 3], bs + bias[c:], atol=1e-6)
    wp = ops.pack_conv_weight(ops.interleave_gamma_beta(wg, wb))
    hf = hid.float().permute(0, 3, 1, 2)
    gi = nhwc(F.conv2d(hf, wg.to(torch.bfloat16).float(), None, 1, 1))
    bi = nhwc(F.conv2d(hf, wb.to(torch.bfloat16).float(), None, 1, 1))
    xu = xf.repeat_interleave(up, dim=1).repeat_interleave(up, dim=2)
    xh = (xu - mean[:, None, None]) * rstd[:, None, None]
    full = xh * (1 + gs[:, None, None] + bias[:c] + gi) + bs[:, None, None] + bias[c:] + bi
    for act, fn in ((2, lambda t: F.leaky_relu(t, 0.2)), (0, lambda t: t)):
        y = ops.conv_mat_fused(hid, wp, x, tab, up, act)
        ref = fn(full)
        assert torch.isfinite(y.float()).all()
        assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 2e-3
    y2 = ops.conv_mat_fused(hid, wp, x, tab, up, 0)
    assert torch.equal(y, y2)          # bit-stable


WGRAD_CASES = [  # b, h, w, cin, cout, k, stride
    (1, 16, 16, 64, 128, 1, 1), (1, 16, 16, 64, 128, 3, 1), (2, 25, 25, 128, 128, 3, 1), (2, 100, 100, 64, 64, 3, 1),
    (2, 50, 50, 128, 256, 3, 1), (3, 100, 100, 64, 128, 3, 2), (2, 25, 25, 512, 512, 3, 1), (2, 50, 50, 256, 128, 1, 1),
    (1, 7, 9, 64, 64, 3, 1), (2, 20, 36, 64, 72, 3, 1),
]


@pytest.mark.parametrize("case", WGRAD_CASES, ids=lambda c: "-".join(map(str, c)))
def test_conv_wgrad(ops, case):
    """dW of the conv against torch autograd on the same bf16-rounded x and dy (fp32 accumulate:
    only the summation order differs)."""
    b, h, w, cin, cout, k, st = case
    g = torch.Generator().manual_seed(cin + cout + k)
    pad = k // 2
    x = torch.randn(b, h, w, cin, generator=g).cuda().to(torch.bfloat16)
    ho, wo = (h + 2 * pad - k) // st + 1, (w + 2 * pad - k) // st + 1
    dy = torch.randn(b, ho, wo, cout, generator=g).cuda().to(torch.bfloat16)
    dw = ops.conv_wgrad(x, dy, k, st)
    wt = torch.zeros(cout, cin, k, k, device="cuda", requires_grad=True)
    F.conv2d(x.float().permute(0, 3, 1, 2), wt, None, st, pad).backward(dy.float().permute(0, 3, 1, 2))
    ref = wt.grad
    assert torch.isfinite(dw).all()
    assert (dw - ref).abs().max().item() <= 2e-5 * ref.abs().max().item() + 1e-3
    dw2 = ops.conv_wgrad(x, dy, k, st, out=dw.clone(), accumulate=True)
    assert torch.allclose(dw2, 2 * dw, rtol=1e-6, atol=1e-6)
    assert torch.equal(ops.conv_wgrad(x, dy, k, st), dw)      # bit-stable split-K


def test_stem_and_tail(ops):
    g = torch.Generator().manual_seed(11)
    img = (torch.rand(3, 3, 100, 100, generator=g) * 2 - 1).cuda()
    wt = (torch.randn(64, 3, 3, 3, generator=g) / 5).cuda()
    bs = torch.randn(64, generator=g).cuda()
    st = torch.zeros(3, ops.stem_stats_rows(100, 100), 64, 2, device="cuda")
    y = ops.stem_conv(img, wt, bs, 64, None, st)
    ref = nhwc(F.conv2d(img, wt, bs, 1, 1))
    assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-3
    mr = ops.in_stats_finalize(st, 10000)
    assert torch.allclose(mr[..., 0], y.float().mean(dim=(1, 2)), atol=1e-4)
    u8 = torch.randint(0, 256, (2, 100, 100, 3), dtype=torch.uint8, generator=g).cuda()
    y8 = ops.stem_conv(u8, wt, bs, 64)
    ref8 = nhwc(F.conv2d(u8.permute(0, 3, 1, 2).float() / 127.5 - 1, wt, bs, 1, 1))
    assert (y8.float() - ref8).abs().max().item() <= 2.0 ** -8 * ref8.abs().max().item() + 1e-3
    xt = torch.randn(2, 100, 100, 64, generator=g).cuda().to(torch.bfloat16)
    w3 = (torch.randn(3, 64, 3, 3, generator=g) / 24).cuda()
    b3 = (torch.randn(3, generator=g) * 0.1).cuda()
    of = torch.empty(2, 3, 100, 100, device="cuda")
    o8 = torch.empty(2, 100, 100, 3, dtype=torch.uint8, device="cuda")
    ops.torgb_tanh(xt, w3, b3, of, o8)
    ref = torch.tanh(F.conv2d(F.leaky_relu(xt.float().permute(0, 3, 1, 2), 0.2), w3, b3, 1, 1))
    assert (of - ref).abs().max().item() <= 2e-5
    r8 = torch.floor((ref + 1) * 127.5 + 0.5).clamp(0, 255).permute(0, 2, 3, 1)
    assert (o8.float() - r8).abs().max().item() <= 1.0   # ties at x.5 may round either way
    assert (o8.float() - r8).abs().mean().item() <= 1e-3


def test_torgb_igemm(ops):
    g = torch.Generator().manual_seed(17)
    xt = F.leaky_relu(torch.randn(3, 100, 100, 64, generator=g), 0.2).cuda().to(torch.bfloat16)
    w3 = (torch.randn(3, 64, 3, 3, generator=g) / 24).cuda()
    b3 = (torch.randn(3, generator=g) * 0.1).cuda()
    wp = ops.pack_torgb_weight(w3)
    of = torch.empty(3, 3, 100, 100, device="cuda")
    o8 = torch.empty(3, 100, 100, 3, dtype=torch.uint8, device="cuda")
    ops.torgb_igemm(xt, wp, b3, of, o8)
    ref = torch.tanh(F.conv2d(xt.float().permute(0, 3, 1, 2), w3.to(torch.bfloat16).float(), b3, 1, 1))
    assert (of - ref).abs().max().item() <= 2e-5
    r8 = torch.floor((ref + 1) * 127.5 + 0.5).clamp(0, 255).permute(0, 2, 3, 1)
    assert (o8.float() - r8).abs().max().item() <= 1.0
    assert (o8.float() - r8).abs().mean().item() <= 1e-3
    # ragged size + only one of the outputs
    xs = torch.randn(2, 20, 36, 128, generator=g).cuda().to(torch.bfloat16)
    w5 = (torch.randn(3, 128, 3, 3, generator=g) / 34).cuda()
    o = ops.torgb_igemm(xs, ops.pack_torgb_weight(w5), None)
    ref = torch.tanh(F.conv2d(xs.float().permute(0, 3, 1, 2), w5.to(torch.bfloat16).float(), None, 1, 1))
    assert (o - ref).abs().max().item() <= 2e-5


def test_state_encoder_kernels(ops):
    g = torch.Generator().manual_seed(13)
    s = (torch.randn(33, 17, generator=g) * 10).cuda()
    pe = ops.pos_encode(s, 10)
    ref = torch.cat([s] + [f(s * 2.0 ** k) for k in range(10) for f in (torch.sin, torch.cos)], dim=-1)
    assert pe.shape == (33, 357) and (pe - ref).abs().max().item() <= 2e-6
    x = torch.randn(70, 357, generator=g).cuda()
    w = (torch.randn(130, 357, generator=g) / 19).cuda()
    b = torch.randn(130, generator=g).cuda()
    y = ops.linear(x, w, b, 2)
    assert (y - F.leaky_relu(F.linear(x, w, b), 0.2)).abs().max().item() <= 1e-4
