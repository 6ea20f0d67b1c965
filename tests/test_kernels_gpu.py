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
    st = torch.full((b, ops.conv_stats_rows(ho, wo, cout, k, stride), cout, 2), float("nan"), device="cuda") if stats else None
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
    st = torch.full((b, ops.conv_stats_rows(h, h, cout, 3, 1), cout, 2), float("nan"), device="cuda") if stats else None
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
    rows = ops.conv_stats_rows(50, 50, 256, 3, 1)
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
                                       (128, 1, 20, 2), (64, 1, 9, 1), (64, 2, 18, 3), (128, 2, 7, 1), (256, 1, 50, 150)])
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
    # CTA-pair kernel (Cout % 256 == 0, Cin % 128 == 0, 3x3 / 1x1): 256 co x 128 ci per pair, pixel tiles <= 64 rows
    (2, 100, 100, 128, 512, 3, 1), (3, 50, 50, 256, 256, 1, 1), (1, 7, 9, 128, 256, 3, 1), (2, 100, 100, 128, 256, 3, 2),
    (2, 25, 25, 256, 1024, 3, 1), (130, 25, 25, 128, 256, 3, 1),
]


@pytest.mark.parametrize("case", WGRAD_CASES, ids=lambda c: "-".join(map(str, c)))
def test_conv_wgrad(ops, case):
    """dW of the conv against torch autograd IN FLOAT64 on the same bf16-rounded x and dy (fp32 accumulation of exact
    bf16 products here: measured 1e-7 .. 4e-6 of the largest gradient; torch's own fp32 backward is 10x further from
    float64 at the large-K shapes, so it cannot be the yardstick)."""
    b, h, w, cin, cout, k, st = case
    g = torch.Generator().manual_seed(cin + cout + k)
    pad = k // 2
    x = torch.randn(b, h, w, cin, generator=g).cuda().to(torch.bfloat16)
    ho, wo = (h + 2 * pad - k) // st + 1, (w + 2 * pad - k) // st + 1
    dy = torch.randn(b, ho, wo, cout, generator=g).cuda().to(torch.bfloat16)
    dw = ops.conv_wgrad(x, dy, k, st)
    wt = torch.zeros(cout, cin, k, k, device="cuda", dtype=torch.float64, requires_grad=True)
    F.conv2d(x.double().permute(0, 3, 1, 2), wt, None, st, pad).backward(dy.double().permute(0, 3, 1, 2))
    ref = wt.grad
    assert torch.isfinite(dw).all()
    assert (dw.double() - ref).abs().max().item() <= 2e-5 * ref.abs().max().item() + 1e-4
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
    # frames narrower than every candidate tile (tile_w in {30, 14, 6}: the row pitch must divide 32), odd sizes
    for b, hh, ww in ((1, 5, 5), (2, 3, 9), (1, 33, 31), (3, 64, 64)):
        xs = torch.randn(b, hh, ww, 64, generator=g).cuda().to(torch.bfloat16)
        o = ops.torgb_igemm(xs, wp, b3)
        ref = torch.tanh(F.conv2d(xs.float().permute(0, 3, 1, 2), w3.to(torch.bfloat16).float(), b3, 1, 1))
        assert (o - ref).abs().max().item() <= 2e-5, (b, hh, ww)


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


# --------------------------------------------------------------------------------------------------------------
# wide (bf16 hi + lo pair) activations and the split-precision convolution ("high" precision schedule)
def _split(t):
    hi = t.to(torch.bfloat16)
    return hi, (t - hi.float()).to(torch.bfloat16)


@pytest.mark.parametrize("b,h,cin,cout,stride,residual", [(2, 25, 256, 512, 1, False), (2, 25, 128, 512, 1, True),
                                                        (3, 100, 64, 128, 2, False), (2, 50, 64, 64, 1, True)])
def test_conv_wide_output_and_residual(ops, b, h, cin, cout, stride, residual):
    """y stored as a bf16 pair: hi + lo reproduces the fp32 result to ~2^-16 (not 2^-8), hi alone is the plain
    kernel's output bit for bit, the statistics are those of the fp32 values, and a wide residual is added whole."""
    g = torch.Generator().manual_seed(cin + cout + h)
    xb = torch.randn(b, h, h, cin, generator=g).cuda().to(torch.bfloat16)
    wt = (torch.randn(cout, cin, 3, 3, generator=g) / (9 * cin) ** 0.5).cuda()
    bs = torch.randn(cout, generator=g).cuda()
    wp = ops.pack_conv_weight(wt)
    ho = (h + 2 - 3) // stride + 1
    res_hi = res_lo = None
    res32 = None
    if residual:
        res32 = torch.randn(b, ho, ho, cout, generator=g).cuda() * 3
        pair = torch.empty(b, ho, ho, 2 * cout, dtype=torch.bfloat16, device="cuda")
        pair[..., :cout], pair[..., cout:] = _split(res32)
        res_hi, res_lo = ops.wide(pair)
    out = torch.full((b, ho, ho, 2 * cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    y_hi, y_lo = ops.wide(out)
    st = torch.full((b, ops.conv_stats_rows(ho, ho, cout, 3, stride), cout, 2), float("nan"), device="cuda")
    ops.conv_igemm(xb, wp, bs, cout, 3, stride, 0, res_hi, y_hi, st, out_lo=y_lo, residual_lo=res_lo)
    ref = nhwc(F.conv2d(xb.float().permute(0, 3, 1, 2), wt.to(torch.bfloat16).float(), bs, stride, 1))
    if residual:
        ref = ref + (res_hi.float() + res_lo.float())
    y = y_hi.float() + y_lo.float()
    scale = ref.abs().max().item()
    assert torch.isfinite(y).all()
    assert (y - ref).abs().max().item() <= 2.0 ** -15 * scale + 2e-5 * scale      # fp32 accumulation order + 2^-17 pair rounding
    assert torch.equal(y_hi, ref.to(torch.bfloat16)) or (y_hi.float() - ref).abs().max().item() <= 2.0 ** -8 * scale
    mr = ops.in_stats_finalize(st, ho * ho)
    assert torch.allclose(mr[..., 0], ref.mean(dim=(1, 2)), atol=2e-5 * scale)
    assert torch.allclose(mr[..., 1], 1.0 / torch.sqrt(ref.var(dim=(1, 2), unbiased=False) + 1e-5), rtol=1e-4)
    # the plain (narrow) call on the same operands stores exactly the high halves
    if not residual:
        y_narrow = ops.conv_igemm(xb, wp, bs, cout, 3, stride, 0)
        assert torch.equal(y_narrow, y_hi)


@pytest.mark.parametrize("b,h,cin,cout,stride", [(2, 100, 64, 128, 2), (2, 50, 128, 256, 2), (2, 25, 256, 512, 1)])
def test_conv_split_precision(ops, b, h, cin, cout, stride):
    """operand [hi | lo | hi] x weights [W_hi | W_hi | W_lo]: the fp32 convolution of UN-rounded inputs to ~2^-15
    (the image encoder's main-line convs in the "high" precision schedule)"""
    g = torch.Generator().manual_seed(cin * 3 + cout)
    x = torch.randn(b, h, h, cin, generator=g).cuda()
    wt = (torch.randn(cout, cin, 3, 3, generator=g) / (9 * cin) ** 0.5).cuda()
    hi, lo = _split(x)
    pair = torch.cat([hi, lo], dim=3).contiguous()          # the wide operand, walked as [hi | lo | hi] by the kernel
    wp = ops.pack_conv_weight_split(wt)
    assert tuple(wp.shape) == (cout, 9 * 3 * cin)
    ho = (h + 2 - 3) // stride + 1
    out = torch.empty(b, ho, ho, 2 * cout, dtype=torch.bfloat16, device="cuda")
    y_hi, y_lo = ops.wide(out)
    ops.conv_igemm(pair, wp, None, cout, 3, stride, 0, None, y_hi, None, out_lo=y_lo, split=True)
    # the same thing with the third segment materialised is the same bytes
    out3 = torch.empty_like(out)
    ops.conv_igemm(torch.cat([hi, lo, hi], dim=3).contiguous(), wp, None, cout, 3, stride, 0, None, out3[..., :cout], None,
                   out_lo=out3[..., cout:])
    assert torch.equal(out, out3)
    ref = nhwc(F.conv2d(x.permute(0, 3, 1, 2), wt, None, stride, 1))
    plain = nhwc(F.conv2d(hi.float().permute(0, 3, 1, 2), wt.to(torch.bfloat16).float(), None, stride, 1))
    err = (y_hi.float() + y_lo.float() - ref).pow(2).mean().sqrt().item()
    err_plain = (plain - ref).pow(2).mean().sqrt().item()
    print(f"split conv rms error {err:.3e} vs single-bf16 operands {err_plain:.3e}")
    assert err <= err_plain / 50 and err <= 3e-5 * ref.abs().max().item()


@pytest.mark.parametrize("c,up,hx,gb", [(64, 1, 100, False), (128, 1, 50, False), (256, 2, 25, True), (512, 1, 25, True)])
def test_mat_norm_wide_in_and_split_out(ops, c, up, hx, gb):
    """mat_norm reading a wide x and writing a wide y (the [hi | lo] split-precision conv operand)"""
    g = torch.Generator().manual_seed(c + up + 11)
    b = 2
    x32 = (torch.randn(b, hx, hx, c, generator=g) * 2 + 0.5).cuda()
    pair = torch.empty(b, hx, hx, 2 * c, dtype=torch.bfloat16, device="cuda")
    pair[..., :c], pair[..., c:] = _split(x32)
    x_hi, x_lo = ops.wide(pair)
    xw = x_hi.float() + x_lo.float()
    mean = xw.mean(dim=(1, 2))
    rstd = 1.0 / torch.sqrt(xw.var(dim=(1, 2), unbiased=False) + 1e-5)
    mr = torch.stack([mean, rstd], dim=2).contiguous()
    h = hx * up
    gs = bs_ = gbm = None
    if gb:
        gs, bs_ = torch.randn(b, c, generator=g).cuda() * 0.3, torch.randn(b, c, generator=g).cuda() * 0.3
        gbm = (torch.randn(b, h, h, 2 * c, generator=g) * 0.3).cuda().to(torch.bfloat16)
    out3 = torch.full((b, h, h, 2 * c), float("nan"), dtype=torch.bfloat16, device="cuda")
    ops.mat_norm(x_hi, mr, gs, bs_, gbm, up, 2, out3[..., :c], x_lo=x_lo, out_lo=out3[..., c:])
    xu = xw.repeat_interleave(up, 1).repeat_interleave(up, 2)
    xh = (xu - mean[:, None, None]) * rstd[:, None, None]
    if gb:
        ref = xh * (1 + gs[:, None, None] + gbm[..., :c].float()) + bs_[:, None, None] + gbm[..., c:].float()
    else:
        ref = xh
    ref = F.leaky_relu(ref, 0.2)
    y = out3[..., :c].float() + out3[..., c:].float()
    assert (y - ref).abs().max().item() <= 1e-5 * ref.abs().max().item() + 2e-5
    # a narrow input with the same high halves differs at the bf16 level: the low halves are really read
    y_narrow = ops.mat_norm(x_hi, mr, gs, bs_, gbm, up, 2)
    assert (y_narrow.float() - ref).abs().max().item() > 10 * (y - ref).abs().max().item()


@pytest.mark.parametrize("c,up,h", [(64, 1, 100), (128, 2, 100), (512, 1, 25), (256, 2, 50), (64, 2, 36), (128, 1, 7)])
def test_conv_mat_fused_wide_x(ops, c, up, h):
    """the fused gamma/beta conv + MAT norm epilogue with a wide pre-norm activation == mat_norm on the same pair"""
    g = torch.Generator().manual_seed(c + up + h)
    b, nh = 2, 128
    hx = h // up
    hs = F.relu(torch.randn(b, h, h, nh, generator=g)).cuda().to(torch.bfloat16)
    wg = (torch.randn(c, nh, 3, 3, generator=g) / (9 * nh) ** 0.5).cuda()
    wb = (torch.randn(c, nh, 3, 3, generator=g) / (9 * nh) ** 0.5).cuda()
    x32 = (torch.randn(b, hx, hx, c, generator=g) * 2 + 0.5).cuda()
    pair = torch.empty(b, hx, hx, 2 * c, dtype=torch.bfloat16, device="cuda")
    pair[..., :c], pair[..., c:] = _split(x32)
    x_hi, x_lo = ops.wide(pair)
    xw = x_hi.float() + x_lo.float()
    mean = xw.mean(dim=(1, 2))
    rstd = 1.0 / torch.sqrt(xw.var(dim=(1, 2), unbiased=False) + 1e-5)
    gs, bs_ = torch.randn(b, c, generator=g).cuda() * 0.3, torch.randn(b, c, generator=g).cuda() * 0.3
    tab = torch.stack([rstd, -mean * rstd, 1 + gs, bs_], dim=2).contiguous()
    y = ops.conv_mat_fused(hs, ops.pack_conv_weight(ops.interleave_gamma_beta(wg, wb)), x_hi, tab, up, 2, 3, x_lo=x_lo)
    gi = nhwc(F.conv2d(hs.float().permute(0, 3, 1, 2), wg.to(torch.bfloat16).float(), None, 1, 1))
    bi = nhwc(F.conv2d(hs.float().permute(0, 3, 1, 2), wb.to(torch.bfloat16).float(), None, 1, 1))
    xu = xw.repeat_interleave(up, 1).repeat_interleave(up, 2)
    xh = xu * rstd[:, None, None] - (mean * rstd)[:, None, None]
    ref = F.leaky_relu(xh * (1 + gs[:, None, None] + gi) + bs_[:, None, None] + bi, 0.2)
    assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-3
    # rms error is the output's own bf16 rounding only (ulp/sqrt(12) with ulp/|y| in (2^-8, 2^-7]: ~1.66e-3 of the
    # rms value), not two roundings: reading only the high halves of x adds its 2^-9 rounding on top
    rel = ((y.float() - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()).item()
    y_narrow = ops.conv_mat_fused(hs, ops.pack_conv_weight(ops.interleave_gamma_beta(wg, wb)), x_hi, tab, up, 2, 3)
    rel_narrow = ((y_narrow.float() - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()).item()
    print(f"fused MAT rel rms error: wide x {rel:.3e}, high halves only {rel_narrow:.3e}")
    assert rel <= 1.9e-3 and rel_narrow >= 1.15 * rel


def test_stem_wide_output(ops):
    g = torch.Generator().manual_seed(8)
    img = (torch.rand(2, 3, 100, 100, generator=g) * 2 - 1).cuda()
    wt = (torch.randn(64, 3, 3, 3, generator=g) / 27 ** 0.5).cuda()
    bs = torch.randn(64, generator=g).cuda() * 0.1
    out = torch.empty(2, 100, 100, 128, dtype=torch.bfloat16, device="cuda")
    y_hi, y_lo = ops.wide(out)
    st = torch.empty(2, ops.stem_stats_rows(100, 100), 64, 2, device="cuda")
    ops.stem_conv(img, wt, bs, 64, y_hi, st, out_lo=y_lo)
    ref = nhwc(F.conv2d(img, wt, bs, 1, 1))
    assert (y_hi.float() + y_lo.float() - ref).abs().max().item() <= 3e-5 * ref.abs().max().item()
    mr = ops.in_stats_finalize(st, 100 * 100)
    assert torch.allclose(mr[..., 0], ref.mean(dim=(1, 2)), atol=1e-5)
    narrow = ops.stem_conv(img, wt, bs, 64)
    assert torch.equal(narrow, y_hi)


@pytest.mark.parametrize("b,s_dim", [(1, 17), (2, 17), (16, 17), (33, 24), (64, 17), (200, 24)])
def test_state_encoder_fused_is_bit_identical_to_the_chain(ops, b, s_dim):
    """the one-launch state path (persistent cooperative kernel, grid barrier between layers) == pos_encode + 9
    linear launches, BIT FOR BIT (same split-K slices and summation order), and close to torch fp32"""
    g = torch.Generator().manual_seed(b + s_dim)
    state = (torch.randn(b, s_dim, generator=g) * 3).cuda()
    k0 = s_dim * 21
    layers = []
    for i in range(8):
        k = k0 if i == 0 else 512
        layers.append(((torch.randn(512, k, generator=g) / k ** 0.5).cuda(), (torch.randn(512, generator=g) * 0.1).cuda()))
    hw = (torch.randn(8576, 512, generator=g) / 512 ** 0.5).cuda()
    hb = (torch.randn(8576, generator=g) * 0.1).cuda()
    h = ops.pos_encode(state, 10)
    ref32 = h
    for w, bias in layers:
        h = ops.linear(h, w, bias, 2)
        ref32 = F.leaky_relu(F.linear(ref32, w, bias), 0.2)
    chain = ops.linear(h, hw, hb, 0)
    ref32 = F.linear(ref32, hw, hb)
    from s2p_b200 import _lib
    l0 = _lib.launch_count()
    fused = ops.state_encoder_fused(state, layers, hw, hb, 10)
    assert _lib.launch_count() - l0 == 1
    torch.cuda.synchronize()
    assert torch.equal(fused, chain)
    assert (fused - ref32).abs().max().item() <= 2e-4 * ref32.abs().max().item()
    # the barrier counters are left zeroed: a second launch on the same workspace gives the same bytes
    assert torch.equal(ops.state_encoder_fused(state, layers, hw, hb, 10), chain)


@pytest.mark.parametrize("b,h,cin,cout,stride,k", [(2, 100, 64, 128, 2, 3), (2, 25, 256, 512, 1, 3), (3, 50, 128, 384, 1, 3), (2, 25, 256, 64, 1, 1)])
def test_conv_fp16_operands(ops, b, h, cin, cout, stride, k):
    """fp16 activation x fp16 weights (tcgen05 kind::f16, format F16): exact products of the fp16-rounded operands,
    and ~8x closer to the un-rounded fp32 convolution than bf16 operands"""
    g = torch.Generator().manual_seed(cin + cout + k)
    x = torch.randn(b, h, h, cin, generator=g).cuda()
    wt = (torch.randn(cout, cin, k, k, generator=g) / (k * k * cin) ** 0.5).cuda()
    bs = torch.randn(cout, generator=g).cuda()
    xh = x.to(torch.float16)
    wp = ops.pack_conv_weight(wt, dtype=torch.float16)
    assert wp.dtype == torch.float16 and torch.equal(wp.view(cout, k, k, cin).permute(0, 3, 1, 2), wt.to(torch.float16))
    ho = (h + 2 * (k // 2) - k) // stride + 1
    out = torch.empty(b, ho, ho, 2 * cout, dtype=torch.bfloat16, device="cuda")
    y_hi, y_lo = ops.wide(out)
    ops.conv_igemm(xh, wp, bs, cout, k, stride, 0, None, y_hi, None, out_lo=y_lo)
    y = y_hi.float() + y_lo.float()
    ref16 = nhwc(F.conv2d(xh.float().permute(0, 3, 1, 2), wt.to(torch.float16).float(), bs, stride, k // 2))
    ref32 = nhwc(F.conv2d(x.permute(0, 3, 1, 2), wt, bs, stride, k // 2))
    refbf = nhwc(F.conv2d(x.to(torch.bfloat16).float().permute(0, 3, 1, 2), wt.to(torch.bfloat16).float(), bs, stride, k // 2))
    scale = ref32.abs().max().item()
    assert (y - ref16).abs().max().item() <= 3e-5 * scale
    e16 = (y - ref32).pow(2).mean().sqrt().item()
    ebf = (refbf - ref32).pow(2).mean().sqrt().item()
    print(f"fp16 operands rms error {e16:.3e} vs bf16 operands {ebf:.3e}")
    assert e16 <= ebf / 6
    with pytest.raises(RuntimeError):          # operands must agree on the dtype
        ops.conv_igemm(xh, ops.pack_conv_weight(wt), bs, cout, k, stride)


def test_mat_norm_fp16_output(ops):
    g = torch.Generator().manual_seed(41)
    b, h, c = 2, 50, 128
    x32 = (torch.randn(b, h, h, c, generator=g) * 2 + 0.5).cuda()
    pair = torch.empty(b, h, h, 2 * c, dtype=torch.bfloat16, device="cuda")
    pair[..., :c], pair[..., c:] = _split(x32)
    x_hi, x_lo = ops.wide(pair)
    xw = x_hi.float() + x_lo.float()
    mean, var = xw.mean(dim=(1, 2)), xw.var(dim=(1, 2), unbiased=False)
    mr = torch.stack([mean, 1.0 / torch.sqrt(var + 1e-5)], dim=2).contiguous()
    out = torch.empty(b, h, h, c, dtype=torch.float16, device="cuda")
    ops.mat_norm(x_hi, mr, act=2, out=out, x_lo=x_lo)
    ref = F.leaky_relu((xw - mean[:, None, None]) * mr[:, None, None, :, 1], 0.2)
    assert torch.equal(out, ref.to(torch.float16)) or (out.float() - ref).abs().max().item() <= 2.0 ** -11 * ref.abs().max().item() + 1e-6


@pytest.mark.parametrize("case", [
    # b, h, w, cin, cout, act, bias, residual, stats, x2 channels
    (2, 100, 100, 128, 64, 2, True, False, True, 0), (3, 100, 100, 64, 64, 0, True, False, False, 128),
    (2, 50, 50, 64, 64, 0, True, True, True, 0), (2, 25, 25, 256, 64, 1, False, False, True, 0),
    (1, 7, 9, 64, 64, 0, True, False, True, 64), (2, 20, 36, 64, 64, 0, True, False, True, 0),
    (150, 25, 25, 64, 64, 0, True, False, True, 0)], ids=lambda c: "-".join(map(str, c)))
def test_conv_horizontal_halo_mode(ops, case):
    """Cout <= 64, 3x3 stride 1: the conv runs in the horizontal-halo layout (one activation box per filter row
    feeds its three taps through row-shifted descriptors; M rows in pitch tile_w + 2, halo rows discarded): same
    results as the plain geometry -- fp32 accumulation of exact bf16 products -- incl. statistics, residual, ragged
    sizes, the second (1x1) operand and the persistent loop"""
    b, h, w, cin, cout, act, bias, residual, stats, c2 = case
    g = torch.Generator().manual_seed(cin + cout + h + c2)
    xb = torch.randn(b, h, w, cin, generator=g).cuda().to(torch.bfloat16)
    wt = (torch.randn(cout, cin, 3, 3, generator=g) / (9 * cin) ** 0.5).cuda()
    bs = torch.randn(cout, generator=g).cuda() if bias else None
    res = torch.randn(b, h, w, cout, generator=g).cuda().to(torch.bfloat16) if residual else None
    wp = ops.pack_conv_weight(wt)
    x2 = w1 = None
    if c2:
        x2 = torch.randn(b, h, w, c2, generator=g).cuda().to(torch.bfloat16)
        w1 = (torch.randn(cout, c2, 1, 1, generator=g) / c2 ** 0.5).cuda()
        wp = torch.cat([wp, ops.pack_conv_weight(w1)], dim=1).contiguous()
    rows = ops.conv_stats_rows(h, w, cout, 3, 1)
    st = torch.full((b, rows, cout, 2), float("nan"), device="cuda") if stats else None
    y = ops.conv_igemm(xb, wp, bs, cout, 3, 1, act, res, None, st, x2=x2)
    ref = F.conv2d(xb.float().permute(0, 3, 1, 2), wt.to(torch.bfloat16).float(), bs, 1, 1)
    if c2:
        ref = ref + F.conv2d(x2.float().permute(0, 3, 1, 2), w1.to(torch.bfloat16).float(), None, 1, 0)
    ref = nhwc(ref)
    if residual:
        ref = ref + res.float()
    ref = F.relu(ref) if act == 1 else F.leaky_relu(ref, 0.2) if act == 2 else ref
    assert torch.isfinite(y.float()).all()
    assert (y.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-3
    if stats:
        mr = ops.in_stats_finalize(st, h * w)
        assert torch.isfinite(st).all()
        assert torch.allclose(mr[..., 0], y.float().mean(dim=(1, 2)), atol=1e-4)
        assert torch.allclose(mr[..., 1], 1.0 / torch.sqrt(y.float().var(dim=(1, 2), unbiased=False) + 1e-5), rtol=1e-4, atol=1e-5)
    assert torch.equal(ops.conv_igemm(xb, wp, bs, cout, 3, 1, act, res, None, None, x2=x2), y)      # bit-stable


@pytest.mark.parametrize("shape", [(4, 6, 100, 100), (2, 3, 7, 9), (1, 1, 1, 1), (3, 2, 50, 51)])
def test_avgpool3s2_matches_torch(ops, shape):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(*shape, generator=g).cuda().requires_grad_(True)
    ref = F.avg_pool2d(x, 3, 2, 1, count_include_pad=False)
    dy = torch.randn(*ref.shape, generator=g).cuda()
    ref.backward(dy)
    from s2p_b200 import autograd as A
    x2 = x.detach().clone().requires_grad_(True)
    out = A.avgpool3s2(x2)
    out.backward(dy)
    assert out.shape == ref.shape and (out - ref).abs().max().item() <= 1e-6
    assert (x2.grad - x.grad).abs().max().item() <= 1e-6


def test_colsum_rows(ops):
    g = torch.Generator().manual_seed(4)
    for rows, cols in ((128, 512), (1, 17), (333, 70), (16, 8576)):
        x = torch.randn(rows, cols, generator=g).cuda()
        got = ops.colsum_rows(x)
        ref = x.double().sum(0).float()
        assert (got - ref).abs().max().item() <= 1e-5 * max(1.0, ref.abs().max().item())
    xs = torch.randn(64, 100, generator=g).cuda()[:, :40]          # a column slice: pitch 100
    assert (ops.colsum_rows(xs) - xs.double().sum(0).float()).abs().max().item() <= 1e-5


@pytest.mark.parametrize("b,hw,cin,cout,k,act,bias", [(2, 50, 64, 256, 3, 1, True), (1, 25, 128, 192, 3, 2, True),
                                                      (2, 100, 64, 64, 3, 0, False), (3, 37, 64, 128, 1, 1, True),
                                                      (2, 51, 64, 128, 4, 2, True)])
def test_simple_epilogue_kernel_is_bit_identical_to_the_general_one(ops, b, hw, cin, cout, k, act, bias):
    # bias / activation-only launches run a kernel of their own (kEpiSimple); asking for statistics selects the general
    # epilogue: same accumulators, same roundings -> the stored bf16 values must be the same bits
    g = torch.Generator().manual_seed(23)
    x = torch.randn(b, hw, hw, cin, generator=g).cuda().to(torch.bfloat16)
    w = ops.pack_conv_weight((torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).cuda())
    bv = torch.randn(cout, generator=g).cuda() if bias else None
    stride, pad = (2, 2) if k == 4 else (1, k // 2)
    y_simple = ops.conv_igemm(x, w, bv, cout, k, stride, act, pad=pad)
    ho = y_simple.shape[1]
    st = torch.empty(b, ops.conv_stats_rows(ho, ho, cout, k, stride), cout, 2, device="cuda")
    y_general = ops.conv_igemm(x, w, bv, cout, k, stride, act, stats=st, pad=pad)
    assert torch.equal(y_simple, y_general)
