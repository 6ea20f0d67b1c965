"""Tensor-level wrappers over the C ABI (one function per exported kernel).

PyTorch is plumbing here: it owns device memory and the stream; every function
passes raw pointers + sizes to libs2p_sm100.so.  Activations are NHWC bf16
tensors ``[B, H, W, C]``; a channel slice ``t[..., a:b]`` of such a tensor is a
valid argument (the pixel pitch is read from ``t.stride(2)``).
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import ACT_LRELU, ACT_NONE, ACT_RELU, IMG_F32_NCHW, IMG_U8_NHWC  # noqa: F401


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return 0 if t is None else t.data_ptr()


def _require_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("s2p_b200 ops need CUDA tensors: there is no CPU path (use oracle/ for a CPU run)")


def _nhwc_pitch(t: torch.Tensor, what: str, f16_ok: bool = False) -> int:
    """pixel pitch of an NHWC (possibly channel-sliced) bf16 tensor (fp16 where the callee takes an fp16 operand)"""
    if (t.dtype != torch.bfloat16 and not (f16_ok and t.dtype == torch.float16)) or t.dim() != 4:
        raise RuntimeError(f"{what}: expected a 4-D bf16 NHWC tensor, got {tuple(t.shape)} {t.dtype}")
    b, h, w, c = t.shape
    pitch = t.stride(2)
    if t.stride(3) != 1 or t.stride(1) != pitch * w or (b > 1 and t.stride(0) != pitch * w * h):
        raise RuntimeError(f"{what}: tensor is not pixel-contiguous NHWC (strides {t.stride()})")
    return pitch


def conv_stats_rows(hout: int, wout: int, cout: int = 256, ksize: int = 3, stride: int = 1) -> int:
    """partial-statistics rows per image written by the conv with this output size / width / filter (the library
    picks the tile geometry from them; the default arguments describe a wide conv in the standard geometry)"""
    return int(_lib.load().s2p_conv_stats_rows(hout, wout, cout, ksize, stride))


def stem_stats_rows(h: int, w: int) -> int:
    return int(_lib.load().s2p_stem_stats_rows(h, w))


def pack_conv_weight(w_oihw: torch.Tensor, out: torch.Tensor | None = None, scale: float = 1.0,
                     dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """OIHW fp32 -> [Cout, R*S*Cin] bf16 (tap-major, channel-minor); dtype=torch.float16 for the fp16-operand convs."""
    _require_cuda(w_oihw)
    w = w_oihw.detach().contiguous().float()
    co, ci, r, s = w.shape
    if out is None:
        out = torch.empty(co, r * s * ci, dtype=dtype, device=w.device)
    assert out.is_contiguous() and out.numel() == w.numel() and out.dtype in (torch.bfloat16, torch.float16)
    _lib.check(_lib.load().s2p_pack_conv_weight(w.data_ptr(), out.data_ptr(), co, ci, r, s, float(scale),
                                                1 if out.dtype == torch.float16 else 0, _stream()),
               "s2p_pack_conv_weight")
    return out


def fits_f16(w: torch.Tensor) -> bool:
    """can these fp32 values be fp16 tensor-core operands (finite after rounding)?  One host sync, at pack time."""
    return bool(torch.isfinite(w.detach().float().to(torch.float16)).all())


def _lo_of(t, lo, what):
    """pointer of the low halves of a wide (bf16 hi + lo pair) tensor: same shape and pixel pitch as the high halves"""
    if lo is None:
        return 0
    if tuple(lo.shape) != tuple(t.shape) or _nhwc_pitch(lo, what) != _nhwc_pitch(t, what):
        raise RuntimeError(f"{what}: the low halves must have the shape and pitch of the high halves")
    return lo.data_ptr()


def wide(t):
    """(hi, lo) channel slices of a wide activation tensor [B, H, W, 2C] = [hi | lo]"""
    c = t.shape[3] // 2
    return t[..., :c], t[..., c:]


def pack_conv_weight_split(w_oihw: torch.Tensor) -> torch.Tensor:
    """packed weight of the split-precision conv: [Cout, R*S*3*Cin] bf16 with [W_hi | W_hi | W_lo] along Cin
    (W_hi = bf16(W), W_lo = bf16(W - W_hi)), for a wide operand [x_hi | x_lo] walked as [x_hi | x_lo | x_hi]
    (conv_igemm split=True): one tensor-core pass computes W_hi x_hi + W_hi x_lo + W_lo x_hi, i.e. (W x) to
    ~2^-16 instead of 2^-8."""
    w = w_oihw.detach().float()
    hi = w.to(torch.bfloat16).float()
    return pack_conv_weight(torch.cat([hi, hi, w - hi], dim=1))


def conv_igemm(x, w_packed, bias, cout, ksize, stride=1, act=ACT_NONE, residual=None, out=None, stats=None, pad=None,
               x2=None, out_lo=None, residual_lo=None, split=False):
    """y = act(conv(x) + bias + residual); returns y.  ``stats`` (fp32
    [B, conv_stats_rows, Cout, 2]) receives (sum, sumsq) partials when given.  ``x2`` (NHWC bf16 of the output's
    spatial size): a second operand whose 1x1 convolution is accumulated into the same tile; ``w_packed`` is then
    [Cout, k*k*Cin + Cin2] (the two packed weights concatenated along the row).  ``out_lo`` / ``residual_lo``: low
    halves of a wide output / residual (bf16 hi + lo pair, see include/s2p_b200.h).  ``split``: x is a wide tensor
    [x_hi | x_lo] (2C channels) and ``w_packed`` comes from pack_conv_weight_split (split-precision convolution)."""
    _require_cuda(x, w_packed, bias, residual, out, stats, x2, out_lo, residual_lo)
    b, hin, win, cin = x.shape
    cin_wrap = 0
    if split:
        if cin % 128:
            raise RuntimeError("conv_igemm: a split operand is [hi | lo] with C % 64 == 0 channels each")
        cin_wrap, cin = cin, cin // 2 * 3
    pad = ksize // 2 if pad is None else pad
    hout = (hin + 2 * pad - ksize) // stride + 1
    wout = (win + 2 * pad - ksize) // stride + 1
    if out is None:
        out = torch.empty(b, hout, wout, cout, dtype=torch.bfloat16, device=x.device)
    if tuple(out.shape) != (b, hout, wout, cout):
        raise RuntimeError(f"conv_igemm: out shape {tuple(out.shape)} != {(b, hout, wout, cout)}")
    cin2 = 0
    if x2 is not None:
        cin2 = x2.shape[3]
        if tuple(x2.shape[:3]) != (b, hout, wout):
            raise RuntimeError(f"conv_igemm: x2 must have the output's spatial size, got {tuple(x2.shape)}")
    op_f16 = x.dtype == torch.float16        # fp16 operands: activation(s) and packed weight all fp16
    if w_packed.dtype != x.dtype or (x2 is not None and x2.dtype != x.dtype):
        raise RuntimeError(f"conv_igemm: operands disagree on dtype (x {x.dtype}, w {w_packed.dtype})")
    if w_packed.numel() != cout * (ksize * ksize * cin + cin2) or not w_packed.is_contiguous():
        raise RuntimeError("conv_igemm: packed weight has the wrong size")
    if bias is not None and (bias.dtype != torch.float32 or bias.numel() != cout or not bias.is_contiguous()):
        raise RuntimeError("conv_igemm: bias must be contiguous fp32 [Cout]")
    if stats is not None:
        rows = conv_stats_rows(hout, wout, cout, ksize, stride)
        if stats.dtype != torch.float32 or stats.numel() != b * rows * cout * 2 or not stats.is_contiguous():
            raise RuntimeError("conv_igemm: stats buffer has the wrong size")
    res_pitch = 0
    if residual is not None:
        if tuple(residual.shape) != tuple(out.shape):
            raise RuntimeError("conv_igemm: residual shape mismatch")
        res_pitch = _nhwc_pitch(residual, "residual")
    _lib.check(
        _lib.load().s2p_conv_igemm_fwd2(
            x.data_ptr(), _nhwc_pitch(x, "x", True), _ptr(x2), _nhwc_pitch(x2, "x2", True) if x2 is not None else 0, cin2,
            w_packed.data_ptr(), _ptr(bias), _ptr(residual), _lo_of(residual, residual_lo, "residual_lo"), res_pitch,
            out.data_ptr(), _lo_of(out, out_lo, "out_lo"), _nhwc_pitch(out, "out"), _ptr(stats), b, hin, win, cin, cout,
            ksize, stride, pad, act, cin_wrap, 1 if op_f16 else 0, _stream()),
        "s2p_conv_igemm_fwd2")
    return out


_wgrad_ws = {}
_retired_ws = []


def _grow_workspace(cache, device, need, dtype, floor=1):
    """per-device scratch shared by successive launches on one stream.  It only ever grows (at least doubling), and
    an outgrown buffer is kept alive instead of freed: a captured CUDA graph of the training step holds its raw
    pointer, and freed memory handed to another tensor would be scribbled on by the next replay."""
    ws = cache.get(device)
    if ws is None or ws.numel() < need:
        if ws is not None:
            _retired_ws.append(ws)
        ws = torch.empty(max(need, floor, 2 * ws.numel() if ws is not None else 0), dtype=dtype, device=device)
        cache[device] = ws
    return ws


def conv_wgrad(x, dy, ksize, stride=1, pad=None, out=None, accumulate=False):
    """dW (fp32 OIHW) of conv(x, W) given dy; x, dy NHWC bf16."""
    _require_cuda(x, dy, out)
    b, hin, win, cin = x.shape
    cout = dy.shape[3]
    pad = ksize // 2 if pad is None else pad
    if out is None:
        out = torch.empty(cout, cin, ksize, ksize, dtype=torch.float32, device=x.device)
        accumulate = False
    if tuple(out.shape) != (cout, cin, ksize, ksize) or out.dtype != torch.float32 or not out.is_contiguous():
        raise RuntimeError("conv_wgrad: out must be contiguous fp32 [Cout, Cin, k, k]")
    lib = _lib.load()
    need = int(lib.s2p_conv_wgrad_workspace_bytes(b, hin, win, cin, cout, ksize, stride, pad))
    if need < 0:
        raise RuntimeError("conv_wgrad: unsupported shape")
    ws = _grow_workspace(_wgrad_ws, x.device, need, torch.uint8, 1 << 20)
    _lib.check(lib.s2p_conv_wgrad(x.data_ptr(), _nhwc_pitch(x, "x"), dy.data_ptr(), _nhwc_pitch(dy, "dy"), out.data_ptr(),
                                  ws.data_ptr(), ws.numel(), b, hin, win, cin, cout, ksize, stride, pad,
                                  1 if accumulate else 0, _stream()), "s2p_conv_wgrad")
    return out


def in_stats_finalize(stats: torch.Tensor, hw: int, eps: float = 1e-5, out=None) -> torch.Tensor:
    """partials [B, rows, C, 2] -> (mean, rstd) [B, C, 2] fp32"""
    _require_cuda(stats)
    b, rows, c, _ = stats.shape
    if out is None:
        out = torch.empty(b, c, 2, dtype=torch.float32, device=stats.device)
    _lib.check(_lib.load().s2p_in_stats_finalize(stats.data_ptr(), rows, out.data_ptr(), b, c, hw, float(eps), _stream()),
               "s2p_in_stats_finalize")
    return out


def mat_norm(x, mean_rstd, gs=None, bs=None, gb_spatial=None, up=1, act=ACT_LRELU, out=None, x_lo=None, out_lo=None):
    """y = act(((x_up - mean) * rstd) * (1 + gs + gI) + bs + bI); gs/bs are [B, C]
    fp32 views (row pitch taken from stride(0)); gb_spatial is NHWC bf16 with 2C
    channels (gI | bI) at the OUTPUT resolution.  ``x_lo``: low halves of a wide x; ``out_lo``: low halves of a
    wide output (with ``out`` the two halves of one tensor: the [hi | lo] split-precision conv operand)."""
    _require_cuda(x, mean_rstd, gs, bs, gb_spatial, out, x_lo, out_lo)
    b, hx, wx, c = x.shape
    h, w = hx * up, wx * up
    if out is None:
        out = torch.empty(b, h, w, c, dtype=torch.bfloat16, device=x.device)
    if tuple(out.shape) != (b, h, w, c):
        raise RuntimeError("mat_norm: out shape mismatch")
    y_f16 = out.dtype == torch.float16       # an fp16 output: the operand of an fp16-operand conv
    if tuple(mean_rstd.shape) != (b, c, 2) or mean_rstd.dtype != torch.float32 or not mean_rstd.is_contiguous():
        raise RuntimeError("mat_norm: mean_rstd must be contiguous fp32 [B, C, 2]")
    gpitch = 0
    if gs is not None:
        if gs.dtype != torch.float32 or tuple(gs.shape) != (b, c) or gs.stride(1) != 1 or bs.stride() != gs.stride():
            raise RuntimeError("mat_norm: gs/bs must be fp32 [B, C] views with equal strides")
        gpitch = gs.stride(0)
    gbp = 0
    if gb_spatial is not None:
        if tuple(gb_spatial.shape) != (b, h, w, 2 * c):
            raise RuntimeError("mat_norm: gb_spatial must be [B, H, W, 2C]")
        gbp = _nhwc_pitch(gb_spatial, "gb_spatial")
    _lib.check(
        _lib.load().s2p_mat_norm_fwd(
            x.data_ptr(), _lo_of(x, x_lo, "x_lo"), _nhwc_pitch(x, "x"), up, mean_rstd.data_ptr(), _ptr(gs), _ptr(bs), gpitch,
            _ptr(gb_spatial), gbp, out.data_ptr(), _lo_of(out, out_lo, "out_lo"), _nhwc_pitch(out, "out", True), b, h, w, c, act,
            1 if y_f16 else 0, _stream()),
        "s2p_mat_norm_fwd")
    return out


def mat_norm_bwd(dy, x, mean_rstd, gs=None, bs=None, gb_spatial=None, up=1, act=ACT_LRELU, add=None):
    """backward of mat_norm: returns (dx [B,H/up,W/up,C] bf16 (+add), d_gb [B,H,W,2C] bf16 or None,
    sums [B,C,4] fp32 = (d gs, d bs, S1, S2))."""
    _require_cuda(dy, x, mean_rstd, gs, bs, gb_spatial, add)
    b, h, w, c = dy.shape
    hx, wx = h // up, w // up
    if tuple(x.shape) != (b, hx, wx, c):
        raise RuntimeError("mat_norm_bwd: x shape mismatch")
    dev = dy.device
    lib = _lib.load()
    chunks = int(lib.s2p_mat_norm_bwd_chunks(b, h, w, c))
    d_gb = torch.empty(b, h, w, 2 * c, dtype=torch.bfloat16, device=dev) if gb_spatial is not None else None
    dxh = torch.empty(b, h, w, c, dtype=torch.bfloat16, device=dev)
    partial = torch.empty(b, chunks, c, 4, dtype=torch.float32, device=dev)
    sums = torch.empty(b, c, 4, dtype=torch.float32, device=dev)
    dx = torch.empty(b, hx, wx, c, dtype=torch.bfloat16, device=dev)
    gpitch = gs.stride(0) if gs is not None else 0
    _lib.check(
        lib.s2p_mat_norm_bwd(
            dy.data_ptr(), _nhwc_pitch(dy, "dy"), x.data_ptr(), _nhwc_pitch(x, "x"), up, mean_rstd.data_ptr(), _ptr(gs),
            _ptr(bs), gpitch, _ptr(gb_spatial), 0 if gb_spatial is None else _nhwc_pitch(gb_spatial, "gb"), act,
            _ptr(d_gb), 2 * c, dxh.data_ptr(), c, partial.data_ptr(), sums.data_ptr(), _ptr(add),
            0 if add is None else _nhwc_pitch(add, "add"), dx.data_ptr(), c, b, h, w, c, _stream()),
        "s2p_mat_norm_bwd")
    return dx, d_gb, sums


def mat_fused_ct(c: int) -> int:
    return int(_lib.load().s2p_mat_fused_ct(c))


def interleave_gamma_beta(w_gamma: torch.Tensor, w_beta: torch.Tensor) -> torch.Tensor:
    """[C, ...] gamma rows and [C, ...] beta rows -> [2C, ...] in the tile-interleaved row
    order s2p_conv_mat_fused_fwd expects (ct gamma rows, ct beta rows, ...)."""
    c = w_gamma.shape[0]
    ct = mat_fused_ct(c)
    parts = []
    for j in range(0, c, ct):
        parts += [w_gamma[j:j + ct], w_beta[j:j + ct]]
    return torch.cat(parts, dim=0).contiguous()


def mat_prepare(stats, hw, gs, bs, bias_gb, eps: float = 1e-5, out=None):
    """partials [B, rows, C, 2] + state gamma/beta [B, C] + conv bias [2C] -> table [B, C, 4] fp32"""
    _require_cuda(stats, gs, bs, bias_gb, out)
    b, rows, c, _ = stats.shape
    if out is None:
        out = torch.empty(b, c, 4, dtype=torch.float32, device=stats.device)
    if tuple(out.shape) != (b, c, 4) or out.dtype != torch.float32 or not out.is_contiguous():
        raise RuntimeError("mat_prepare: table must be contiguous fp32 [B, C, 4]")
    gpitch = 0
    if gs is not None:
        if gs.dtype != torch.float32 or tuple(gs.shape) != (b, c) or gs.stride(1) != 1 or bs.stride() != gs.stride():
            raise RuntimeError("mat_prepare: gs/bs must be fp32 [B, C] views with equal strides")
        gpitch = gs.stride(0)
    if bias_gb is not None and (bias_gb.dtype != torch.float32 or bias_gb.numel() != 2 * c or not bias_gb.is_contiguous()):
        raise RuntimeError("mat_prepare: bias_gb must be contiguous fp32 [2C]")
    _lib.check(_lib.load().s2p_mat_prepare(stats.data_ptr(), rows, _ptr(gs), _ptr(bs), gpitch, _ptr(bias_gb),
                                           out.data_ptr(), b, c, hw, float(eps), _stream()), "s2p_mat_prepare")
    return out


def conv_mat_fused(h, w_packed_gb, x, mod_table, up=1, act=ACT_LRELU, ksize=3, out=None, x_lo=None):
    """y = act(IN(x_up) * (1 + gs + conv_gamma(h)) + bs + conv_beta(h)) in one kernel: the
    gamma/beta conv's accumulators are consumed by its own epilogue.  ``x_lo``: low halves of a wide x."""
    _require_cuda(h, w_packed_gb, x, mod_table, out, x_lo)
    b, hh, ww, cin = h.shape
    bx, hx, wx, c = x.shape
    if bx != b or hx * up != hh or wx * up != ww:
        raise RuntimeError(f"conv_mat_fused: x {tuple(x.shape)} (up={up}) does not match h {tuple(h.shape)}")
    if out is None:
        out = torch.empty(b, hh, ww, c, dtype=torch.bfloat16, device=h.device)
    if tuple(out.shape) != (b, hh, ww, c):
        raise RuntimeError("conv_mat_fused: out shape mismatch")
    if w_packed_gb.dtype != torch.bfloat16 or w_packed_gb.numel() != 2 * c * ksize * ksize * cin or not w_packed_gb.is_contiguous():
        raise RuntimeError("conv_mat_fused: packed weight has the wrong size/dtype")
    if tuple(mod_table.shape) != (b, c, 4) or mod_table.dtype != torch.float32 or not mod_table.is_contiguous():
        raise RuntimeError("conv_mat_fused: mod_table must be contiguous fp32 [B, C, 4]")
    _lib.check(
        _lib.load().s2p_conv_mat_fused_fwd(
            h.data_ptr(), _nhwc_pitch(h, "h"), w_packed_gb.data_ptr(), x.data_ptr(), _lo_of(x, x_lo, "x_lo"),
            _nhwc_pitch(x, "x"), up,
            mod_table.data_ptr(), out.data_ptr(), _nhwc_pitch(out, "out"), b, hh, ww, cin, c, ksize, act, _stream()),
        "s2p_conv_mat_fused_fwd")
    return out


def stem_conv(image, w_oihw, bias, cout, out=None, stats=None, out_lo=None):
    """Cin=3 3x3 conv from the boundary image: fp32 NCHW [B,3,H,W] or uint8 NHWC [B,H,W,3].  ``out_lo``: low halves
    of a wide output."""
    _require_cuda(image, w_oihw, bias, out, stats, out_lo)
    if image.dtype == torch.uint8:
        b, h, w, _ = image.shape
        fmt = IMG_U8_NHWC
    elif image.dtype == torch.float32:
        b, _, h, w = image.shape
        fmt = IMG_F32_NCHW
    else:
        raise RuntimeError(f"stem_conv: unsupported image dtype {image.dtype}")
    if not image.is_contiguous():
        raise RuntimeError("stem_conv: image must be contiguous")
    if out is None:
        out = torch.empty(b, h, w, cout, dtype=torch.bfloat16, device=image.device)
    _lib.check(
        _lib.load().s2p_stem_conv_fwd(image.data_ptr(), fmt, w_oihw.data_ptr(), _ptr(bias), out.data_ptr(),
                                      _lo_of(out, out_lo, "out_lo"), _nhwc_pitch(out, "out"), _ptr(stats), b, h, w, cout,
                                      _stream()),
        "s2p_stem_conv_fwd")
    return out


def torgb_tanh(x, w_oihw, bias, out_f32=None, out_u8=None):
    _require_cuda(x, w_oihw, bias, out_f32, out_u8)
    b, h, w, cin = x.shape
    if out_f32 is None and out_u8 is None:
        out_f32 = torch.empty(b, 3, h, w, dtype=torch.float32, device=x.device)
    _lib.check(
        _lib.load().s2p_torgb_tanh_fwd(x.data_ptr(), _nhwc_pitch(x, "x"), w_oihw.data_ptr(), _ptr(bias),
                                       _ptr(out_f32), _ptr(out_u8), b, h, w, cin, _stream()),
        "s2p_torgb_tanh_fwd")
    return out_f32 if out_f32 is not None else out_u8


def pack_torgb_weight(w_oihw: torch.Tensor) -> torch.Tensor:
    """[3, Cin, 3, 3] fp32 -> [48, 3*Cin] bf16 for torgb_igemm: row 16*s + c holds tap s (filter column) of output
    channel c (rows 16*s + 3 .. 16*s + 15 zero), column r*Cin + ci filter row r of input channel ci -- the three taps
    of a filter row are 16-column groups of ONE 48-column UMMA (see s2p_torgb_igemm_fwd)."""
    co, ci, r, s_ = w_oihw.shape
    if co > 16 or r != 3 or s_ != 3:
        raise RuntimeError("pack_torgb_weight: a 3x3 filter with at most 16 output channels")
    out = torch.zeros(3, 16, 3, ci, dtype=torch.float32, device=w_oihw.device)
    out[:, :co] = w_oihw.detach().float().permute(3, 0, 2, 1)        # [s, c, r, ci]
    return out.reshape(48, 3 * ci).to(torch.bfloat16).contiguous()


def torgb_igemm(x_activated, w_packed16, bias, out_f32=None, out_u8=None):
    """tanh(conv3x3(x) + bias) on the tensor cores; x is already LeakyReLU'd."""
    _require_cuda(x_activated, w_packed16, bias, out_f32, out_u8)
    b, h, w, cin = x_activated.shape
    if out_f32 is None and out_u8 is None:
        out_f32 = torch.empty(b, 3, h, w, dtype=torch.float32, device=x_activated.device)
    if w_packed16.dtype != torch.bfloat16 or tuple(w_packed16.shape) != (48, 3 * cin) or not w_packed16.is_contiguous():
        raise RuntimeError("torgb_igemm: packed weight must be contiguous bf16 [48, 3*Cin] (pack_torgb_weight)")
    _lib.check(
        _lib.load().s2p_torgb_igemm_fwd(x_activated.data_ptr(), _nhwc_pitch(x_activated, "x"), w_packed16.data_ptr(),
                                        _ptr(bias), _ptr(out_f32), _ptr(out_u8), b, h, w, cin, _stream()),
        "s2p_torgb_igemm_fwd")
    return out_f32 if out_f32 is not None else out_u8


def pos_encode(state, n_freqs, out=None):
    _require_cuda(state, out)
    b, s = state.shape
    d = s * (1 + 2 * n_freqs)
    if out is None:
        out = torch.empty(b, d, dtype=torch.float32, device=state.device)
    _lib.check(_lib.load().s2p_pos_encode_fwd(state.data_ptr(), out.data_ptr(), out.stride(0), b, s, n_freqs, _stream()),
               "s2p_pos_encode_fwd")
    return out


def linear(x, w, bias, act=ACT_NONE, out=None):
    _require_cuda(x, w, bias, out)
    b, k = x.shape
    n = w.shape[0]
    if out is None:
        out = torch.empty(b, n, dtype=torch.float32, device=x.device)
    if x.dtype != torch.float32 or w.dtype != torch.float32 or not w.is_contiguous() or x.stride(1) != 1:
        raise RuntimeError("linear: fp32 row-major operands required")
    lib = _lib.load()
    need = int(lib.s2p_linear_workspace_floats(b, k, n))
    ws = _grow_workspace(_linear_ws, x.device, need, torch.float32) if need > 0 else None
    _lib.check(lib.s2p_linear_fwd(x.data_ptr(), x.stride(0), w.data_ptr(), _ptr(bias), out.data_ptr(),
                                  out.stride(0), b, k, n, act, _ptr(ws), need, _stream()), "s2p_linear_fwd")
    return out


_linear_ws = {}
_state_ws = {}


def state_encoder_fused(state, layers, heads_w, heads_b, n_freqs, out=None):
    """gsbs = heads(MLP(pos_encode(state))) in ONE launch (s2p_state_encoder_fwd); bit-identical to pos_encode +
    linear chained.  layers: [(w [width, K] fp32, b [width] fp32)], heads_w [N, width], heads_b [N]."""
    import ctypes
    _require_cuda(state, heads_w, heads_b, out, *[t for wb in layers for t in wb])
    b, s = state.shape
    width, n = layers[0][0].shape[0], heads_w.shape[0]
    if state.dtype != torch.float32 or not state.is_contiguous():
        raise RuntimeError("state_encoder_fused: contiguous fp32 state required")
    for w, bias in layers:
        if w.dtype != torch.float32 or not w.is_contiguous() or w.shape[0] != width or not bias.is_contiguous():
            raise RuntimeError("state_encoder_fused: contiguous fp32 [width, K] layer weights required")
    if not heads_w.is_contiguous() or heads_w.shape[1] != width or heads_w.dtype != torch.float32:
        raise RuntimeError("state_encoder_fused: heads_w must be contiguous fp32 [N, width]")
    if out is None:
        out = torch.empty(b, n, dtype=torch.float32, device=state.device)
    lib = _lib.load()
    need = int(lib.s2p_state_encoder_workspace_bytes(b, width))
    key = (state.device, torch.cuda.current_stream().cuda_stream)    # one workspace per stream: launches on a stream are ordered
    ws = _state_ws.get(key)
    if ws is None or ws.numel() < need:
        if ws is not None:
            _retired_ws.append(ws)        # a captured CUDA graph may still hold its pointer
        ws = torch.zeros(max(need, 2 * (ws.numel() if ws is not None else 0)), dtype=torch.uint8, device=state.device)
        _state_ws[key] = ws
    k = len(layers)
    wp = (ctypes.c_void_p * k)(*[w.data_ptr() for w, _ in layers])
    bp = (ctypes.c_void_p * k)(*[bias.data_ptr() for _, bias in layers])
    _lib.check(lib.s2p_state_encoder_fwd(state.data_ptr(), b, s, n_freqs, wp, bp, k, width, heads_w.data_ptr(), _ptr(heads_b), n,
                                         out.data_ptr(), out.stride(0), ws.data_ptr(), ws.numel(), _stream()),
               "s2p_state_encoder_fwd")
    return out


# Kernels of this library update parameters through raw pointers (fused Adam, a replayed CUDA graph of the whole
# training step), which torch's per-tensor version counters do not see.  Every such update bumps this epoch, and the
# modules that cache packed bf16 weights (the generator's inference schedule) key their cache on it.
_PARAM_EPOCH = [0]


def bump_param_epoch():
    _PARAM_EPOCH[0] += 1


def param_epoch() -> int:
    return _PARAM_EPOCH[0]


def adam_step_multi(params, grads, exp_avgs, exp_avg_sqs, lr, beta1, beta2, eps, step_dev):
    """torch.optim.Adam update of many fp32 tensors, 64 per launch"""
    import ctypes
    lib = _lib.load()
    for i in range(0, len(params), 64):
        chunk = range(i, min(i + 64, len(params)))
        cnt = len(chunk)
        arr = lambda ts: (ctypes.c_void_p * cnt)(*[ts[j].data_ptr() for j in chunk])
        ns = (ctypes.c_longlong * cnt)(*[params[j].numel() for j in chunk])
        for j in chunk:
            if grads[j].dtype != torch.float32 or not grads[j].is_contiguous() or not params[j].is_contiguous():
                raise RuntimeError("adam_step_multi: contiguous fp32 tensors required")
        _lib.check(lib.s2p_adam_step_multi(arr(params), arr(grads), arr(exp_avgs), arr(exp_avg_sqs), ns, cnt, float(lr),
                                           float(beta1), float(beta2), float(eps), step_dev.data_ptr(), _stream()),
                   "s2p_adam_step_multi")


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, step):
    """in-place torch.optim.Adam update of one fp32 tensor"""
    _require_cuda(param, grad, exp_avg, exp_avg_sq)
    g = grad.contiguous().float()
    if param.dtype != torch.float32 or not param.is_contiguous():
        raise RuntimeError("adam_step: parameters must be contiguous fp32")
    _lib.check(_lib.load().s2p_adam_step(param.data_ptr(), g.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
                                         param.numel(), float(lr), float(beta1), float(beta2), float(eps), int(step),
                                         _stream()), "s2p_adam_step")


def adam_step_dev(param, grad, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, step_dev):
    """Adam update with the step count in a device int32 tensor (graph-capturable)."""
    _require_cuda(param, grad, exp_avg, exp_avg_sq, step_dev)
    g = grad if (grad.is_contiguous() and grad.dtype == torch.float32) else grad.contiguous().float()
    if param.dtype != torch.float32 or not param.is_contiguous() or step_dev.dtype != torch.int32:
        raise RuntimeError("adam_step_dev: contiguous fp32 parameters and an int32 step tensor are required")
    _lib.check(_lib.load().s2p_adam_step_dev(param.data_ptr(), g.data_ptr(), exp_avg.data_ptr(), exp_avg_sq.data_ptr(),
                                             param.numel(), float(lr), float(beta1), float(beta2), float(eps),
                                             step_dev.data_ptr(), _stream()), "s2p_adam_step_dev")


def pack_conv_weights_multi(weights, want_dgrad):
    """[(fwd, dgrad | None)] for a list of OIHW fp32 conv weights: pack_conv_weight / pack_conv_weight_dgrad of all of
    them in one launch per 48 tensors, into ONE bf16 allocation (views are 256-byte aligned)."""
    import ctypes
    lib = _lib.load()
    ws = []
    for w in weights:
        _require_cuda(w)
        w = w.detach()
        if w.dtype != torch.float32 or not w.is_contiguous():
            w = w.contiguous().float()
        ws.append(w)
    if not ws:
        return []
    al = lambda n: (n + 127) // 128 * 128  # noqa: E731
    total = sum(al(w.numel()) * (2 if d else 1) for w, d in zip(ws, want_dgrad))
    flat = torch.empty(total, dtype=torch.bfloat16, device=ws[0].device)
    out, off = [], 0
    for w, d in zip(ws, want_dgrad):
        co, ci, r, s = w.shape
        f = flat[off:off + w.numel()].view(co, r * s * ci)
        off += al(w.numel())
        g = None
        if d:
            g = flat[off:off + w.numel()].view(ci, r * s * co)
            off += al(w.numel())
        out.append((f, g))
    for i in range(0, len(ws), 48):
        ch = range(i, min(i + 48, len(ws)))
        n = len(ch)
        vp = lambda xs: (ctypes.c_void_p * n)(*xs)  # noqa: E731
        ia = lambda k: (ctypes.c_int * n)(*[ws[j].shape[k] for j in ch])  # noqa: E731
        _lib.check(lib.s2p_pack_conv_weights_multi(vp([ws[j].data_ptr() for j in ch]), vp([out[j][0].data_ptr() for j in ch]),
                                                   vp([out[j][1].data_ptr() if out[j][1] is not None else None for j in ch]),
                                                   ia(0), ia(1), ia(2), ia(3), n, _stream()), "s2p_pack_conv_weights_multi")
    return out


def pack_conv_weight_dgrad(w_oihw):
    """OIHW fp32 -> [Cin, R*S*Cout] bf16: the packed weight of the data-gradient conv."""
    _require_cuda(w_oihw)
    w = w_oihw.detach().contiguous().float()
    co, ci, r, s = w.shape
    out = torch.empty(ci, r * s * co, dtype=torch.bfloat16, device=w.device)
    _lib.check(_lib.load().s2p_pack_conv_weight_dgrad(w.data_ptr(), out.data_ptr(), co, ci, r, s, _stream()),
               "s2p_pack_conv_weight_dgrad")
    return out


def act_bwd_bias_grad(dy, y, act, want_bias_grad=True):
    """(dy * act'(y), db): gradient through ReLU/LeakyReLU given the activated output, and the
    bias gradient sum_{pixels} in the same pass.  dy, y NHWC bf16."""
    _require_cuda(dy, y)
    b, h, w, c = dy.shape
    lib = _lib.load()
    out = torch.empty(b, h, w, c, dtype=torch.bfloat16, device=dy.device) if act != ACT_NONE else dy
    db = partial = None
    if want_bias_grad:
        db = torch.empty(c, dtype=torch.float32, device=dy.device)
        partial = torch.empty(int(lib.s2p_act_bwd_bias_grad_chunks()) * c, dtype=torch.float32, device=dy.device)
    elif act == ACT_NONE:
        return dy, None
    _lib.check(lib.s2p_act_bwd_bias_grad(dy.data_ptr(), _nhwc_pitch(dy, "dy"), _ptr(y) if act != ACT_NONE else 0,
                                         _nhwc_pitch(y, "y") if act != ACT_NONE else 0, act,
                                         out.data_ptr() if act != ACT_NONE else 0, c, _ptr(db), _ptr(partial),
                                         b * h * w, c, _stream()), "s2p_act_bwd_bias_grad")
    return out, db


def colsum_rows(x):
    """x.sum(dim=0) of an fp32 [rows, cols] matrix (unit column stride)"""
    _require_cuda(x)
    if x.dtype != torch.float32 or x.dim() != 2 or x.stride(1) != 1:
        raise RuntimeError("colsum_rows: fp32 [rows, cols] with unit column stride required")
    out = torch.empty(x.shape[1], dtype=torch.float32, device=x.device)
    _lib.check(_lib.load().s2p_colsum_rows_f32(x.data_ptr(), x.stride(0), x.shape[0], x.shape[1], out.data_ptr(), _stream()),
               "s2p_colsum_rows_f32")
    return out


def avgpool3s2(x):
    """F.avg_pool2d(x, 3, 2, 1, count_include_pad=False) for contiguous fp32 [..., H, W]"""
    _require_cuda(x)
    if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() < 2:
        raise RuntimeError("avgpool3s2: contiguous fp32 [..., H, W] required")
    h, w = x.shape[-2:]
    y = torch.empty(*x.shape[:-2], (h - 1) // 2 + 1, (w - 1) // 2 + 1, dtype=torch.float32, device=x.device)
    if x.numel():
        _lib.check(_lib.load().s2p_avgpool3s2_fwd(x.data_ptr(), y.data_ptr(), x.numel() // (h * w), h, w, _stream()),
                   "s2p_avgpool3s2_fwd")
    return y


def avgpool3s2_bwd(dy, h, w):
    _require_cuda(dy)
    if dy.dtype != torch.float32 or not dy.is_contiguous():
        raise RuntimeError("avgpool3s2_bwd: contiguous fp32 gradient required")
    dx = torch.empty(*dy.shape[:-2], h, w, dtype=torch.float32, device=dy.device)
    if dx.numel():
        _lib.check(_lib.load().s2p_avgpool3s2_bwd(dy.data_ptr(), dx.data_ptr(), dx.numel() // (h * w), h, w, _stream()),
                   "s2p_avgpool3s2_bwd")
    return dx


def maxpool2(x, out=None):
    """nn.MaxPool2d(2, 2) on NHWC bf16"""
    _require_cuda(x, out)
    b, h, w, c = x.shape
    if out is None:
        out = torch.empty(b, h // 2, w // 2, c, dtype=torch.bfloat16, device=x.device)
    _lib.check(_lib.load().s2p_maxpool2_fwd(x.data_ptr(), _nhwc_pitch(x, "x"), out.data_ptr(), _nhwc_pitch(out, "out"),
                                            b, h, w, c, _stream()), "s2p_maxpool2_fwd")
    return out


def maxpool2_bwd(dy, x):
    """gradient of maxpool2 w.r.t. x (same shape as x, fully written)"""
    _require_cuda(dy, x)
    b, h, w, c = x.shape
    dx = torch.empty(b, h, w, c, dtype=torch.bfloat16, device=x.device)
    _lib.check(_lib.load().s2p_maxpool2_bwd(dy.data_ptr(), _nhwc_pitch(dy, "dy"), x.data_ptr(), _nhwc_pitch(x, "x"),
                                            dx.data_ptr(), _nhwc_pitch(dx, "dx"), b, h, w, c, _stream()), "s2p_maxpool2_bwd")
    return dx


def zero_insert2(dy, hout, wout):
    """[B,Ho,Wo,C] bf16 -> [B,hout,wout,C] with dy at the even positions and zeros elsewhere (stride-2 dgrad input)"""
    _require_cuda(dy)
    b, ho, wo, c = dy.shape
    out = torch.empty(b, hout, wout, c, dtype=torch.bfloat16, device=dy.device)
    if b:
        _lib.check(_lib.load().s2p_zero_insert2(dy.data_ptr(), _nhwc_pitch(dy, "dy"), out.data_ptr(), b, ho, wo, hout, wout, c,
                                                _stream()), "s2p_zero_insert2")
    return out


def nchw_to_nhwc64(x):
    """fp32 NCHW [B,c<=8,H,W] -> bf16 NHWC [B,H,W,64], zero-padded channels"""
    _require_cuda(x)
    if x.dtype != torch.float32 or x.dim() != 4:
        raise RuntimeError("nchw_to_nhwc64: fp32 NCHW tensor required")
    x = x.contiguous()
    b, c, h, w = x.shape
    out = torch.empty(b, h, w, 64, dtype=torch.bfloat16, device=x.device)
    if b:
        _lib.check(_lib.load().s2p_nchw_to_nhwc64(x.data_ptr(), out.data_ptr(), b, c, h, w, _stream()), "s2p_nchw_to_nhwc64")
    return out


def nhwc64_to_nchw(dy, c):
    """the first c channels of a bf16 NHWC tensor as fp32 NCHW (backward of nchw_to_nhwc64)"""
    _require_cuda(dy)
    b, h, w, _ = dy.shape
    out = torch.empty(b, c, h, w, dtype=torch.float32, device=dy.device)
    if b:
        _lib.check(_lib.load().s2p_nhwc64_to_nchw(dy.data_ptr(), _nhwc_pitch(dy, "dy"), out.data_ptr(), b, c, h, w, _stream()),
                   "s2p_nhwc64_to_nchw")
    return out


def tanh_nhwc_to_nchw(x, c):
    """tanh of the first c channels of a bf16 NHWC tensor, as fp32 NCHW"""
    _require_cuda(x)
    b, h, w, _ = x.shape
    out = torch.empty(b, c, h, w, dtype=torch.float32, device=x.device)
    if b:
        _lib.check(_lib.load().s2p_tanh_nhwc_to_nchw(x.data_ptr(), _nhwc_pitch(x, "x"), out.data_ptr(), b, c, h, w, _stream()),
                   "s2p_tanh_nhwc_to_nchw")
    return out


def tanh_bwd_nchw_to_nhwc64(dy, y):
    """dy * (1 - y^2) (fp32 NCHW, <= 8 channels) as bf16 NHWC with 64 zero-padded channels"""
    _require_cuda(dy, y)
    dy, y = dy.contiguous().float(), y.contiguous()
    b, c, h, w = dy.shape
    out = torch.empty(b, h, w, 64, dtype=torch.bfloat16, device=dy.device)
    if b:
        _lib.check(_lib.load().s2p_tanh_bwd_nchw_to_nhwc64(dy.data_ptr(), y.data_ptr(), out.data_ptr(), b, c, h, w, _stream()),
                   "s2p_tanh_bwd_nchw_to_nhwc64")
    return out


def multi_scale_blocks() -> int:
    return int(_lib.load().s2p_multi_scale_blocks())


def _ptr_array(ptrs):
    import ctypes
    return (ctypes.c_void_p * len(ptrs))(*ptrs)


def multi_scale_fwd(src_ptrs, dst_ptrs, numels, scale_idx, scales):
    """dst_i = src_i * scales[scale_idx[i]] for raw fp32 device pointers, 64 tensors per launch"""
    import ctypes
    lib = _lib.load()
    for i in range(0, len(src_ptrs), 64):
        j = min(i + 64, len(src_ptrs))
        _lib.check(lib.s2p_multi_scale_fwd(_ptr_array(src_ptrs[i:j]), _ptr_array(dst_ptrs[i:j]),
                                           (ctypes.c_longlong * (j - i))(*numels[i:j]), (ctypes.c_int * (j - i))(*scale_idx[i:j]),
                                           j - i, scales.data_ptr(), _stream()), "s2p_multi_scale_fwd")


def multi_scale_bwd(dout_ptrs, src_ptrs, dsrc_ptrs, numels, scale_idx, scales, partial):
    """dsrc_i = dout_i * scale; partial[i, :] = per-block sums of dout_i * src_i (fp32 [count, multi_scale_blocks()])"""
    import ctypes
    lib = _lib.load()
    nb = multi_scale_blocks()
    for i in range(0, len(src_ptrs), 64):
        j = min(i + 64, len(src_ptrs))
        _lib.check(lib.s2p_multi_scale_bwd(_ptr_array(dout_ptrs[i:j]), _ptr_array(src_ptrs[i:j]), _ptr_array(dsrc_ptrs[i:j]),
                                           (ctypes.c_longlong * (j - i))(*numels[i:j]), (ctypes.c_int * (j - i))(*scale_idx[i:j]),
                                           j - i, scales.data_ptr(), partial.data_ptr() + i * nb * 4, _stream()),
                   "s2p_multi_scale_bwd")


def l1_mean(a, b):
    """mean |a - b| of two NHWC bf16 tensors (channel slices allowed) -> fp32 scalar tensor"""
    _require_cuda(a, b)
    if a.shape != b.shape:
        raise RuntimeError("l1_mean: shape mismatch")
    bb, h, w, c = a.shape
    lib = _lib.load()
    partial = torch.empty(int(lib.s2p_l1_mean_chunks()), dtype=torch.float32, device=a.device)
    out = torch.empty((), dtype=torch.float32, device=a.device)
    _lib.check(lib.s2p_l1_mean_fwd(a.data_ptr(), _nhwc_pitch(a, "a"), b.data_ptr(), _nhwc_pitch(b, "b"), bb * h * w, c,
                                   partial.data_ptr(), out.data_ptr(), _stream()), "s2p_l1_mean_fwd")
    return out


def l1_mean_bwd(a, b, gout):
    """gradient of l1_mean w.r.t. a: sign(a - b) * gout / numel, NHWC bf16"""
    _require_cuda(a, b, gout)
    bb, h, w, c = a.shape
    da = torch.empty(bb, h, w, c, dtype=torch.bfloat16, device=a.device)
    g = gout.detach().float().contiguous()
    _lib.check(_lib.load().s2p_l1_mean_bwd(a.data_ptr(), _nhwc_pitch(a, "a"), b.data_ptr(), _nhwc_pitch(b, "b"), g.data_ptr(),
                                           bb * h * w, c, da.data_ptr(), _nhwc_pitch(da, "da"), _stream()), "s2p_l1_mean_bwd")
    return da
