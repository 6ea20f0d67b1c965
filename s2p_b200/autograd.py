"""Differentiable wrappers of the C-ABI kernels for the GAN training step
(the reference's train.py --netG=s2p, README.md:59).

torch.autograd is the tape only: every forward and backward below is one of this
repo's sm_100a kernels (conv forward / dgrad = the tcgen05 implicit GEMM, wgrad = the
MN-major tcgen05 kernel, norm backward = the fused reduce + dx kernels).  Activations
and their gradients are NHWC bf16; parameters and their gradients are fp32 in the
nn.Conv2d / nn.Linear layouts, so a stock optimiser state_dict stays compatible.
"""
from __future__ import annotations

import torch

from . import ops
from .ops import ACT_LRELU, ACT_NONE, ACT_RELU


def _act_grad(dy: torch.Tensor, y: torch.Tensor, act: int) -> torch.Tensor:
    """gradient through ReLU / LeakyReLU(0.2) given the activated output y (sign(y) = sign(pre))."""
    if act == ACT_NONE:
        return dy
    if act == ACT_RELU:
        return dy * (y > 0)
    return torch.where(y > 0, dy, dy * 0.2)


class ConvFn(torch.autograd.Function):
    """y = act(conv(x, W) + b + residual); x, residual, y NHWC bf16; W fp32 OIHW, b fp32."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual, stride, act, stats, pad, packed=None):
        cout, cin, k, _ = weight.shape
        pad = k // 2 if pad is None else pad
        ctx.packed = packed          # (fwd, dgrad) packed weights of a FROZEN layer, built once by the caller
        wp = packed[0] if packed is not None else ops.pack_conv_weight(weight)
        y = ops.conv_igemm(x, wp, None if bias is None else bias.detach().float().contiguous(), cout, k, stride, act,
                           residual, None, stats, pad)
        ctx.save_for_backward(x, weight, y if act != ACT_NONE else None)
        ctx.cfg = (k, stride, act, bias is not None, residual is not None, pad)
        ctx.grad_slot = getattr(x, "_s2p_grad_slot", None)     # x is a channel slice made by split_channels
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        k, stride, act, has_bias, has_res, pad = ctx.cfg
        dy = dy.contiguous()
        want_db = has_bias and ctx.needs_input_grad[2]
        # activation gradient and bias gradient in one pass over dy
        dy, db = ops.act_bwd_bias_grad(dy, y, act, want_db)
        dx = dw = dres = None
        if ctx.needs_input_grad[0]:
            cout, cin = weight.shape[0], weight.shape[1]
            # dgrad = the same implicit GEMM with the filter flipped and in/out channels swapped
            g = dy
            if stride == 2:   # transposed conv: zero-insert the gradient, then a stride-1 conv
                g = ops.zero_insert2(dy, x.shape[1] + 2 * pad - k + 1, x.shape[2] + 2 * pad - k + 1)
            wd = ctx.packed[1] if ctx.packed is not None and ctx.packed[1] is not None else ops.pack_conv_weight_dgrad(weight)
            # the gradient of a split_channels slice goes straight into its place in the gradient of the whole tensor
            out = ctx.grad_slot[0].slice_for(ctx.grad_slot[1], x) if ctx.grad_slot is not None else None
            dx = ops.conv_igemm(g, wd, None, cin, k, 1, ACT_NONE, out=out, pad=k - 1 - pad)
        if ctx.needs_input_grad[1]:
            dw = ops.conv_wgrad(x, dy, k, stride, pad)
        if has_res and ctx.needs_input_grad[3]:
            dres = dy
        return dx, dw, db, dres, None, None, None, None, None


def conv(x, weight, bias=None, residual=None, stride=1, act=ACT_NONE, stats=None, pad=None, packed=None):
    return ConvFn.apply(x, weight, bias, residual, stride, act, stats, pad, packed)


class _GradArena:
    """The gradient of a tensor whose channel slices feed separate convs: each conv's data gradient is written by the
    conv kernel into its slice of ONE buffer (the kernel takes a pixel pitch), so the backward of the split returns the
    buffer as it is.  torch.split's backward concatenates the slice gradients: 1.9 GB read + 1.9 GB written for the
    768-channel merged mlp_shared tensor at 100 x 100 (612 us of a batch-128 step, ncu)."""

    def __init__(self, width):
        self.width = width
        self.buf = None

    def slice_for(self, index, like):
        b, h, w, c = like.shape
        if self.buf is None or self.buf.shape[:3] != (b, h, w) or self.buf.device != like.device:
            self.buf = torch.empty(b, h, w, self.width, dtype=like.dtype, device=like.device)
        return self.buf[..., index * c:(index + 1) * c]


class SplitChannelsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, nh):
        n = x.shape[3] // nh
        ctx.arena = _GradArena(x.shape[3])
        ctx.nh, ctx.meta = nh, (x.shape, x.dtype, x.device)
        return tuple(x[..., i * nh:(i + 1) * nh] for i in range(n))

    @staticmethod
    def backward(ctx, *grads):
        arena, nh = ctx.arena, ctx.nh
        buf = arena.buf
        in_place = buf is not None and all(
            g is not None and g.data_ptr() == buf.data_ptr() + i * nh * buf.element_size() and g.stride() == buf[..., :nh].stride()
            and g.shape[3] == nh for i, g in enumerate(grads))
        arena.buf = None                 # one use per backward (a captured graph replays the same allocation)
        if in_place:
            return buf, None
        shape, dtype, device = ctx.meta
        parts = [g if g is not None else torch.zeros(*shape[:3], nh, dtype=dtype, device=device) for g in grads]
        return torch.cat(parts, dim=3), None


def split_channels(x, nh):
    """torch.split(x, nh, dim=3) whose slices carry the slot their gradient goes to (see _GradArena)."""
    outs = SplitChannelsFn.apply(x, nh)
    arena = getattr(outs[0].grad_fn, "arena", None)      # the Function's ctx is the backward node
    if arena is not None:
        for i, o in enumerate(outs):
            o._s2p_grad_slot = (arena, i)
    return outs


class MatNormFn(torch.autograd.Function):
    """y = act(IN(up(x)) * (1 + gs + gI) + bs + bI); statistics come precomputed (mean_rstd)."""

    @staticmethod
    def forward(ctx, x, mean_rstd, gs, bs, gb, up, act):
        y = ops.mat_norm(x, mean_rstd, gs, bs, gb, up, act)
        ctx.save_for_backward(x, mean_rstd, gs, bs, gb)
        ctx.cfg = (up, act)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, mean_rstd, gs, bs, gb = ctx.saved_tensors
        up, act = ctx.cfg
        dx, d_gb, sums = ops.mat_norm_bwd(dy.contiguous(), x, mean_rstd, gs, bs, gb, up, act)
        dgs = sums[..., 0] if gs is not None else None
        dbs = sums[..., 1] if bs is not None else None
        return dx, None, dgs, dbs, d_gb, None, None


def mat_norm(x, mean_rstd, gs=None, bs=None, gb=None, up=1, act=ACT_LRELU):
    return MatNormFn.apply(x, mean_rstd, gs, bs, gb, up, act)


class LinearFn(torch.autograd.Function):
    """y = act(x W^T + b), fp32, on the repo's own SGEMM kernel (state encoder / MAT state heads)."""

    @staticmethod
    def forward(ctx, x, weight, bias, act):
        y = ops.linear(x.contiguous(), weight.contiguous(), bias, act)
        ctx.save_for_backward(x, weight, y if act != ACT_NONE else None)
        ctx.act = act
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y = ctx.saved_tensors
        dy = _act_grad(dy.contiguous(), y, ctx.act) if ctx.act != ACT_NONE else dy.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = ops.linear(dy, weight.t().contiguous(), None, ACT_NONE)             # [B,N] x [N,K]
        if ctx.needs_input_grad[1]:
            dw = ops.linear(dy.t().contiguous(), x.t().contiguous(), None, ACT_NONE)  # [N,B] x [B,K]
        if ctx.needs_input_grad[2]:
            db = ops.colsum_rows(dy)
        return dx, dw, db, None


def linear(x, weight, bias=None, act=ACT_NONE):
    return LinearFn.apply(x, weight, bias, act)


class MultiScaleFn(torch.autograd.Function):
    """Scaled copies of many fp32 parameter tensors written straight into concatenated operand buffers:
    out[g][off_i : off_i + n_i] = src_i * scales[k_i].  `plan` = (group_shapes, [(group, elem_offset, k)] per source,
    onehot [K, n_src] fp32 on the device).  Backward: one multi-tensor launch for the parameter gradients plus
    fixed-order partial dot products for d scales (oracle MAT.forward's `weight * sigmoid(blend)` and its MulBackward)."""

    @staticmethod
    def forward(ctx, scales, plan, *srcs):
        shapes, where, onehot = plan
        dev = scales.device
        outs = [torch.empty(sh, dtype=torch.float32, device=dev) for sh in shapes]
        scales = scales.contiguous()
        srcs_c = [t.detach().contiguous() for t in srcs]
        ops.multi_scale_fwd([t.data_ptr() for t in srcs_c], [outs[g].data_ptr() + 4 * off for g, off, _ in where],
                            [t.numel() for t in srcs_c], [k for _, _, k in where], scales)
        ctx.save_for_backward(scales, *srcs_c)
        ctx.plan = plan
        return tuple(outs)

    @staticmethod
    def backward(ctx, *gouts):
        scales, *srcs = ctx.saved_tensors
        shapes, where, onehot = ctx.plan
        gouts = [torch.zeros(sh, dtype=torch.float32, device=scales.device) if g is None else g.contiguous().float()
                 for g, sh in zip(gouts, shapes)]
        dsrcs = [torch.empty_like(t) for t in srcs]
        partial = torch.empty(len(srcs), ops.multi_scale_blocks(), dtype=torch.float32, device=scales.device)
        ops.multi_scale_bwd([gouts[g].data_ptr() + 4 * off for g, off, _ in where], [t.data_ptr() for t in srcs],
                            [t.data_ptr() for t in dsrcs], [t.numel() for t in srcs], [k for _, _, k in where], scales, partial)
        dscales = onehot @ partial.sum(dim=1)
        return (dscales, None, *dsrcs)


def multi_scale(scales, plan, srcs):
    return MultiScaleFn.apply(scales, plan, *srcs)


class L1MeanFn(torch.autograd.Function):
    """mean |a - b| over NHWC bf16 feature maps; b is a constant (detached target)."""

    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return ops.l1_mean(a, b)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        return ops.l1_mean_bwd(a, b, g), None


def _as_nhwc(t):
    """the NHWC bf16 tensor behind `t` (itself, or an NCHW-permuted view of one), else None"""
    if t.dtype != torch.bfloat16 or t.dim() != 4:
        return None
    for cand in (t, t.permute(0, 2, 3, 1)):
        b, h, w, c = cand.shape
        p = cand.stride(2)
        if cand.stride(3) == 1 and c % 8 == 0 and p % 8 == 0 and cand.stride(1) == p * w and (b == 1 or cand.stride(0) == p * w * h) \
                and cand.data_ptr() % 16 == 0:
            return cand
    return None


def l1_mean(a, b):
    """F.l1_loss(a.float(), b.float()) for bf16 feature maps on the fused kernel; other inputs take the torch path"""
    an, bn = _as_nhwc(a), _as_nhwc(b.detach())
    if an is None or bn is None or an.shape != bn.shape:
        return torch.nn.functional.l1_loss(a.float(), b.detach().float())
    return L1MeanFn.apply(an, bn)


class MaxPool2Fn(torch.autograd.Function):
    """nn.MaxPool2d(2, 2) on NHWC bf16 (VGG19 feature extractor of the perceptual loss)."""

    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return ops.maxpool2(x)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return ops.maxpool2_bwd(dy.contiguous(), x)


def maxpool2(x):
    return MaxPool2Fn.apply(x)


class AvgPool3s2Fn(torch.autograd.Function):
    """F.avg_pool2d(x, 3, stride=2, padding=1, count_include_pad=False) on fp32 NCHW (between the discriminator scales)"""

    @staticmethod
    def forward(ctx, x):
        ctx.hw = tuple(x.shape[-2:])
        return ops.avgpool3s2(x.contiguous())

    @staticmethod
    def backward(ctx, dy):
        return ops.avgpool3s2_bwd(dy.contiguous(), *ctx.hw)


def avgpool3s2(x):
    return AvgPool3s2Fn.apply(x)


class PadChannelsFn(torch.autograd.Function):
    """fp32 NCHW image (<= 8 channels) -> bf16 NHWC with 64 zero-padded channels, the layout of every first conv"""

    @staticmethod
    def forward(ctx, x):
        ctx.c = x.shape[1]
        return ops.nchw_to_nhwc64(x)

    @staticmethod
    def backward(ctx, dy):
        return ops.nhwc64_to_nchw(dy if dy.stride(-1) == 1 else dy.contiguous(), ctx.c)


def pad_channels(x):
    return PadChannelsFn.apply(x)


class TanhOutFn(torch.autograd.Function):
    """image = tanh(rgb[..., :c]) as fp32 NCHW from the 64-channel-padded bf16 NHWC output of the ToRGB conv"""

    @staticmethod
    def forward(ctx, rgb, c):
        if rgb.shape[-1] != 64:
            raise RuntimeError("tanh_out: expects the 64-channel padded ToRGB output")
        y = ops.tanh_nhwc_to_nchw(rgb, c)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return ops.tanh_bwd_nchw_to_nhwc64(dy, y), None


def tanh_out(rgb, c=3):
    return TanhOutFn.apply(rgb, c)


def conv_stats(y_shape, device, ksize=3, stride=1):
    """scratch for the conv epilogue's statistics partials of an output of shape [B,H,W,C] (the row count follows the
    tile geometry of the conv that writes them: ops.conv_stats_rows)"""
    b, h, w, c = y_shape
    return torch.empty(b, ops.conv_stats_rows(h, w, c, ksize, stride), c, 2, dtype=torch.float32, device=device)
