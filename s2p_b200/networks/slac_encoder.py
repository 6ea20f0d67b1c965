"""SLAC image encoder (SURVEY.md section 8 f, rank 3): the first consumer of the generated frames.

Mirror of ``Encoder`` (/root/reference/rlkit/torch/slac/network/latent.py:117-171): same constructor, same
parameter names (``net.0.weight`` ... so ``load_state_dict(torch.load('encoder.pth'))`` of
rlkit/torch/slac/algo.py:149-150 works), ``forward(x[B,S,C,H,W] in [0,1]) -> [B,S,output_dim]``.

B200-first: layer 1 (3->32, 5x5 s2, K = 75) is a CUDA-core kernel that can read the HBM-resident uint8
dataset directly (window gather + /255 fused, ``forward_from_buffer``), every other layer is the tcgen05
implicit-GEMM kernel of the generator with the LeakyReLU(0.2) in its epilogue, and the final 4x4 "valid"
conv on a 4x4 map runs as a 1x1 conv over [frames, 4096] so no MMA row is wasted.  Activations bf16,
accumulation fp32 (the reference is fp32: tolerance stated in tests/test_slac_encoder_gpu.py).
Inference only (the latent-model training loop is outside the hot path).  No CPU path.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from .. import _lib, ops
from ..ops import ACT_LRELU, _stream


def _specs(image_size, input_dim, output_dim):
    """(cin, cout, k, stride, pad) per conv, latent.py:123-160"""
    if image_size == 64:
        return [(input_dim, 32, 5, 2, 2), (32, 64, 3, 2, 1), (64, 128, 3, 2, 1), (128, 256, 3, 2, 1),
                (256, output_dim, 4, 1, 0)]
    if image_size == 100:
        return [(input_dim, 32, 5, 2, 2), (32, 64, 3, 2, 1), (64, 128, 3, 2, 1), (128, 256, 3, 2, 1),
                (256, 256, 3, 2, 1), (256, output_dim, 4, 1, 0)]
    raise RuntimeError(f"SlacEncoderB200: image_size={image_size} (the reference defines 64 and 100)")


class SlacEncoderB200(nn.Module):
    def __init__(self, input_dim=3, output_dim=256, image_size=64):
        super().__init__()
        if input_dim != 3 or output_dim % 64:
            raise RuntimeError("SlacEncoderB200: input_dim must be 3 and output_dim a multiple of 64")
        _lib.load()
        self.image_size, self.output_dim = image_size, output_dim
        self.specs = _specs(image_size, input_dim, output_dim)
        mods = []
        for cin, cout, k, st, pad in self.specs:     # parameter container with the reference's key names
            conv = nn.Conv2d(cin, cout, k, st, pad)
            nn.init.xavier_uniform_(conv.weight, gain=1.0)   # initialize_weight, network/initializer.py:4-8
            nn.init.constant_(conv.bias, 0)
            mods += [conv, nn.LeakyReLU(0.2, inplace=True)]
        self.net = nn.Sequential(*mods)
        self._packed = None
        self._key = None

    def _convs(self):
        return [m for m in self.net if isinstance(m, nn.Conv2d)]

    def _pack(self):
        convs = self._convs()
        key = tuple((c.weight._version, c.bias._version, c.weight.data_ptr()) for c in convs)
        if self._packed is not None and key == self._key:
            return self._packed
        packed = []
        for i, c in enumerate(convs):
            w = c.weight.detach().float()
            b = c.bias.detach().float().contiguous()
            if i == 0:
                packed.append((w.contiguous(), b))
                continue
            if w.shape[1] % 64:                         # layer 2 reads the 64-pitch output of layer 1
                w = torch.nn.functional.pad(w, (0, 0, 0, 0, 0, 64 - w.shape[1] % 64))
            packed.append((ops.pack_conv_weight(w), b))
        self._packed, self._key = packed, key
        return packed

    def _tail(self, y, n):
        """layers 2.. on the implicit-GEMM kernel; y: NHWC bf16 [n, Ho, Wo, 64]"""
        packed = self._pack()
        last = len(self.specs) - 1
        for i in range(1, last):
            _, cout, k, st, pad = self.specs[i]
            y = ops.conv_igemm(y, packed[i][0], packed[i][1], cout, k, stride=st, act=ACT_LRELU, pad=pad)
        _, cout, k, _, _ = self.specs[last]
        if y.shape[1] != k or y.shape[2] != k:
            raise RuntimeError(f"SlacEncoderB200: final map is {tuple(y.shape[1:3])}, expected {k}x{k}")
        flat = y.reshape(1, 1, n, k * k * y.shape[3])   # 4x4 valid conv == linear over (h, w, c)
        out = ops.conv_igemm(flat, packed[last][0], packed[last][1], cout, 1, stride=1, act=ACT_LRELU, pad=0)
        return out.reshape(n, cout).float()

    def _conv1(self, n, h, w, frames=None, frames_tp1=None, codes=None, x=None):
        dev = self.net[0].weight.device
        w1, b1 = self._pack()[0]
        ho, wo = (h + 4 - 5) // 2 + 1, (w + 4 - 5) // 2 + 1
        y = torch.empty(n, ho, wo, 64, dtype=torch.bfloat16, device=dev)
        p = lambda t: 0 if t is None else t.data_ptr()
        _lib.check(_lib.load().s2p_slac_conv1_fwd(p(frames), p(frames_tp1), 0 if frames is None else frames.shape[0], p(codes), p(x), w1.data_ptr(), b1.data_ptr(),
                                                  y.data_ptr(), 64, n, h, w, _stream()), "s2p_slac_conv1_fwd")
        return y

    @torch.no_grad()
    def forward(self, x):
        """x: float [B, S, 3, H, W] in [0,1] (latent.py:162-170) -> [B, S, output_dim] fp32"""
        if not x.is_cuda:
            raise RuntimeError("SlacEncoderB200 needs CUDA tensors: there is no CPU path")
        B, S, C, H, W = x.shape
        if C != 3 or H != self.image_size or W != self.image_size:
            raise RuntimeError(f"SlacEncoderB200: expected [B,S,3,{self.image_size},{self.image_size}], got {tuple(x.shape)}")
        xf = x.float().contiguous().view(B * S, C, H, W)
        return self._tail(self._conv1(B * S, H, W, x=xf), B * S).view(B, S, -1)

    @torch.no_grad()
    def forward_from_buffer(self, buffer, idxes):
        """Encoder(buffer.gather_states(idxes)) without materialising the float windows: layer 1 reads the
        uint8 HWC frames of a SlacReplayBufferB200 through its window codes."""
        idx = torch.as_tensor(np.asarray(idxes), dtype=torch.int64, device=buffer.device)
        codes = buffer.codes[idx].contiguous()
        b, s = codes.shape
        c, h, w = buffer.state_shape
        y = self._conv1(b * s, h, w, frames=buffer.frames, frames_tp1=buffer.frames_tp1, codes=codes)
        return self._tail(y, b * s).view(b, s, -1)


def _decoder_specs(image_size, input_dim, output_dim):
    """(cin, cout, k, stride, pad, output_padding) per ConvTranspose2d, latent.py:63-101"""
    if image_size == 64:
        return [(input_dim, 256, 4, 1, 0, 0), (256, 128, 3, 2, 1, 1), (128, 64, 3, 2, 1, 1), (64, 32, 3, 2, 1, 1),
                (32, output_dim, 5, 2, 2, 1)]
    if image_size == 100:
        return [(input_dim, 256, 4, 1, 0, 0), (256, 256, 3, 2, 1, 0), (256, 128, 3, 2, 1, 0), (128, 64, 3, 2, 1, 0),
                (64, 32, 3, 2, 1, 1), (32, output_dim, 5, 2, 2, 1)]
    raise RuntimeError(f"SlacDecoderB200: image_size={image_size} (the reference defines 64 and 100)")


def _pad64(n):
    return (n + 63) // 64 * 64


class SlacDecoderB200(nn.Module):
    """Mirror of ``Decoder`` (/root/reference/rlkit/torch/slac/network/latent.py:55-115): same constructor and
    parameter names (ConvTranspose2d weights [Cin, Cout, k, k]), ``forward(x[B,S,latent]) -> (mean [B,S,3,H,W], std)``.

    Every transposed conv runs on the generator's tcgen05 implicit-GEMM kernel in its data-gradient form: the input is
    zero-inserted (stride 2; output_padding = extra zero rows/columns at the far edge), the filter is flipped with
    in/out channels swapped, padding k-1-p; the first layer (4x4 on a 1x1 map) is a 1x1 conv latent -> 4*4*256.
    Channel counts are zero-padded to multiples of 64 (288 -> 320, 32 -> 64, 3 -> 64).  Inference only."""

    def __init__(self, input_dim=288, output_dim=3, std=1.0, image_size=64):
        super().__init__()
        _lib.load()
        self.image_size, self.std, self.output_dim = image_size, std, output_dim
        self.specs = _decoder_specs(image_size, input_dim, output_dim)
        mods = []
        for cin, cout, k, st, pad, op in self.specs:
            m = nn.ConvTranspose2d(cin, cout, k, st, pad, op)
            nn.init.xavier_uniform_(m.weight, gain=1.0)
            nn.init.constant_(m.bias, 0)
            mods += [m, nn.LeakyReLU(0.2, inplace=True)]
        self.net = nn.Sequential(*mods)
        self._packed, self._key = None, None

    def _pack(self):
        convs = [m for m in self.net if isinstance(m, nn.ConvTranspose2d)]
        key = tuple((c.weight._version, c.bias._version, c.weight.data_ptr()) for c in convs)
        if self._packed is not None and key == self._key:
            return self._packed
        packed = []
        for i, c in enumerate(convs):
            w = c.weight.detach().float()                    # [Cin, Cout, k, k]
            b = c.bias.detach().float()
            cin, cout, k, _ = w.shape
            if i == 0:
                # y[n, r, s, co] = sum_ci x[n, ci] * w[ci, co, r, s]: a 1x1 conv with Cout' = k*k*Cout in (r, s, co) order
                w1 = w.permute(2, 3, 1, 0).reshape(k * k * cout, cin, 1, 1)
                w1 = torch.nn.functional.pad(w1, (0, 0, 0, 0, 0, _pad64(cin) - cin))
                packed.append((ops.pack_conv_weight(w1.contiguous()), b.repeat(k * k).contiguous()))
                continue
            weq = w.permute(1, 0, 2, 3).flip(2, 3)           # equivalent regular conv weight [Cout, Cin, k, k]
            weq = torch.nn.functional.pad(weq, (0, 0, 0, 0, 0, _pad64(cin) - cin, 0, _pad64(cout) - cout))
            packed.append((ops.pack_conv_weight(weq.contiguous()), torch.nn.functional.pad(b, (0, _pad64(cout) - cout)).contiguous()))
        self._packed, self._key = packed, key
        return packed

    @torch.no_grad()
    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("SlacDecoderB200 needs CUDA tensors: there is no CPU path")
        B, S, L = x.shape
        if L != self.specs[0][0]:
            raise RuntimeError(f"SlacDecoderB200: latent dim {L} != {self.specs[0][0]}")
        n = B * S
        packed = self._pack()
        dev = x.device
        cin0, cout0, k0 = self.specs[0][0], self.specs[0][1], self.specs[0][2]
        z = torch.zeros(1, 1, n, _pad64(cin0), dtype=torch.bfloat16, device=dev)
        z[0, 0, :, :cin0] = x.reshape(n, L)
        y = ops.conv_igemm(z, packed[0][0], packed[0][1], k0 * k0 * cout0, 1, stride=1, act=ACT_LRELU, pad=0)
        y = y.reshape(n, k0, k0, cout0)
        for i in range(1, len(self.specs)):
            cin, cout, k, st, pad, op = self.specs[i]
            h = y.shape[1]
            g = torch.zeros(n, (h - 1) * st + 1 + op, (h - 1) * st + 1 + op, y.shape[3], dtype=torch.bfloat16, device=dev)
            g[:, 0:(h - 1) * st + 1:st, 0:(h - 1) * st + 1:st] = y
            y = ops.conv_igemm(g, packed[i][0], packed[i][1], _pad64(cout), k, stride=1, act=ACT_LRELU, pad=k - 1 - pad)
        mean = y[..., :self.output_dim].permute(0, 3, 1, 2).float().reshape(B, S, self.output_dim, y.shape[1], y.shape[2])
        return mean, torch.ones_like(mean).mul_(self.std)
