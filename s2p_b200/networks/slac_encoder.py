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
        _lib.check(_lib.load().s2p_slac_conv1_fwd(p(frames), p(frames_tp1), p(codes), p(x), w1.data_ptr(), b1.data_ptr(),
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
