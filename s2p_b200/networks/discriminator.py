"""B200-native multiscale patch discriminator for train.py --netG=s2p (README.md:59; GAN loss
per rebuttal.md:71,131-135).  Same parameter names/shapes as ORACLE_SPEC v1.1
(oracle/s2p_oracle.py MultiscaleDiscriminator), so state_dicts are interchangeable.  Every
conv (4x4, stride 2/1, pad 2) runs on the tcgen05 implicit-GEMM kernel and differentiates
through the dgrad / wgrad kernels; IN + LeakyReLU is the fused norm kernel and its backward.
The 6 input channels and the single logit channel are zero-padded to 64 channels."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import autograd as A
from .. import ops
from ..ops import ACT_LRELU, ACT_NONE


class _NLayerParams(nn.Module):
    def __init__(self, input_nc=6, ndf=64, n_layers=4):
        super().__init__()
        nf = ndf
        specs = [(input_nc, nf, 2, False)]
        for n in range(1, n_layers):
            nf_prev, nf = nf, min(nf * 2, 512)
            specs.append((nf_prev, nf, 1 if n == n_layers - 1 else 2, True))
        specs.append((nf, 1, 1, False))
        self.specs = specs
        for i, (ci, co, st, norm) in enumerate(specs):
            # nn.Sequential with the conv at index 0: parameter names "model{i}.0.weight" as in the oracle
            self.add_module(f"model{i}", nn.Sequential(nn.Conv2d(ci, co, 4, st, 2)))

    def padded_params(self):
        """[(weight, bias)] with in / out channels zero-padded to multiples of 64 (what the kernels run on)"""
        out = []
        for i, (ci, co, st, norm) in enumerate(self.specs):
            conv = getattr(self, f"model{i}")[0]
            w, b = conv.weight, conv.bias
            if ci % 64:
                w = F.pad(w, (0, 0, 0, 0, 0, 64 - ci % 64))
            if co % 64:
                w = F.pad(w, (0, 0, 0, 0, 0, 0, 0, 64 - co % 64))
                b = F.pad(b, (0, 64 - co % 64))
            out.append((w, b))
        return out

    def forward(self, x_nhwc, params=None, packs=None):
        feats = []
        x = x_nhwc
        last = len(self.specs) - 1
        params = self.padded_params() if params is None else params
        for i, (ci, co, st, norm) in enumerate(self.specs):
            w, b = params[i]
            pk = None if packs is None else packs[i]
            B, h, wd, _ = x.shape
            ho, wo = (h + 4 - 4) // st + 1, (wd + 4 - 4) // st + 1
            if norm:
                stt = A.conv_stats((B, ho, wo, w.shape[0]), x.device, 4, st)
                y = A.conv(x, w, b, None, st, ACT_NONE, stt, 2, packed=pk)
                x = A.mat_norm(y, ops.in_stats_finalize(stt, ho * wo))
            else:
                x = A.conv(x, w, b, None, st, ACT_NONE if i == last else ACT_LRELU, None, 2, packed=pk)
            feats.append(x[..., :co].permute(0, 3, 1, 2))      # NCHW view, like the oracle's outputs
        return feats


class MultiscaleDiscriminatorB200(nn.Module):
    """forward(cat(prev_image, image)[B,6,H,W]) -> list[num_D] of lists of feature maps (last = logits)."""

    def __init__(self, opt=None, input_nc=6, ndf=64, n_layers=4, num_D=2):
        super().__init__()
        self.num_D = num_D
        for i in range(num_D):
            self.add_module(f"discriminator_{i}", _NLayerParams(input_nc, ndf, n_layers))
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.xavier_normal_(m.weight, gain=0.02)
                nn.init.zeros_(m.bias)

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("MultiscaleDiscriminatorB200 needs CUDA tensors; there is no CPU fallback")
        out = []
        # both packed forms of every conv weight of every scale in one launch (the first layer's data gradient only
        # when the input carries one: the generator step)
        nets = [getattr(self, f"discriminator_{i}") for i in range(self.num_D)]
        params = [n.padded_params() for n in nets]
        flat = [(w, x.requires_grad or li > 0) for ps in params for li, (w, _) in enumerate(ps)]
        packed = ops.pack_conv_weights_multi([w for w, _ in flat], [d for _, d in flat])
        for i in range(self.num_D):
            x0 = A.pad_channels(x.float())
            nl = len(params[i])
            out.append(nets[i](x0, params[i], packed[i * nl:(i + 1) * nl]))
            if i + 1 < self.num_D:
                x = A.avgpool3s2(x.float())
        return out


def define_D(opt=None):
    gpu = getattr(opt, "gpu_ids", [0]) if opt is not None else [0]
    if not gpu:
        raise RuntimeError("--gpu_ids=-1 (CPU) is not served by s2p_b200: it has no CPU fallback")
    return MultiscaleDiscriminatorB200(opt).to(torch.device("cuda", gpu[0]))
