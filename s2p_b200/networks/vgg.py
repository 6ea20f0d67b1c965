"""VGG19 feature extractor + perceptual loss of the training step (SURVEY.md section 8 a8;
/root/reference/rebuttal.md:71,131-135: "L1 + GAN + perceptual (VGG pretrained on ImageNet)").

Same parameter names as torchvision's ``vgg19().features`` (``features.0.weight`` ...) and as the oracle's
VGG19Features (ORACLE_SPEC v1.2), so either state_dict loads.  The ImageNet checkpoint cannot be fetched
offline: the extractor is random-init under the spec's seed (SURVEY.md section 8 d cfg 2) -- the arithmetic
of the step is complete, perceptual quality is not claimed.

Every conv (3x3 + ReLU in the epilogue) is the tcgen05 implicit-GEMM kernel and differentiates through its
dgrad form (the extractor is frozen: no wgrad); 2x2 max pooling is ``maxpool2_{fwd,bwd}_kernel``.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import autograd as A
from .. import ops
from ..ops import ACT_RELU

VGG19_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512]
VGG_SLICE_ENDS = (2, 7, 12, 21, 30)
VGG_WEIGHTS = (1.0 / 32, 1.0 / 16, 1.0 / 8, 1.0 / 4, 1.0)
VGG_SEED = 4321
LAMBDA_VGG = 10.0


class VGG19FeaturesB200(nn.Module):
    def __init__(self):
        super().__init__()
        layers, cin = [], 3
        for v in VGG19_CFG:
            if v == "M":
                layers.append(nn.MaxPool2d(2, 2))
            else:
                layers += [nn.Conv2d(cin, v, 3, padding=1), nn.ReLU(inplace=False)]   # parameter containers only
                cin = v
        self.features = nn.Sequential(*layers)
        gen = torch.Generator().manual_seed(VGG_SEED)
        for m in self.features:
            if isinstance(m, nn.Conv2d):
                std = math.sqrt(2.0 / (m.out_channels * 9))
                with torch.no_grad():
                    m.weight.copy_(torch.randn(m.weight.shape, generator=gen) * std)
                    m.bias.zero_()
        for p in self.parameters():
            p.requires_grad_(False)
        self._packs, self._packs_key = None, None

    def _packed(self):
        """(padded weight, (fwd pack, dgrad pack)) per conv -- the extractor is frozen, so packed once per state_dict"""
        convs = [m for m in self.features if isinstance(m, nn.Conv2d)]
        key = tuple((c.weight.data_ptr(), c.weight._version) for c in convs)
        if self._packs is None or key != self._packs_key:
            packs = []
            for c in convs:
                w = c.weight.detach()
                if w.shape[1] % 64:
                    w = F.pad(w, (0, 0, 0, 0, 0, 64 - w.shape[1] % 64))
                packs.append((w, (ops.pack_conv_weight(w), ops.pack_conv_weight_dgrad(w))))
            self._packs, self._packs_key = packs, key
        return self._packs

    def forward(self, image):
        """image: fp32 NCHW [B,3,H,W] in [-1,1] -> five NHWC bf16 feature maps (relu1_1 ... relu5_1)"""
        if not image.is_cuda:
            raise RuntimeError("VGG19FeaturesB200 needs CUDA tensors; there is no CPU fallback")
        x = A.pad_channels(image.float())   # NHWC bf16, 3 -> 64 zero-padded channels, one kernel (and one for its backward)
        outs = []
        packs = iter(self._packed())
        i = 0
        while i < len(self.features):
            m = self.features[i]
            if isinstance(m, nn.MaxPool2d):
                x = A.maxpool2(x)
                i += 1
            else:
                wgt, pk = next(packs)
                x = A.conv(x, wgt, m.bias, None, 1, ACT_RELU, None, 1, pk)
                i += 2                                                             # conv + its ReLU
            if i in VGG_SLICE_ENDS:
                outs.append(x)
        return outs


def vgg_loss(vgg, fake, real):
    """sum_i w_i * L1(vgg(fake)_i, vgg(real)_i.detach())  (SPADE VGGLoss; oracle vgg_loss)"""
    fx = vgg(fake)
    with torch.no_grad():
        fy = vgg(real)
    return sum(w * A.l1_mean(a, b) for w, a, b in zip(VGG_WEIGHTS, fx, fy))
