"""ORACLE_SPEC v1 -- fp32 PyTorch restatement of the S2P generator hot path.

TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this
file.  ``s2p_b200`` never imports it (tests/test_layout.py greps for that).

PARITY UNPINNED.  The reference's generator source (``simple_test.py``,
``train.py``, ``models/``, ``options/``) is absent from /root/reference
(SURVEY.md section 0): the only in-tree evidence for this path is prose
(/root/reference/README.md:33,59 for the CLI; /root/reference/rebuttal.md:6,71,
146-154,170-175 for the architecture) plus the boundary shapes on the consumer
side (/root/reference/rlkit/torch/slac/algo.py:174-175,189-190: frames are
[N,100,100,3] uint8 HWC; /root/reference/state_transition_rollout.py:44 and the
normalize_configs_dict.pkl shapes: cheetah state is 17 x f32).  There are no
golden vectors, tests or fixtures for the path in the reference (SURVEY.md
section 4).  Everything architectural below is therefore BUILDER-DEFINED and
frozen here as ORACLE_SPEC v1; the lineage designs the
README names (README.md:72-75: NVlabs/SPADE resnet block + param-free norm
modulation, nerf-pytorch sin/cos embedding, StyleGAN mapping MLP) guide the op
semantics.  Do not edit silently: any change is a spec bump and needs new
golden vectors (tests/golden/make_golden.py).

Spec (all fp32):

  opt: netG='s2p', env_type in {cheetah: S=17, walker: S=24}, ngf=64,
       nhidden=128, w_dim=512, pe_freqs=10 (include input), mapping_layers=8,
       lrelu slope 0.2, IN eps 1e-5.

  forward(image[B,3,100,100] in [-1,1], state[B,S]) -> [B,3,100,100] in [-1,1]
    StateEncoder : pe = cat(s, sin(2^0 s), cos(2^0 s), ..., sin(2^9 s), cos(2^9 s))
                   w  = (Linear + LeakyReLU(0.2)) x 8, width 512
    ImageEncoder : f100 = LReLU(IN(conv3x3(3->64)))          @100
                   f50  = LReLU(IN(conv3x3 s2 (64->128)))     @50
                   f25  = LReLU(IN(conv3x3 s2 (128->256)))    @25
    Trunk        : x = conv3x3(256->512)(f25)
                   x = MATResBlk(512->512) x2                 @25  (cond f25)
                   x = up2(x); x = MATResBlk(512->256)        @50  (cond f50)
                   x = up2(x); x = MATResBlk(256->128)        @100 (cond f100)
                   x = MATResBlk(128->64)                     @100 (cond f100)
                   y = tanh(conv3x3(64->3)(LReLU(x)))
    MATResBlk(fin->fout), fmid = min(fin, fout):
                   dx  = conv3x3(fin->fmid)(LReLU(MAT0(x)))
                   dx  = conv3x3(fmid->fout)(LReLU(MAT1(dx)))
                   xs  = conv1x1(fin->fout, no bias)(MATs(x)) if fin != fout else x
                   out = xs + dx
    MAT_C(x | w, f): xhat = InstanceNorm(x, eps=1e-5, affine=False)
                   gs, bs = Linear(512->C)(w) each            (state branch; "SAT")
                   h = ReLU(conv3x3(Cf->128)(f)); gI, bI = conv3x3(128->C)(h) each
                   g = a*gs + (1-a)*gI ; b = c*bs + (1-c)*bI ; a, c = sigmoid(scalars)
                   out = xhat * (1 + g) + b
"""
from __future__ import annotations

import argparse
import math
from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F

SPEC_VERSION = "ORACLE_SPEC v1"   # v1.3 added the 'trained_like' init only (no change to the forward)

# state dimensionality per --env_type.  cheetah=17 is pinned by the reference
# (normalize_configs_dict.pkl, state_transition_rollout.py:44); walker=24 is the
# DeepMind-Control walker observation size (not in tree).
STATE_DIM = {"cheetah": 17, "walker": 24}
IMAGE_SIZE = 100  # examples/iql/mujoco_finetune.py:242,327 (image_size=100)


def make_options(argv=None, **overrides):
    """SPADE-style options namespace; flag names from README.md:33,59."""
    p = argparse.ArgumentParser("s2p-oracle", add_help=False)
    p.add_argument("--netG", type=str, default="s2p")
    p.add_argument("--env_type", type=str, default="cheetah", choices=sorted(STATE_DIM))
    p.add_argument("--gpu_ids", type=str, default="-1")
    p.add_argument("--batchSize", type=int, default=16)
    p.add_argument("--dataroot", type=str, default="./datasets")
    p.add_argument("--start_idx", type=int, default=0)
    p.add_argument("--seq_len", type=int, default=5)
    p.add_argument("--ngf", type=int, default=64)
    p.add_argument("--nhidden", type=int, default=128)
    p.add_argument("--w_dim", type=int, default=512)
    p.add_argument("--pe_freqs", type=int, default=10)
    p.add_argument("--mapping_layers", type=int, default=8)
    p.add_argument("--init_type", type=str, default="xavier")
    p.add_argument("--init_variance", type=float, default=0.02)
    opt, _ = p.parse_known_args(argv if argv is not None else [])
    for k, v in overrides.items():
        setattr(opt, k, v)
    if isinstance(opt.gpu_ids, str):
        opt.gpu_ids = [int(t) for t in opt.gpu_ids.split(",") if t.strip() and int(t) >= 0]
    opt.state_dim = STATE_DIM[opt.env_type]
    return opt


def positional_encoding(s: torch.Tensor, n_freqs: int) -> torch.Tensor:
    """nerf-pytorch style embedding, include_input=True, log-sampled freqs 2^k."""
    outs = [s]
    for k in range(n_freqs):
        f = float(2 ** k)
        outs.append(torch.sin(s * f))
        outs.append(torch.cos(s * f))
    return torch.cat(outs, dim=-1)


class StateEncoder(nn.Module):
    def __init__(self, state_dim, n_freqs=10, w_dim=512, n_layers=8):
        super().__init__()
        self.n_freqs = n_freqs
        in_dim = state_dim * (1 + 2 * n_freqs)
        dims = [in_dim] + [w_dim] * n_layers
        self.layers = nn.ModuleList(nn.Linear(dims[i], dims[i + 1]) for i in range(n_layers))

    def forward(self, state):
        h = positional_encoding(state, self.n_freqs)
        for lin in self.layers:
            h = F.leaky_relu(lin(h), 0.2)
        return h


class ImageEncoder(nn.Module):
    def __init__(self, ngf=64):
        super().__init__()
        self.conv1 = nn.Conv2d(3, ngf, 3, 1, 1)
        self.conv2 = nn.Conv2d(ngf, ngf * 2, 3, 2, 1)
        self.conv3 = nn.Conv2d(ngf * 2, ngf * 4, 3, 2, 1)

    def forward(self, img):
        f100 = F.leaky_relu(F.instance_norm(self.conv1(img), eps=1e-5), 0.2)
        f50 = F.leaky_relu(F.instance_norm(self.conv2(f100), eps=1e-5), 0.2)
        f25 = F.leaky_relu(F.instance_norm(self.conv3(f50), eps=1e-5), 0.2)
        return f100, f50, f25


class MAT(nn.Module):
    """Multi-modal affine transformation (rebuttal.md:146-154)."""

    def __init__(self, norm_nc, feat_nc, w_dim=512, nhidden=128):
        super().__init__()
        self.fc_gamma = nn.Linear(w_dim, norm_nc)
        self.fc_beta = nn.Linear(w_dim, norm_nc)
        self.mlp_shared = nn.Conv2d(feat_nc, nhidden, 3, 1, 1)
        self.mlp_gamma = nn.Conv2d(nhidden, norm_nc, 3, 1, 1)
        self.mlp_beta = nn.Conv2d(nhidden, norm_nc, 3, 1, 1)
        self.blend_gamma = nn.Parameter(torch.zeros(()))
        self.blend_beta = nn.Parameter(torch.zeros(()))

    def forward(self, x, w, feat):
        xhat = F.instance_norm(x, eps=1e-5)
        gs = self.fc_gamma(w)[:, :, None, None]
        bs = self.fc_beta(w)[:, :, None, None]
        h = F.relu(self.mlp_shared(feat))
        gi = self.mlp_gamma(h)
        bi = self.mlp_beta(h)
        a = torch.sigmoid(self.blend_gamma)
        c = torch.sigmoid(self.blend_beta)
        g = a * gs + (1 - a) * gi
        b = c * bs + (1 - c) * bi
        return xhat * (1 + g) + b


class MATResBlk(nn.Module):
    def __init__(self, fin, fout, feat_nc, w_dim=512, nhidden=128):
        super().__init__()
        fmid = min(fin, fout)
        self.learned_shortcut = fin != fout
        self.conv_0 = nn.Conv2d(fin, fmid, 3, 1, 1)
        self.conv_1 = nn.Conv2d(fmid, fout, 3, 1, 1)
        self.norm_0 = MAT(fin, feat_nc, w_dim, nhidden)
        self.norm_1 = MAT(fmid, feat_nc, w_dim, nhidden)
        if self.learned_shortcut:
            self.conv_s = nn.Conv2d(fin, fout, 1, bias=False)
            self.norm_s = MAT(fin, feat_nc, w_dim, nhidden)

    def forward(self, x, w, feat):
        if self.learned_shortcut:
            xs = self.conv_s(self.norm_s(x, w, feat))
        else:
            xs = x
        dx = self.conv_0(F.leaky_relu(self.norm_0(x, w, feat), 0.2))
        dx = self.conv_1(F.leaky_relu(self.norm_1(dx, w, feat), 0.2))
        return xs + dx


class S2PGenerator(nn.Module):
    """netG for --netG=s2p: forward(image, state) -> next image."""

    def __init__(self, opt):
        super().__init__()
        self.opt = opt
        ngf, wd, nh = opt.ngf, opt.w_dim, opt.nhidden
        self.state_encoder = StateEncoder(opt.state_dim, opt.pe_freqs, wd, opt.mapping_layers)
        self.image_encoder = ImageEncoder(ngf)
        self.head = nn.Conv2d(ngf * 4, ngf * 8, 3, 1, 1)
        self.block_0 = MATResBlk(ngf * 8, ngf * 8, ngf * 4, wd, nh)   # @25, cond f25
        self.block_1 = MATResBlk(ngf * 8, ngf * 8, ngf * 4, wd, nh)   # @25, cond f25
        self.block_2 = MATResBlk(ngf * 8, ngf * 4, ngf * 2, wd, nh)   # @50, cond f50
        self.block_3 = MATResBlk(ngf * 4, ngf * 2, ngf, wd, nh)       # @100, cond f100
        self.block_4 = MATResBlk(ngf * 2, ngf, ngf, wd, nh)           # @100, cond f100
        self.conv_img = nn.Conv2d(ngf, 3, 3, 1, 1)
        init_weights(self, opt.init_type, opt.init_variance)

    def forward(self, image, state):
        w = self.state_encoder(state)
        f100, f50, f25 = self.image_encoder(image)
        x = self.head(f25)
        x = self.block_0(x, w, f25)
        x = self.block_1(x, w, f25)
        x = F.interpolate(x, scale_factor=2, mode="nearest")
        x = self.block_2(x, w, f50)
        x = F.interpolate(x, scale_factor=2, mode="nearest")
        x = self.block_3(x, w, f100)
        x = self.block_4(x, w, f100)
        return torch.tanh(self.conv_img(F.leaky_relu(x, 0.2)))

    @torch.no_grad()
    def rollout(self, image, states):
        """N-step autoregressive synthesis (README.md:30-33 --seq_len):
        img_{k+1} = netG(img_k, states[:, k]); float output fed back unquantised."""
        outs = []
        img = image
        for k in range(states.shape[1]):
            img = self.forward(img, states[:, k])
            outs.append(img)
        return torch.stack(outs, dim=1)


def init_weights(net, init_type="xavier", gain=0.02):
    """SPADE BaseNetwork.init_weights flavour: xavier_normal_(gain) on conv/linear
    weights, zero bias.  'scaled' is a variance-preserving (He-style, fan_in) init
    used by the parity tests so that activations and outputs are O(1) and the bf16
    path is actually exercised (gain-0.02 nets emit ~0 everywhere).  'trained_like'
    (v1.3) = 'scaled' with a non-saturating output head and blend scalars away from 0."""
    for m in net.modules():
        if isinstance(m, (nn.Conv2d, nn.Linear)):
            if init_type == "xavier":
                nn.init.xavier_normal_(m.weight, gain=gain)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif init_type in ("scaled", "trained_like"):
                fan_in = m.weight[0].numel()
                nn.init.normal_(m.weight, 0.0, math.sqrt(1.0 / fan_in))
                if m.bias is not None:
                    nn.init.normal_(m.bias, 0.0, 0.1)
            else:
                raise ValueError(f"unknown init_type {init_type}")
    if init_type == "scaled":
        for m in net.modules():
            if isinstance(m, MAT):
                nn.init.normal_(m.blend_gamma, 0.0, 1.0)
                nn.init.normal_(m.blend_beta, 0.0, 1.0)
    if init_type == "trained_like":
        # ORACLE_SPEC v1.3: the third parity init.  'scaled' weights, but (i) the ToRGB conv is
        # scaled by 0.3 so that tanh works in its near-linear range (output std ~0.3, nothing saturated: every
        # upstream rounding error reaches the output undamped), and (ii) every MAT blend scalar sits well away from
        # sigmoid = 0.5: |blend| ~ U(1, 2.5) with a random sign, i.e. a state-dominated or an image-dominated mix.
        for m in net.modules():
            if isinstance(m, MAT):
                for p in (m.blend_gamma, m.blend_beta):
                    mag = 1.0 + 1.5 * torch.rand(())
                    sign = 1.0 if torch.rand(()) < 0.5 else -1.0
                    with torch.no_grad():
                        p.fill_(float(mag * sign))
        conv_img = getattr(net, "conv_img", None)
        if conv_img is not None:
            with torch.no_grad():
                conv_img.weight.mul_(0.3)


def define_G(opt):
    if opt.netG != "s2p":
        raise ValueError(f"oracle implements --netG=s2p only, got {opt.netG}")
    return S2PGenerator(opt)


def to_uint8_hwc(img_nchw: torch.Tensor) -> torch.Tensor:
    """[-1,1] float NCHW -> uint8 NHWC, the on-disk layout of
    image_observations_tp1 (rlkit/torch/slac/algo.py:175,190)."""
    q = torch.floor((img_nchw + 1.0) * 127.5 + 0.5).clamp_(0, 255).to(torch.uint8)
    return q.permute(0, 2, 3, 1).contiguous()


def from_uint8_hwc(frames_nhwc: torch.Tensor) -> torch.Tensor:
    """uint8 NHWC -> [-1,1] float NCHW (inverse of the above up to rounding)."""
    return frames_nhwc.permute(0, 3, 1, 2).to(torch.float32) * (1.0 / 127.5) - 1.0


def flops_per_frame(opt=None) -> dict:
    """Algorithmic MACs per frame by part."""
    opt = opt or make_options()
    g = define_G(opt)
    macs = {}
    hw = {100: 100 * 100, 50: 50 * 50, 25: 25 * 25}

    def conv_macs(conv, res):
        k = conv.kernel_size[0] * conv.kernel_size[1]
        return hw[res] * k * conv.in_channels * conv.out_channels

    enc = g.image_encoder
    macs["encoder"] = conv_macs(enc.conv1, 100) + conv_macs(enc.conv2, 50) + conv_macs(enc.conv3, 25)
    macs["head"] = conv_macs(g.head, 25)
    trunk = 0
    mod = 0
    for blk, res in ((g.block_0, 25), (g.block_1, 25), (g.block_2, 50), (g.block_3, 100), (g.block_4, 100)):
        trunk += conv_macs(blk.conv_0, res) + conv_macs(blk.conv_1, res)
        mats = [blk.norm_0, blk.norm_1]
        if blk.learned_shortcut:
            trunk += conv_macs(blk.conv_s, res)
            mats.append(blk.norm_s)
        for m in mats:
            mod += conv_macs(m.mlp_shared, res) + conv_macs(m.mlp_gamma, res) + conv_macs(m.mlp_beta, res)
    macs["trunk"] = trunk
    macs["modulation"] = mod
    macs["torgb"] = conv_macs(g.conv_img, 100)
    se = sum(l.in_features * l.out_features for l in g.state_encoder.layers)
    se += sum(m.fc_gamma.in_features * m.fc_gamma.out_features * 2 for m in g.modules() if isinstance(m, MAT))
    macs["state"] = se
    macs["total"] = sum(macs.values())
    return macs


if __name__ == "__main__":
    m = flops_per_frame()
    for k, v in m.items():
        print(f"{k:12s} {v / 1e9:8.3f} GMAC")


# ======================================================================================
# ORACLE_SPEC v1.1 -- training side, BUILDER-DEFINED like the
# generator above (the reference's train.py, discriminator and losses are absent from the
# snapshot; semantics from rebuttal.md:71,131-135,188 "L1 + GAN + perceptual" and the SPADE
# lineage README.md:72-75).  v1.1 only ADDS classes; the generator spec is unchanged.
#
#   NLayerDiscriminator(input_nc=6): conv4x4 s2 p2 (6->64) + LReLU(0.2)
#       conv4x4 s2 p2 (64->128) + IN + LReLU ; conv4x4 s2 p2 (128->256) + IN + LReLU
#       conv4x4 s1 p2 (256->512) + IN + LReLU ; conv4x4 s1 p2 (512->1)
#       forward returns the 5 intermediate outputs (the last one = patch logits)
#   MultiscaleDiscriminator(num_D=2): D_0 on the input, D_1 on AvgPool2d(3, 2, 1,
#       count_include_pad=False)(input); input = cat(prev_image, image) on channels
#   losses: hinge GAN (D: mean relu(1 - D(real)) + mean relu(1 + D(fake)); G: -mean D(fake)),
#       averaged over scales; feature matching L1 over the intermediate outputs (lambda 10,
#       / num_D); L1(fake, real) * lambda_l1 (10).  The VGG19 perceptual term of the paper
#       needs ImageNet weights (not available offline): v1.1 had lambda_vgg = 0; v1.2 (below) adds the term with
#       a seeded random-init extractor, enabled by passing `vgg=` to training_step.
#   optimiser: Adam(lr 2e-4, betas (0.0, 0.9)) for G and for D (SPADE TTUR off).
# ======================================================================================
class NLayerDiscriminator(nn.Module):
    def __init__(self, input_nc=6, ndf=64, n_layers=4):
        super().__init__()
        kw, padw = 4, 2
        nf = ndf
        seq = [[nn.Conv2d(input_nc, nf, kw, 2, padw), nn.LeakyReLU(0.2)]]
        for n in range(1, n_layers):
            nf_prev, nf = nf, min(nf * 2, 512)
            stride = 1 if n == n_layers - 1 else 2
            seq.append([nn.Conv2d(nf_prev, nf, kw, stride, padw), nn.InstanceNorm2d(nf, eps=1e-5, affine=False),
                        nn.LeakyReLU(0.2)])
        seq.append([nn.Conv2d(nf, 1, kw, 1, padw)])
        for i, layers in enumerate(seq):
            self.add_module(f"model{i}", nn.Sequential(*layers))
        self.n_models = len(seq)

    def forward(self, x):
        feats = []
        for i in range(self.n_models):
            x = getattr(self, f"model{i}")(x)
            feats.append(x)
        return feats


class MultiscaleDiscriminator(nn.Module):
    def __init__(self, opt=None, input_nc=6, ndf=64, n_layers=4, num_D=2):
        super().__init__()
        self.num_D = num_D
        for i in range(num_D):
            self.add_module(f"discriminator_{i}", NLayerDiscriminator(input_nc, ndf, n_layers))
        init_weights(self, getattr(opt, "init_type", "xavier") if opt is not None else "xavier",
                     getattr(opt, "init_variance", 0.02) if opt is not None else 0.02)

    def forward(self, x):
        out = []
        for i in range(self.num_D):
            out.append(getattr(self, f"discriminator_{i}")(x))
            x = F.avg_pool2d(x, 3, 2, 1, count_include_pad=False)
        return out


def define_D(opt=None):
    return MultiscaleDiscriminator(opt)


LAMBDA_FEAT = 10.0
LAMBDA_L1 = 10.0
LAMBDA_VGG = 10.0
ADAM = dict(lr=2e-4, betas=(0.0, 0.9))

# ---- ORACLE_SPEC v1.2: VGG19 perceptual loss (rebuttal.md:71,131-135 "L1 + GAN + perceptual (VGG pretrained on
# ImageNet)"; lineage NVlabs/SPADE VGGLoss).  The ImageNet weights cannot be fetched offline, so the extractor is
# RANDOM-INIT under a fixed seed (SURVEY.md section 8 d, cfg 2: "VGG weights random-init (quality n/a)"): the
# arithmetic of the training step is complete, the perceptual quality of the loss is not claimed.
VGG19_CFG = [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512]  # ... conv5_1
VGG_SLICE_ENDS = (2, 7, 12, 21, 30)            # torchvision vgg19.features indices: relu1_1 ... relu5_1
VGG_WEIGHTS = (1.0 / 32, 1.0 / 16, 1.0 / 8, 1.0 / 4, 1.0)
VGG_SEED = 4321


class VGG19Features(nn.Module):
    """features[0:30] of torchvision's vgg19 (same module indices, so a real checkpoint would load);
    forward returns the five relu*_1 maps.  Inputs are the [-1,1] images as they are (SPADE does the same)."""

    def __init__(self):
        super().__init__()
        layers, cin = [], 3
        for v in VGG19_CFG:
            if v == "M":
                layers.append(nn.MaxPool2d(2, 2))
            else:
                layers += [nn.Conv2d(cin, v, 3, padding=1), nn.ReLU(inplace=False)]
                cin = v
        self.features = nn.Sequential(*layers)
        assert len(self.features) == 30
        gen = torch.Generator().manual_seed(VGG_SEED)
        for m in self.features:
            if isinstance(m, nn.Conv2d):       # torchvision's init: kaiming_normal_(fan_out, relu), zero bias
                std = math.sqrt(2.0 / (m.out_channels * 9))
                with torch.no_grad():
                    m.weight.copy_(torch.randn(m.weight.shape, generator=gen) * std)
                    m.bias.zero_()
        for p in self.parameters():
            p.requires_grad_(False)

    def forward(self, x):
        outs = []
        for i, m in enumerate(self.features):
            x = m(x)
            if i + 1 in VGG_SLICE_ENDS:
                outs.append(x)
        return outs


def vgg_loss(vgg, fake, real):
    """SPADE VGGLoss: sum_i w_i * L1(vgg(fake)_i, vgg(real)_i.detach())"""
    fx = vgg(fake)
    with torch.no_grad():
        fy = vgg(real)
    return sum(w * F.l1_loss(a, b) for w, a, b in zip(VGG_WEIGHTS, fx, fy))


def hinge_d_loss(pred_real, pred_fake):
    loss = 0.0
    for r, f in zip(pred_real, pred_fake):
        loss = loss + F.relu(1.0 - r[-1]).mean() + F.relu(1.0 + f[-1]).mean()
    return loss / len(pred_real)


def generator_losses(pred_fake, pred_real, fake, real, vgg=None):
    """dict of the G-step loss terms (SPADE Pix2PixModel.compute_generator_loss shape)."""
    num_d = len(pred_fake)
    gan = sum(-f[-1].mean() for f in pred_fake) / num_d
    feat = 0.0
    for f, r in zip(pred_fake, pred_real):
        for a, b in zip(f[:-1], r[:-1]):
            feat = feat + F.l1_loss(a, b.detach()) * LAMBDA_FEAT / num_d
    losses = {"GAN": gan, "GAN_Feat": feat, "L1": F.l1_loss(fake, real) * LAMBDA_L1}
    if vgg is not None:
        losses["VGG"] = vgg_loss(vgg, fake, real) * LAMBDA_VGG
    return losses


def training_step(netG, netD, optG, optD, prev_image, state, real_image, vgg=None):
    """one G step then one D step (SPADE Pix2PixTrainer order); returns (g_losses, d_loss, fake)."""
    fake = netG(prev_image, state)
    pred = netD(torch.cat([torch.cat([prev_image, fake], 1), torch.cat([prev_image, real_image], 1)], 0))
    pred_fake = [[t[: t.shape[0] // 2] for t in p] for p in pred]
    pred_real = [[t[t.shape[0] // 2:] for t in p] for p in pred]
    g_losses = generator_losses(pred_fake, pred_real, fake, real_image, vgg)
    optG.zero_grad(set_to_none=True)
    sum(g_losses.values()).backward()
    optG.step()
    with torch.no_grad():
        fake_d = fake.detach()
    pred = netD(torch.cat([torch.cat([prev_image, fake_d], 1), torch.cat([prev_image, real_image], 1)], 0))
    pred_fake = [[t[: t.shape[0] // 2] for t in p] for p in pred]
    pred_real = [[t[t.shape[0] // 2:] for t in p] for p in pred]
    d_loss = hinge_d_loss(pred_real, pred_fake)
    optD.zero_grad(set_to_none=True)
    d_loss.backward()
    optD.step()
    return g_losses, d_loss, fake_d
