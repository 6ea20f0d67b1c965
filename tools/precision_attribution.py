"""CPU emulation of the B200 forward's rounding points, for error attribution.

The product path stores activations as bf16 between kernels and feeds bf16 operands to the tensor cores (fp32
accumulate).  This tool replays the ORACLE's arithmetic in fp32 on the CPU with a bf16 rounding inserted at each
site the product rounds at, and reports the PSNR against the un-rounded oracle with (a) every site on, (b) every
site on but one (what removing that rounding would buy), (c) only one site on (that site's own noise floor).
It is a planning tool: the numbers that count are the GPU parity tests.

    python tools/precision_attribution.py [--env cheetah] [--init scaled] [--sites all]
"""
import argparse
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import common  # noqa: E402

SITES = ["w_conv", "w_gb", "e", "f", "hs", "gb", "x", "t", "dx", "ts", "xlast"]


def bf(t):
    return t.to(torch.bfloat16).to(torch.float32)


class Emu:
    """sites: set of rounding sites that are ON.  hi_lo: set of sites stored as a bf16 hi+lo pair instead (error 2^-17)."""

    def __init__(self, net, sites, split=()):
        self.net, self.on, self.split = net, set(sites), set(split)

    def R(self, site, t):
        if site in self.split:
            hi = bf(t)
            return hi + bf(t - hi)
        return bf(t) if site in self.on else t

    def conv(self, x, conv, wsite="w_conv", scale=None):
        w = conv.weight if scale is None else conv.weight * scale
        b = conv.bias if (scale is None or conv.bias is None) else conv.bias * scale
        return F.conv2d(x, self.R(wsite, w), b, conv.stride, conv.padding)

    def inorm(self, x):
        return F.instance_norm(x, eps=1e-5)

    def mat(self, m, x, w, feat_h):
        a, c = torch.sigmoid(m.blend_gamma), torch.sigmoid(m.blend_beta)
        gs, bs = m.fc_gamma(w)[:, :, None, None], m.fc_beta(w)[:, :, None, None]
        gi = self.R("gb", self.conv(feat_h, m.mlp_gamma, "w_gb", 1 - a))
        bi = self.R("gb", self.conv(feat_h, m.mlp_beta, "w_gb", 1 - c))
        return self.inorm(x) * (1 + a * gs + gi) + (c * bs + bi)

    def block(self, blk, x, w, feat, up):
        if up:
            x = F.interpolate(x, scale_factor=2, mode="nearest")
        hsh = lambda m: self.R("hs", F.relu(self.conv(feat, m.mlp_shared)))  # noqa: E731
        if blk.learned_shortcut:
            ts = self.R("ts", self.mat(blk.norm_s, x, w, hsh(blk.norm_s)))
            xs = self.conv(ts, blk.conv_s)
        else:
            xs = x
        t = self.R("t", F.leaky_relu(self.mat(blk.norm_0, x, w, hsh(blk.norm_0)), 0.2))
        dx = self.R("dx", self.conv(t, blk.conv_0))
        t = self.R("t", F.leaky_relu(self.mat(blk.norm_1, dx, w, hsh(blk.norm_1)), 0.2))
        return xs + self.conv(t, blk.conv_1)

    @torch.no_grad()
    def forward(self, image, state):
        n = self.net
        w = n.state_encoder(state)
        enc = n.image_encoder
        e1 = self.R("e", F.conv2d(image, enc.conv1.weight, enc.conv1.bias, 1, 1))   # stem: fp32 weights
        f1 = self.R("f", F.leaky_relu(self.inorm(e1), 0.2))
        e2 = self.R("e", self.conv(f1, enc.conv2))
        f2 = self.R("f", F.leaky_relu(self.inorm(e2), 0.2))
        e3 = self.R("e", self.conv(f2, enc.conv3))
        f4 = self.R("f", F.leaky_relu(self.inorm(e3), 0.2))
        x = self.R("x", self.conv(f4, n.head))
        x = self.R("x", self.block(n.block_0, x, w, f4, False))
        x = self.R("x", self.block(n.block_1, x, w, f4, False))
        x = self.R("x", self.block(n.block_2, x, w, f2, True))
        x = self.R("x", self.block(n.block_3, x, w, f1, True))
        x = self.R("xlast", F.leaky_relu(self.block(n.block_4, x, w, f1, False), 0.2))
        return torch.tanh(self.conv(x, n.conv_img))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--env", default="cheetah")
    ap.add_argument("--init", default="scaled")
    ap.add_argument("--batch", type=int, default=2)
    ap.add_argument("--quick", action="store_true", help="only the all-on line and the leave-one-out table")
    a = ap.parse_args()
    torch.set_num_threads(os.cpu_count())
    net = common.build_oracle(a.env, a.init)
    image, states = common.make_inputs(a.env, a.batch, 1)
    st = states[:, 0]
    with torch.no_grad():
        ref = net(image, st)
    base = Emu(net, []).forward(image, st)
    print(f"emulator with no rounding vs oracle: {common.psnr01(base, ref):.1f} dB (restatement check)")
    allp = common.psnr01(Emu(net, SITES).forward(image, st), ref)
    print(f"all sites on (product today, fused MAT => 'gb' off): "
          f"{common.psnr01(Emu(net, [s for s in SITES if s != 'gb']).forward(image, st), ref):.2f} dB; with bf16 gb maps {allp:.2f} dB")
    today = [s for s in SITES if s != "gb"]
    print(f"{'site':8s} {'leave-one-out':>14s} {'only-this':>10s} {'as hi+lo':>9s}")
    for s in today:
        loo = common.psnr01(Emu(net, [t for t in today if t != s]).forward(image, st), ref)
        only = float("nan") if a.quick else common.psnr01(Emu(net, [s]).forward(image, st), ref)
        print(f"{s:8s} {loo:14.2f} {only:10.2f}")
    combos = [("x,dx fp32", ["x", "dx"]), ("x,dx,e fp32", ["x", "dx", "e"]), ("x,dx,e,xlast", ["x", "dx", "e", "xlast"]),
              ("x,dx,e,hs", ["x", "dx", "e", "hs"]), ("only weights rounded", [s for s in today if not s.startswith("w_")]),
              ("only w_conv+w_gb+t+f+ts+hs (operands only)", ["x", "dx", "e", "xlast"])]
    for name, off in combos:
        p = common.psnr01(Emu(net, [t for t in today if t not in off]).forward(image, st), ref)
        print(f"off: {name:45s} {p:.2f} dB")


if __name__ == "__main__":
    main()
