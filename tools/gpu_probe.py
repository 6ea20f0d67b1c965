"""Bring-up probe: runs each kernel of libs2p_sm100.so against a plain torch fp32
reference on the GPU and prints diagnostics (not a test; tests/ has the assertions).
usage: python tools/gpu_probe.py <group>   group in {small, conv, conv_s2, conv_big}
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from s2p_b200 import ops

dev = "cuda"
torch.manual_seed(0)


def nhwc(x_nchw):
    return x_nchw.permute(0, 2, 3, 1).contiguous()


def report(name, got, ref, tol):
    err = (got.float() - ref.float()).abs()
    ok = bool(err.max() <= tol) and bool(torch.isfinite(got.float()).all())
    print(f"[{'OK ' if ok else 'BAD'}] {name}: max_abs={err.max().item():.4e} mean_abs={err.mean().item():.4e} "
          f"ref_absmax={ref.abs().max().item():.3f} tol={tol}", flush=True)
    return ok, err


def probe_conv(b, h, w, cin, cout, k, stride, act=0, bias=True, residual=False, stats=False, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(b, cin, h, w, generator=g).to(dev)
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(dev)
    bs = torch.randn(cout, generator=g).to(dev) if bias else None
    xb = nhwc(x).to(torch.bfloat16)
    wp = ops.pack_conv_weight(wt)
    pad = k // 2
    ho = (h + 2 * pad - k) // stride + 1
    wo = (w + 2 * pad - k) // stride + 1
    res = torch.randn(b, ho, wo, cout, generator=g).to(dev).to(torch.bfloat16) if residual else None
    st = None
    if stats:
        st = torch.full((b, ops.conv_stats_rows(ho, wo, cout, k, stride), cout, 2), float("nan"), device=dev)
    y = ops.conv_igemm(xb, wp, bs, cout, k, stride, act, res, None, st)
    torch.cuda.synchronize()
    ref = F.conv2d(xb.float().permute(0, 3, 1, 2), wp.float().view(cout, k, k, cin).permute(0, 3, 1, 2), bs, stride, pad)
    ref = nhwc(ref)
    if residual:
        ref = ref + res.float()
    if act == 1:
        ref = F.relu(ref)
    elif act == 2:
        ref = F.leaky_relu(ref, 0.2)
    name = f"conv b{b} {h}x{w} {cin}->{cout} k{k} s{stride} act{act} bias{int(bias)} res{int(residual)} stats{int(stats)}"
    ok, err = report(name, y, ref, 0.03 + 0.01 * ref.abs().max().item())
    if not ok:
        e = err.view(b, ho, wo, cout)
        print("   err by image :", e.amax(dim=(1, 2, 3)).tolist()[:8])
        print("   err by row h :", [round(v, 3) for v in e.amax(dim=(0, 2, 3)).tolist()[:30]])
        print("   err by col w :", [round(v, 3) for v in e.amax(dim=(0, 1, 3)).tolist()[:30]])
        ec = e.amax(dim=(0, 1, 2))
        print("   err by channel (first 16):", [round(v, 3) for v in ec.tolist()[:16]], " bad channels:", int((ec > 0.05).sum()))
        print("   got[0,0,0,:8]", y[0, 0, 0, :8].float().tolist())
        print("   ref[0,0,0,:8]", ref[0, 0, 0, :8].tolist())
        print("   got[0,1,1,:8]", y[0, min(1, ho - 1), min(1, wo - 1), :8].float().tolist())
        print("   ref[0,1,1,:8]", ref[0, min(1, ho - 1), min(1, wo - 1), :8].tolist())
    if stats:
        yf = y.float()
        s1 = st[..., 0].sum(dim=1)
        s2 = st[..., 1].sum(dim=1)
        report("   stats sum", s1, yf.sum(dim=(1, 2)), 1e-2 * max(1.0, yf.sum(dim=(1, 2)).abs().max().item()))
        report("   stats sumsq", s2, (yf * yf).sum(dim=(1, 2)), 1e-3 * (yf * yf).sum(dim=(1, 2)).abs().max().item())
        mr = ops.in_stats_finalize(st, ho * wo)
        report("   mean", mr[..., 0], yf.mean(dim=(1, 2)), 1e-4)
        report("   rstd", mr[..., 1], 1.0 / torch.sqrt(yf.var(dim=(1, 2), unbiased=False) + 1e-5), 1e-3)
    return ok


def group_small():
    # state encoder
    s = torch.randn(5, 17, device=dev) * 10
    pe = ops.pos_encode(s, 10)
    ref = torch.cat([s] + [f(s * 2.0 ** k) for k in range(10) for f in (torch.sin, torch.cos)], dim=-1)
    report("pos_encode", pe, ref, 2e-6)
    x = torch.randn(37, 357, device=dev)
    w = torch.randn(512, 357, device=dev) / 19
    b = torch.randn(512, device=dev)
    report("linear lrelu", ops.linear(x, w, b, 2), F.leaky_relu(F.linear(x, w, b), 0.2), 1e-4)
    # stem
    img = torch.rand(3, 3, 100, 100, device=dev) * 2 - 1
    wt = torch.randn(64, 3, 3, 3, device=dev) / 5
    bs = torch.randn(64, device=dev)
    st = torch.zeros(3, ops.stem_stats_rows(100, 100), 64, 2, device=dev)
    y = ops.stem_conv(img, wt, bs, 64, None, st)
    ref = nhwc(F.conv2d(img, wt, bs, 1, 1))
    report("stem f32", y, ref, 0.03)
    mr = ops.in_stats_finalize(st, 10000)
    report("stem mean", mr[..., 0], y.float().mean(dim=(1, 2)), 1e-4)
    report("stem rstd", mr[..., 1], 1 / torch.sqrt(y.float().var(dim=(1, 2), unbiased=False) + 1e-5), 1e-3)
    u8 = torch.randint(0, 256, (2, 100, 100, 3), dtype=torch.uint8, device=dev)
    y8 = ops.stem_conv(u8, wt, bs, 64)
    ref8 = nhwc(F.conv2d(u8.permute(0, 3, 1, 2).float() / 127.5 - 1, wt, bs, 1, 1))
    report("stem u8", y8, ref8, 0.03)
    # norm
    for c, up in ((64, 1), (256, 2), (512, 1)):
        hx = 25
        xx = (torch.randn(2, hx, hx, c, device=dev) * 2 + 0.5).to(torch.bfloat16)
        h = hx * up
        mean = xx.float().mean(dim=(1, 2))
        rstd = 1 / torch.sqrt(xx.float().var(dim=(1, 2), unbiased=False) + 1e-5)
        mr = torch.stack([mean, rstd], dim=-1).contiguous()
        gsbs = torch.randn(2, 2 * c + 64, device=dev) * 0.3
        gs, bs_ = gsbs[:, :c], gsbs[:, c:2 * c]
        gb = (torch.randn(2, h, h, 2 * c, device=dev) * 0.3).to(torch.bfloat16)
        y = ops.mat_norm(xx, mr, gs, bs_, gb, up, 2)
        xu = xx.float()
        if up == 2:
            xu = xu.repeat_interleave(2, dim=1).repeat_interleave(2, dim=2)
        xh = (xu - mean[:, None, None]) * rstd[:, None, None]
        ref = F.leaky_relu(xh * (1 + gs[:, None, None] + gb[..., :c].float()) + bs_[:, None, None] + gb[..., c:].float(), 0.2)
        report(f"mat_norm C{c} up{up}", y, ref, 0.05)
        y2 = ops.mat_norm(xx, mr, None, None, None, up, 2)
        report(f"plain IN C{c} up{up}", y2, F.leaky_relu(xh, 0.2), 0.05)
    # tail
    xt = torch.randn(2, 100, 100, 64, device=dev).to(torch.bfloat16)
    wt3 = torch.randn(3, 64, 3, 3, device=dev) / 24
    b3 = torch.randn(3, device=dev) * 0.1
    of = torch.empty(2, 3, 100, 100, device=dev)
    o8 = torch.empty(2, 100, 100, 3, dtype=torch.uint8, device=dev)
    ops.torgb_tanh(xt, wt3, b3, of, o8)
    ref = torch.tanh(F.conv2d(F.leaky_relu(xt.float().permute(0, 3, 1, 2), 0.2), wt3, b3, 1, 1))
    report("torgb f32", of, ref, 1e-4)
    r8 = torch.floor((ref + 1) * 127.5 + 0.5).clamp(0, 255).permute(0, 2, 3, 1)
    report("torgb u8", o8, r8, 1.0)


def group_conv():
    probe_conv(1, 25, 25, 64, 64, 1, 1, bias=False)
    probe_conv(1, 25, 25, 64, 64, 3, 1, bias=False)
    probe_conv(2, 25, 25, 128, 128, 3, 1, act=2, bias=True)
    probe_conv(2, 50, 50, 64, 256, 3, 1, act=1, bias=True, stats=True)
    probe_conv(2, 25, 25, 256, 192, 3, 1, residual=True, stats=True)
    probe_conv(3, 100, 100, 128, 64, 3, 1, residual=True, stats=True)
    probe_conv(2, 20, 36, 64, 128, 3, 1)          # ragged tiles (partial boxes)
    probe_conv(4, 25, 25, 512, 1024, 3, 1)        # multiple n tiles, long K


def group_conv_s2():
    probe_conv(1, 100, 100, 64, 128, 3, 2, bias=True, stats=True)
    probe_conv(2, 50, 50, 128, 256, 3, 2, bias=True, stats=True)


def group_conv_big():
    probe_conv(16, 100, 100, 256, 512, 3, 1, bias=True)
    probe_conv(16, 100, 100, 512, 256, 1, 1, bias=False)


if __name__ == "__main__":
    grp = sys.argv[1]
    print(f"=== probe group {grp} on {torch.cuda.get_device_name(0)}", flush=True)
    globals()["group_" + grp]()
    torch.cuda.synchronize()
    print(f"=== group {grp} done", flush=True)
