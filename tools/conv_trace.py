"""Timeline probe of conv_igemm_kernel: runs a few layer shapes against a private build of the library compiled with
-DS2P_CONV_TRACE (tools/conv_trace_build.sh -> tools/_trace/libs2p_trace.so) and prints, for CTA 0, how long the
first producer warp waited for a free ring slot, how long it took to issue a stage, how long the MMA thread waited for
data, how long it took to issue a stage's MMAs, and the tile-level waits.  Clock = clock64() of the CTA's SM.
(The producer rows cover every third stage: the stages of producer warp 0.)

usage (on the GPU box): bash tools/conv_trace_build.sh && python tools/conv_trace.py"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from s2p_b200 import _lib  # noqa: E402

_lib.LIB_PATH = os.path.join(ROOT, "tools", "_trace", "libs2p_trace.so")
from s2p_b200 import ops  # noqa: E402

lib = _lib.load()
lib.s2p_conv_trace_read.restype = C.c_int
lib.s2p_conv_trace_read.argtypes = [C.c_void_p]
N = 1024


def read():
    buf = np.zeros((12, N), dtype=np.int64)
    assert lib.s2p_conv_trace_read(buf.ctypes.data) == 0
    return buf


def report(name, stages_per_tile, skip_tiles=4, tiles=12, paired=True):
    t = read()
    s0, s1 = skip_tiles * stages_per_tile, (skip_tiles + tiles) * stages_per_tile
    p0, p1 = s0 // 3, s1 // 3             # producer warp 0 records its own stages only: every third
    prod_wait = (t[1] - t[0])[p0:p1]
    prod_issue = (t[2] - t[1])[p0:p1]
    # the MMA thread records one entry per loop iteration: two ring stages for tiles of at most 128 columns
    ipt = (stages_per_tile + 1) // 2 if paired else stages_per_tile
    m0, m1 = skip_tiles * ipt, (skip_tiles + tiles) * ipt
    mma_wait = (t[4] - t[3])[m0:m1]
    mma_issue = (t[5] - t[4])[m0:m1]
    per_stage = np.diff(t[5][m0:m1 + 1]) * ipt / stages_per_tile
    tile_wait = (t[7] - t[6])[skip_tiles:skip_tiles + tiles]
    epi_wait = (t[9] - t[8])[skip_tiles:skip_tiles + tiles]
    epi_busy = (t[8][skip_tiles + 1:skip_tiles + tiles + 1] - t[9][skip_tiles:skip_tiles + tiles])
    f = lambda a: f"mean {a.mean():7.0f}  med {np.median(a):7.0f}  max {a.max():7.0f}"
    print(f"== {name}: {stages_per_tile} stages per tile; clk per stage {per_stage.mean():.0f}, per tile {per_stage.mean() * stages_per_tile:.0f}")
    print(f"   producer wait(empty) {f(prod_wait)}\n   producer issue       {f(prod_issue)}")
    print(f"   mma wait(full)       {f(mma_wait)}   (per loop iteration: {2 if paired else 1} stage(s))\n   mma issue            {f(mma_issue)}")
    print(f"   mma wait(tmem_empty) {f(tile_wait)}\n   epi wait(tmem_full)  {f(epi_wait)}\n   epi busy             {f(epi_busy)}")
    if os.environ.get("CONV_TRACE_TIMELINE"):
        base = t[6][skip_tiles]
        print("   timeline of 5 tiles (clk from the first): issuer [wait acc .. got acc | per iteration: wait, got, issued] | epilogue [tile start, accumulator full]")
        for i in range(skip_tiles, skip_tiles + 5):
            its = " ".join(f"({t[3][j] - base},{t[4][j] - base},{t[5][j] - base})" for j in range(i * ipt, (i + 1) * ipt))
            print(f"     tile {i}: issuer {t[6][i] - base:6d} {t[7][i] - base:6d} | {its} | epilogue {t[8][i] - base:6d} {t[9][i] - base:6d}")


def run_plain(b, cin, cout, hw=100, cin2=0, stats=False, wide=False):
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(b, hw, hw, cin, device="cuda", generator=g).to(torch.bfloat16)
    w = torch.randn(cout, cin, 3, 3, device="cuda", generator=g) * 0.05
    wp = ops.pack_conv_weight(w)
    x2 = None
    if cin2:
        x2 = torch.randn(b, hw, hw, cin2, device="cuda", generator=g).to(torch.bfloat16)
        w2 = ops.pack_conv_weight(torch.randn(cout, cin2, 1, 1, device="cuda", generator=g) * 0.05)
        wp = torch.cat([wp.view(cout, -1), w2.view(cout, -1)], dim=1).contiguous()
    bias = torch.zeros(cout, device="cuda")
    st = torch.zeros(b, ops.conv_stats_rows(hw, hw, cout), cout, 2, device="cuda") if stats else None
    out = torch.empty(b, hw, hw, cout * (2 if wide else 1), dtype=torch.bfloat16, device="cuda")
    o, ol = (out[..., :cout], out[..., cout:]) if wide else (out, None)
    for _ in range(2):
        ops.conv_igemm(x, wp, bias, cout, 3, act=ops.ACT_NONE, stats=st, x2=x2, out=o, out_lo=ol)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.conv_igemm(x, wp, bias, cout, 3, act=ops.ACT_NONE, stats=st, x2=x2, out=o, out_lo=ol)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3


def run_torgb(b, cin=64, hw=100):
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(b, hw, hw, cin, device="cuda", generator=g).to(torch.bfloat16)
    wp = ops.pack_torgb_weight(torch.randn(3, cin, 3, 3, device="cuda", generator=g) * 0.05)
    bias = torch.zeros(3, device="cuda")
    u8 = torch.empty(b, hw, hw, 3, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        ops.torgb_igemm(x, wp, bias, out_u8=u8)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.torgb_igemm(x, wp, bias, out_u8=u8)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3


if __name__ == "__main__":
    B = 128
    us = run_torgb(B)
    report(f"ToRGB <48 folded, HH> Cin 64 ({us:.0f} us)", 3)
    us = run_plain(B, 64, 64)
    report(f"<64, HH> Cin 64 ({us:.0f} us)", 3)
    us = run_plain(B, 128, 64, stats=True, wide=True)
    report(f"<64, HH> Cin 128, stats + wide out ({us:.0f} us)", 6)
    us = run_plain(B, 64, 64, cin2=128, stats=True)
    report(f"<64, HH> Cin 64 + 1x1 of 128 ({us:.0f} us)", 5)
    us = run_plain(B, 256, 128, stats=True)
    report(f"<128> Cin 256 ({us:.0f} us)", 36, skip_tiles=2, tiles=6)
    us = run_plain(B, 64, 768)
    report(f"<256> Cin 64 -> 768 (hs1-like, {us:.0f} us)", 9, skip_tiles=4, tiles=12, paired=False)
