"""Bulk forward throughput against the micro-batch size (rows per launch set): python tools/microbatch_sweep.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import s2p_b200

rows, reps = 1024, 8
g = torch.Generator().manual_seed(0)
frames = torch.randint(0, 256, (rows, 100, 100, 3), dtype=torch.uint8, generator=g).cuda()
for mb in (128, 256, 512, 1024):
    torch.manual_seed(1234)
    net = s2p_b200.define_G(s2p_b200.make_options(env_type="cheetah", gpu_ids="0", init_type="scaled", max_microbatch=mb)).eval()
    net.pack()
    states = torch.randn(rows, net.opt.state_dim, generator=torch.Generator().manual_seed(1)).cuda()
    for _ in range(2):
        out = net.forward_u8(frames, states)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        out = net.forward_u8(frames, states)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ms = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))
    print(f"micro-batch {mb:4d}: {ms[len(ms) // 2]:8.3f} ms per {rows} rows  {rows / ms[len(ms) // 2] * 1e3:8.0f} frames/s  "
          f"peak mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB  checksum {int(out.sum().item())}")
    del net
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
