"""window_gather_kernel against the HBM roofline: batch x 9 frames of 100x100x3, 30 kB read + 120 kB written
per frame.  usage: python tools/slac_gather_bench.py [batch]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from s2p_b200.slac_buffer import SlacReplayBufferB200  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
N = 60_000
buf = SlacReplayBufferB200(N, 8, (3, 100, 100), (6,), "cuda")
buf.frames = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device="cuda")
buf.frames_tp1 = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device="cuda")
rows = torch.arange(8, N, device="cuda")
buf.codes = torch.cat([2 * (rows.unsqueeze(1) - 8 + torch.arange(8, device="cuda")), (2 * (rows - 1) + 1).unsqueeze(1)], 1)
buf._n = buf.codes.shape[0]
idx = np.random.RandomState(0).randint(0, buf._n, size=B)
for _ in range(3):
    buf.gather_states(idx)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 10
e0.record()
for _ in range(reps):
    out = buf.gather_states(idx)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
byts = B * 9 * 30000 * 5
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6552.0
print(f"window_gather_kernel: batch {B} x 9 frames in {ms * 1e3:.1f} us = {byts / ms / 1e6:.0f} GB/s algorithmic "
      f"({byts / ms / 1e6 / peak:.2f} of {peak} GB/s measured HBM peak), {B / ms * 1e3:.0f} windows/s")
