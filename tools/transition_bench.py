"""Times ensemble_transition_kernel over the whole-dataset call (1M rows, cheetah dims) with CUDA events.
fp32 CUDA-core roofline: 2 * rows * E * (32*256 + (L-1)*256*256 + 256*64) flops (padded dims) against
148 SMs x 128 FMA/clk x 2 x SM clock.   usage: python tools/transition_bench.py [rows]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from s2p_b200.transition import EnsembleTransitionB200  # noqa: E402
from s2p_b200 import _lib  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
torch.manual_seed(0)
m = EnsembleTransitionB200(17, 6, 256, 3).cuda()
norm = dict(obs_mean=np.zeros(17, np.float32), obs_std=np.ones(17, np.float32), next_obs_mean=np.zeros(17, np.float32),
            next_obs_std=np.ones(17, np.float32), reward_mean=0.0, reward_std=1.0)
obs = torch.randn(T, 17, device="cuda")
act = torch.rand(T, 6, device="cuda") * 2 - 1
idx = torch.randint(0, 7, (T,), device="cuda", dtype=torch.int32)
for _ in range(3):
    m.rollout_rows(obs, act, idx, norm)
torch.cuda.synchronize()
n0 = _lib.launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 5
e0.record()
for _ in range(reps):
    m.rollout_rows(obs, act, idx, norm)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
flops = 2.0 * T * 7 * (32 * 256 + 2 * 256 * 256 + 256 * 64)
print(f"ensemble_transition_kernel: {T} rows in {ms:.2f} ms = {T / ms * 1e3 / 1e6:.2f} M rows/s, "
      f"{flops / ms / 1e9:.1f} TFLOP/s fp32 (padded dims), launches per call {(_lib.launch_count() - n0) // reps}")
