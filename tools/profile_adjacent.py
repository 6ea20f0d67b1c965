"""One call of each adjacent component inside a cudaProfilerStart/Stop range, for ncu:
state-transition producer (131072 rows), SLAC window gather (batch 256) and SLAC encoder (batch 64)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from s2p_b200 import _lib  # noqa: E402
from s2p_b200.networks.slac_encoder import SlacEncoderB200  # noqa: E402
from s2p_b200.slac_buffer import SlacReplayBufferB200  # noqa: E402
from s2p_b200.transition import EnsembleTransitionB200  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
T = 131072
tm = EnsembleTransitionB200(17, 6, 256, 3).to(dev)
norm = dict(obs_mean=np.zeros(17, np.float32), obs_std=np.ones(17, np.float32), next_obs_mean=np.zeros(17, np.float32),
            next_obs_std=np.ones(17, np.float32), reward_mean=0.0, reward_std=1.0)
obs, act = torch.randn(T, 17, device=dev), torch.rand(T, 6, device=dev) * 2 - 1
idx = torch.randint(0, 7, (T,), device=dev, dtype=torch.int32)
N = 20_000
buf = SlacReplayBufferB200(N, 8, (3, 100, 100), (6,), dev)
buf.frames = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device=dev)
buf.frames_tp1 = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device=dev)
rows = torch.arange(8, N, device=dev)
buf.codes = torch.cat([2 * (rows.unsqueeze(1) - 8 + torch.arange(8, device=dev)), (2 * (rows - 1) + 1).unsqueeze(1)], 1)
buf._n = buf.codes.shape[0]
enc = SlacEncoderB200(3, 256, 100).to(dev)
pick = np.random.RandomState(0).randint(0, buf._n, size=256)


def once():
    tm.rollout_rows(obs, act, idx, norm)
    buf.gather_states(pick)
    enc.forward_from_buffer(buf, pick[:64])


for _ in range(2):
    once()
torch.cuda.synchronize()
l0 = _lib.launch_count()
torch.cuda.profiler.start()
once()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(f"profile_adjacent: launches={_lib.launch_count() - l0}")
