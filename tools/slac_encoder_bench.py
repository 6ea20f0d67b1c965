"""SLAC encoder from the resident uint8 dataset (forward_from_buffer) vs the reference-style pipeline on the same
GPU (float windows + eager fp32/TF32 cuDNN convs).  usage: python tools/slac_encoder_bench.py [batch]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from s2p_b200 import _lib  # noqa: E402
from s2p_b200.networks.slac_encoder import SlacEncoderB200  # noqa: E402
from s2p_b200.slac_buffer import SlacReplayBufferB200  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
N = 20_000
buf = SlacReplayBufferB200(N, 8, (3, 100, 100), (6,), "cuda")
buf.frames = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device="cuda")
buf.frames_tp1 = torch.randint(0, 256, (N, 100, 100, 3), dtype=torch.uint8, device="cuda")
rows = torch.arange(8, N, device="cuda")
buf.codes = torch.cat([2 * (rows.unsqueeze(1) - 8 + torch.arange(8, device="cuda")), (2 * (rows - 1) + 1).unsqueeze(1)], 1)
buf._n = buf.codes.shape[0]
enc = SlacEncoderB200(3, 256, 100).cuda()
idx = np.random.RandomState(0).randint(0, buf._n, size=B)


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


n0 = _lib.launch_count()
enc.forward_from_buffer(buf, idx)
launches = _lib.launch_count() - n0
ms = timed(lambda: enc.forward_from_buffer(buf, idx))
ref = torch.nn.Sequential(*[m if isinstance(m, torch.nn.LeakyReLU) else torch.nn.Conv2d(m.in_channels, m.out_channels, m.kernel_size, m.stride, m.padding) for m in enc.net]).cuda()
ms_ref = timed(lambda: ref(buf.gather_states(idx).view(B * 9, 3, 100, 100)))
print(f"SLAC encoder, batch {B} x 9 frames: {ms:.3f} ms ({B * 9 / ms * 1e3:.0f} frames/s, {launches} launches) fused from uint8; "
      f"float windows + eager cuDNN fp32 on the same GPU: {ms_ref:.3f} ms")
