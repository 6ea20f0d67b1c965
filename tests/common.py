"""Shared helpers for the tests: seeded inputs (same recipe as tests/golden/make_golden.py)
and the oracle builder.  Importing the oracle is allowed here (tests are the checker)."""
import json
import os

import numpy as np
import torch

import s2p_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def manifest():
    with open(os.path.join(GOLDEN, "manifest.json")) as f:
        return json.load(f)


def state_stats():
    with open(os.path.join(GOLDEN, "cheetah_state_stats.json")) as f:
        return json.load(f)


def make_inputs(env, batch, steps, seed_img=0, seed_state=1):
    g0 = torch.Generator().manual_seed(seed_img)
    g1 = torch.Generator().manual_seed(seed_state)
    image = torch.rand(batch, 3, 100, 100, generator=g0) * 2 - 1
    z = torch.randn(batch, steps, O.STATE_DIM[env], generator=g1)
    if env == "cheetah":
        st = state_stats()
        states = torch.tensor(st["obs_mean"]) + torch.tensor(st["obs_std"]) * z
    else:
        states = z
    return image, states


def build_oracle(env="cheetah", init_type="scaled", seed=1234):
    opt = O.make_options(env_type=env, init_type=init_type)
    torch.manual_seed(seed)
    return O.define_G(opt).eval()


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def psnr01(a, b):
    """PSNR with images mapped from [-1,1] to [0,1], peak = 1."""
    mse = ((a.double() - b.double()) * 0.5).pow(2).mean().item()
    return float("inf") if mse == 0 else 10.0 * np.log10(1.0 / mse)
