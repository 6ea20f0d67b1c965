"""Generates the golden vectors of ORACLE_SPEC v1 (oracle/s2p_oracle.py, fp32, CPU).

The reference has no golden vectors for this path and its
generator source is absent, so these pin the FROZEN ORACLE, not the reference
("parity unpinned").  Run from the repo root:  python tests/golden/make_golden.py
Inputs follow BASELINE.md section 3: image = rand(seed 0)*2-1, state = obs_mean +
obs_std * randn(seed 1) with the cheetah statistics of the reference's
world_model/.../normalize_configs_dict.pkl (copied into cheetah_state_stats.json by
this script when /root/reference is present), weights = oracle init under
torch.manual_seed(1234).
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import s2p_oracle as O  # noqa: E402

PKL = "/root/reference/world_model/cheetah-run-mixed_first_500k/ensemble/normalize_all/normalize_configs_dict.pkl"
STATS = os.path.join(HERE, "cheetah_state_stats.json")

CASES = [
    # name, env, init_type, batch, steps
    ("cheetah_xavier_b2", "cheetah", "xavier", 2, 1),
    ("cheetah_scaled_b2", "cheetah", "scaled", 2, 1),
    ("walker_scaled_b2", "walker", "scaled", 2, 1),
    ("cheetah_scaled_roll5_b1", "cheetah", "scaled", 1, 5),
    ("cheetah_trained_like_b2", "cheetah", "trained_like", 2, 1),
    ("walker_trained_like_roll5_b1", "walker", "trained_like", 1, 5),
]


def state_stats():
    if os.path.exists(PKL):
        d = torch.load(PKL, weights_only=False)
        js = {"obs_mean": [float(v) for v in d["obs_mean"]], "obs_std": [float(v) for v in d["obs_std"]],
              "source": "reference world_model/cheetah-run-mixed_first_500k/ensemble/normalize_all/normalize_configs_dict.pkl"}
        with open(STATS, "w") as f:
            json.dump(js, f, indent=1)
    with open(STATS) as f:
        return json.load(f)


def make_inputs(env, batch, steps, stats):
    g0 = torch.Generator().manual_seed(0)
    g1 = torch.Generator().manual_seed(1)
    image = torch.rand(batch, 3, 100, 100, generator=g0) * 2 - 1
    S = O.STATE_DIM[env]
    z = torch.randn(batch, steps, S, generator=g1)
    if env == "cheetah":
        mean = torch.tensor(stats["obs_mean"], dtype=torch.float32)
        std = torch.tensor(stats["obs_std"], dtype=torch.float32)
        states = mean + std * z
    else:
        states = z
    return image, states


def build(env, init_type):
    opt = O.make_options(env_type=env, init_type=init_type)
    torch.manual_seed(1234)
    return O.define_G(opt).eval()


def weight_checksum(net):
    return float(sum(p.detach().double().abs().sum() for p in net.parameters()))


def main():
    torch.set_num_threads(os.cpu_count())
    stats = state_stats()
    manifest = {"spec": O.SPEC_VERSION, "torch": torch.__version__, "cases": {}}
    for name, env, init, batch, steps in CASES:
        net = build(env, init)
        image, states = make_inputs(env, batch, steps, stats)
        with torch.no_grad():
            out = net.rollout(image, states)          # [B, steps, 3, 100, 100]
        last = out[:, -1]
        first = out[:, 0]
        np.savez_compressed(os.path.join(HERE, name + ".npz"), first=first.numpy(), last=last.numpy(),
                            u8_last=O.to_uint8_hwc(last).numpy())
        manifest["cases"][name] = dict(env=env, init_type=init, batch=batch, steps=steps,
                                       weight_abs_sum=weight_checksum(net),
                                       out_mean=float(last.mean()), out_std=float(last.std()),
                                       out_range=float(last.max() - last.min()),
                                       saturated_frac=float((last.abs() > 0.98).float().mean()))
        print(name, manifest["cases"][name])
    with open(os.path.join(HERE, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1)


if __name__ == "__main__":
    main()
