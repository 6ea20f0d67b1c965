"""Seeded synthetic dataset in the reference's on-disk schema, shared by the SLAC
buffer golden generator and its tests (same recipe, so fixtures only hold results)."""
import numpy as np

import slac_buffer_oracle as SO
import transition_oracle as TO

H, W, C, A = 12, 10, 3, 6


def make_dataset(seed, lengths, tail, h=H, w=W):
    rs = np.random.RandomState(seed)
    n = sum(lengths) + tail
    d = {
        "image_observations": rs.randint(0, 256, size=(n, h, w, C)).astype(np.uint8),
        "image_observations_tp1": rs.randint(0, 256, size=(n, h, w, C)).astype(np.uint8),
        "actions": rs.uniform(-1, 1, size=(n, A)).astype(np.float32),
        "original_actions": rs.uniform(-1, 1, size=(n, A)).astype(np.float32),
        "rewards": rs.randn(n).astype(np.float32),
        "original_rewards": rs.randn(n).astype(np.float32),
        "aleatoric_uncertainty": np.abs(rs.randn(n, 1)).astype(np.float32),
        "disagreement_uncertainty": np.abs(rs.randn(n, 1)).astype(np.float32),
        "timeouts": np.zeros(n, dtype=np.float32),
    }
    d["timeouts"][np.cumsum(lengths) - 1] = 1
    obs_i, act_i, start = [], [], 0
    for m in lengths:
        o, a = TO.obs_act_indices(m, start)
        obs_i.append(o); act_i.append(a); start += m
    if tail:
        o, a = TO.obs_act_indices(tail + 8, start)  # a trailing partial trajectory
        obs_i.append(o[:tail]); act_i.append(a[:tail])
    d["slac_observation_indices"] = np.concatenate(obs_i)
    d["slac_action_indices"] = np.concatenate(act_i)
    return d


def make_multistep(seed, lengths, tail, n_starts, steps, h=H, w=W, num_sequences=8):
    """a real dataset + a multi-step generated dataset built from it by the ORACLE's row loop with seeded stand-ins
    for the generated pieces (what is generated does not matter to the buffer semantics under test)"""
    real = make_dataset(seed, lengths, tail, h, w)
    n = real["actions"].shape[0]
    rs = np.random.RandomState(seed + 1000)
    real["observations"] = rs.randn(n, 17).astype(np.float32)
    history = num_sequences - steps
    t = real["timeouts"]
    pos = np.arange(n) - np.maximum.accumulate(np.where(np.concatenate([[0], t[:-1]]) == 1, np.arange(n), 0))
    eligible = np.arange(n)[pos >= history]
    starts = eligible[rs.randint(0, len(eligible), size=n_starts)]
    gen = {
        "observations": rs.randn(n_starts, steps, 17).astype(np.float32),
        "actions": rs.uniform(-1, 1, size=(n_starts, steps, A)).astype(np.float32),
        "rewards": rs.randn(n_starts, steps).astype(np.float32),
        "next_observations": rs.randn(n_starts, steps, 17).astype(np.float32),
        "image_observations": rs.randint(0, 256, size=(n_starts, steps, h, w, C)).astype(np.uint8),
        "image_observations_tp1": rs.randint(0, 256, size=(n_starts, steps, h, w, C)).astype(np.uint8),
        "disagreement_uncertainty": np.abs(rs.randn(n_starts, steps, 1)).astype(np.float32),
        "aleatoric_uncertainty": np.abs(rs.randn(n_starts, steps, 1)).astype(np.float32),
    }
    ms = SO.multistep_rows(real, starts, history, gen)
    ms["_starts"], ms["_gen"] = starts, gen
    return real, ms


def loaded(d):
    """what load_data_in_buffer holds after reading the file (algo.py:174-203): CHW frames, squeezed uncertainties"""
    e = {k: v for k, v in d.items() if not k.startswith("_")}
    e["image_observations"] = SO.to_chw(d["image_observations"])
    e["image_observations_tp1"] = SO.to_chw(d["image_observations_tp1"])
    e["aleatoric_uncertainty"] = d["aleatoric_uncertainty"].squeeze(-1)
    e["disagreement_uncertainty"] = d["disagreement_uncertainty"].squeeze(-1)
    return e
