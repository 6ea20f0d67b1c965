"""CPU oracle of the state-transition producer (SURVEY.md section 8 f, rank 1).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline leg; the product (s2p_b200/) never imports it.

numpy fp32 restatement of two pieces of the reference, citations relative to /root/reference:

* ``EnsembleTransition.forward``  gaussian_ensemble.py:83-96 (``EnsembleLinear.forward`` :37-48,
  ``Swish`` :9-11, ``soft_clamp`` :13-19) -- 7 MLPs 23 -> 256 -> 256 -> 256 -> 36, local mode
  (predicted delta + normalised observation), soft-clamped log-std;
* the ``all_state_1step_random_action`` loop of ``generate_rollout_by_dynamics_reward_ftn``
  state_transition_rollout.py:149-229 -- normalise observations, draw uniform actions and a
  random ensemble member per row (np.random, in the reference's call order), de-normalise the
  selected member's prediction, disagreement / aleatoric uncertainty, SLAC index tables
  (``generate_obs_act_indices`` :105-132).

PARITY PINNED: unlike the generator, this code IS in the mounted reference.
tests/golden/make_transition_golden.py imports the reference's own ``gaussian_ensemble`` module,
runs it on seeded inputs and stores its outputs; tests/test_transition_oracle.py holds this
restatement to those vectors.  The rollout script itself cannot be imported (gym, dmc2gym, h5py
are absent), so its loop is restated here line by line around the real model class.
"""
from __future__ import annotations

import numpy as np

HIDDEN = 256
INTEGER_INF = int(1e9)  # state_transition_rollout.py:111
SLAC_NUM_SEQUENCES = 8  # state_transition_rollout.py:16


def layer_names(hidden_layers=3):
    return [f"backbones.{i}" for i in range(hidden_layers)] + ["output_layer"]


def make_weights(obs_dim=17, action_dim=6, hidden=HIDDEN, hidden_layers=3, ensemble=7, seed=0, with_reward=True):
    """Seeded numpy weights with the reference's parameter names and shapes
    (gaussian_ensemble.py:27-28,80-81).  Biases are non-zero on purpose (the reference's zero
    init would not exercise the bias path)."""
    rs = np.random.RandomState(seed)
    out_dim = obs_dim + int(with_reward)
    dims = [obs_dim + action_dim] + [hidden] * hidden_layers + [2 * out_dim]
    sd = {}
    for name, (fi, fo) in zip(layer_names(hidden_layers), zip(dims[:-1], dims[1:])):
        std = 1.0 / (2.0 * fi ** 0.5)  # gaussian_ensemble.py:30
        sd[f"{name}.weight"] = np.clip(rs.randn(ensemble, fi, fo) * std, -2, 2).astype(np.float32)
        sd[f"{name}.bias"] = (rs.randn(ensemble, 1, fo) * 0.1).astype(np.float32)
    # spread the log-std half of the output bias so both sides of soft_clamp are exercised
    sd["output_layer.bias"][..., out_dim:] = (rs.randn(ensemble, 1, out_dim) * 3.0 - 2.0).astype(np.float32)
    sd["max_logstd"] = (np.ones(out_dim) * 1.0 + rs.randn(out_dim) * 0.1).astype(np.float32)
    sd["min_logstd"] = (np.ones(out_dim) * -5.0 + rs.randn(out_dim) * 0.1).astype(np.float32)
    return sd


def softplus(x):
    """F.softplus(beta=1, threshold=20) in fp32."""
    x = x.astype(np.float32)
    return np.where(x > 20.0, x, np.log1p(np.exp(np.minimum(x, 20.0), dtype=np.float32), dtype=np.float32))


def soft_clamp(x, _min, _max):
    """gaussian_ensemble.py:13-19"""
    x = _max - softplus(_max - x)
    x = _min + softplus(x - _min)
    return x.astype(np.float32)


def ensemble_forward(sd, obs_action, obs_dim=17, hidden_layers=3, select=None):
    """gaussian_ensemble.py:83-96.  obs_action [T, obs_dim+action_dim] -> (mu, std) each
    [E, T, obs_dim+1]; mu[..., :obs_dim] is the (normalised) next observation."""
    x = np.asarray(obs_action, dtype=np.float32)
    names = layer_names(hidden_layers)
    E = sd[f"{names[0]}.weight"].shape[0]
    sel = list(range(E)) if select is None else list(select)
    out = np.broadcast_to(x, (len(sel),) + x.shape)
    for name in names[:-1]:
        w, b = sd[f"{name}.weight"][sel], sd[f"{name}.bias"][sel]
        out = np.einsum("bij,bjk->bik", out, w, dtype=np.float32) + b  # :37-48
        out = (out / (1.0 + np.exp(-out, dtype=np.float32))).astype(np.float32)  # Swish :9-11
    w, b = sd["output_layer.weight"][sel], sd["output_layer.bias"][sel]
    out = np.einsum("bij,bjk->bik", out, w, dtype=np.float32) + b
    half = out.shape[-1] // 2
    mu, logstd = out[..., :half].copy(), out[..., half:]  # torch.chunk :87
    logstd = soft_clamp(logstd, sd["min_logstd"], sd["max_logstd"])  # :88
    mu[..., :obs_dim] += x[..., :obs_dim]  # mode == 'local' :90-93
    return mu.astype(np.float32), np.exp(logstd, dtype=np.float32)


def obs_act_indices(traj_length, traj_start_idx, num_seq=SLAC_NUM_SEQUENCES):
    """state_transition_rollout.py:105-132 (vectorised; same values)."""
    i = np.arange(traj_length)[:, None]
    obs = np.where(i < num_seq, INTEGER_INF, i - num_seq + np.arange(num_seq + 1)[None, :] + traj_start_idx)
    act = np.where(i < num_seq, INTEGER_INF, i - num_seq + np.arange(num_seq)[None, :] + traj_start_idx)
    return obs.astype(int), act.astype(int)


def select_and_uncertainty(mu, std, ensemble_idx, norm, obs_dim=17):
    """state_transition_rollout.py:181-204 for one batch of rows."""
    T = mu.shape[1]
    rows = np.arange(T)
    predicted_obs, predicted_rew = mu[:, :, :obs_dim], mu[:, :, -1]
    next_obs = predicted_obs[ensemble_idx, rows] * norm["next_obs_std"] + norm["next_obs_mean"]  # :194
    reward = predicted_rew[ensemble_idx, rows] * norm["reward_std"] + norm["reward_mean"]  # :195
    mode = mu[:, :, :-1]  # :199
    diff = mode - mode.mean(axis=0)  # :201-202
    disagreement = np.sqrt((diff * diff).sum(-1, dtype=np.float32))[..., None].max(axis=0)  # :203
    aleatoric = np.sqrt((std * std).sum(-1, dtype=np.float32))[..., None].max(axis=0)  # :204
    return (next_obs.astype(np.float32), reward.astype(np.float32), disagreement.astype(np.float32),
            aleatoric.astype(np.float32))


def rollout_1step_random_action(sd, observations, timeouts, norm, act_low, act_high, rng, obs_dim=17):
    """state_transition_rollout.py:134-229: one pass over every trajectory of the dataset.
    `rng` is a np.random.RandomState used in the reference's call order (uniform actions, then
    randint member index, per trajectory).  Returns the dict of replaced / added dataset keys."""
    observations = np.asarray(observations, dtype=np.float32)
    timeout_idx = np.sort(np.where(np.asarray(timeouts) == 1)[0])  # :74
    normalized_obs = (observations - norm["obs_mean"]) / norm["obs_std"]  # :149
    A = len(act_low)
    keys = ["actions", "rewards", "next_observations", "disagreement_uncertainty", "aleatoric_uncertainty",
            "slac_action_indices", "slac_observation_indices"]
    acc = {k: [] for k in keys}
    for idx, end_idx in enumerate(timeout_idx):
        start_idx = 0 if idx == 0 else timeout_idx[idx - 1] + 1  # :153-158
        n = end_idx - start_idx + 1
        obs_i, act_i = obs_act_indices(n, start_idx)  # :170
        actions = rng.uniform(low=act_low, high=act_high, size=(n, A)).astype(np.float32)  # :175
        mu, std = ensemble_forward(sd, np.concatenate([normalized_obs[start_idx:end_idx + 1], actions], -1), obs_dim)
        ensemble_idx = rng.randint(0, mu.shape[0], size=n)  # :192
        nxt, rew, dis, ale = select_and_uncertainty(mu, std, ensemble_idx, norm, obs_dim)
        for k, v in zip(keys, [actions, rew, nxt, dis, ale, act_i, obs_i]):
            acc[k].append(v)
    return {k: np.concatenate(v, axis=0) for k, v in acc.items()}  # :222-229
