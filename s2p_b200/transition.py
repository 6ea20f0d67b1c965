"""State-transition producer: host-side mirror of the reference's ensemble world model and of
its ``all_state_1step_random_action`` rollout (SURVEY.md section 8 f, rank 1).

* ``EnsembleTransitionB200`` has the constructor, parameter names/shapes, ``forward``,
  ``set_select`` and ``update_save`` of ``EnsembleTransition`` (/root/reference/gaussian_ensemble.py:59-108,
  ``EnsembleLinear`` :21-57), so ``load_state_dict(torch.load('model_dist_state_dict_590.pkl'))``
  (state_transition_rollout.py:100) works unchanged; ``forward`` returns the same
  ``torch.distributions.Normal`` but is ONE launch of ``ensemble_transition_kernel``.
* ``generate_rollout`` is the body of ``generate_rollout_by_dynamics_reward_ftn``
  (state_transition_rollout.py:134-229) with the per-trajectory model calls replaced by one
  fused launch over the whole dataset; np.random is consumed in the reference's order so the
  same seed gives the same actions and member choices.

No CPU path: CPU tensors or a missing library raise RuntimeError.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .ops import _ptr, _require_cuda, _stream

HIDDEN = 256
INTEGER_INF = int(1e9)        # state_transition_rollout.py:111
SLAC_NUM_SEQUENCES = 8        # state_transition_rollout.py:16
_IN_PAD, _OUT_PAD = 32, 64


class EnsembleLinearB200(nn.Module):
    """Parameter container with the names of the reference's EnsembleLinear (gaussian_ensemble.py:21-35)."""

    def __init__(self, in_features, out_features, ensemble_size=7):
        super().__init__()
        self.ensemble_size = ensemble_size
        self.weight = nn.Parameter(torch.zeros(ensemble_size, in_features, out_features))
        self.bias = nn.Parameter(torch.zeros(ensemble_size, 1, out_features))
        nn.init.trunc_normal_(self.weight, std=1 / (2 * in_features ** 0.5))
        self.saved_weight = nn.Parameter(self.weight.detach().clone())
        self.saved_bias = nn.Parameter(self.bias.detach().clone())
        self.select = list(range(ensemble_size))

    def set_select(self, indexes):  # gaussian_ensemble.py:50-54
        assert len(indexes) <= self.ensemble_size and max(indexes) < self.ensemble_size
        self.select = list(indexes)
        self.weight.data[indexes] = self.saved_weight.data[indexes]
        self.bias.data[indexes] = self.saved_bias.data[indexes]

    def update_save(self, indexes):  # gaussian_ensemble.py:56-58
        self.saved_weight.data[indexes] = self.weight.data[indexes]
        self.saved_bias.data[indexes] = self.bias.data[indexes]


class EnsembleTransitionB200(nn.Module):
    def __init__(self, obs_dim, action_dim, hidden_features=HIDDEN, hidden_layers=3, ensemble_size=7, mode="local",
                 with_reward=True):
        super().__init__()
        if hidden_features != HIDDEN or mode != "local" or not with_reward:
            raise RuntimeError("EnsembleTransitionB200 implements the reference's call site only: hidden_features=256, "
                               "mode='local', with_reward=True (state_transition_rollout.py:44-51)")
        if _lib.load().s2p_transition_member_floats(obs_dim, action_dim, hidden_layers) < 0:
            raise RuntimeError(f"unsupported dims obs_dim={obs_dim} action_dim={action_dim} hidden_layers={hidden_layers}")
        self.obs_dim, self.action_dim, self.mode, self.with_reward = obs_dim, action_dim, mode, with_reward
        self.ensemble_size, self.hidden_layers = ensemble_size, hidden_layers
        dims = [obs_dim + action_dim] + [hidden_features] * hidden_layers
        self.backbones = nn.ModuleList(EnsembleLinearB200(i, o, ensemble_size) for i, o in zip(dims[:-1], dims[1:]))
        self.output_layer = EnsembleLinearB200(hidden_features, 2 * (obs_dim + 1), ensemble_size)
        self.max_logstd = nn.Parameter(torch.ones(obs_dim + 1) * 1)
        self.min_logstd = nn.Parameter(torch.ones(obs_dim + 1) * -5)
        self._packed = None
        self._packed_key = None

    # ---- reference surface ---------------------------------------------------------------
    def set_select(self, indexes):
        for layer in self.backbones:
            layer.set_select(indexes)
        self.output_layer.set_select(indexes)
        self._packed = None

    def update_save(self, indexes):
        for layer in self.backbones:
            layer.update_save(indexes)
        self.output_layer.update_save(indexes)

    @property
    def select(self):
        return self.output_layer.select

    # ---- packing ---------------------------------------------------------------------------
    def _layers(self):
        return list(self.backbones) + [self.output_layer]

    def packed_weights(self) -> torch.Tensor:
        """[len(select), member_floats] fp32 in the layout of include/s2p_b200.h; rebuilt when a
        parameter or the selection changed."""
        layers = self._layers()
        key = (tuple(self.select),) + tuple((l.weight._version, l.bias._version, l.weight.data_ptr()) for l in layers)
        if self._packed is not None and key == self._packed_key:
            return self._packed
        sel = self.select
        parts = []
        for li, l in enumerate(layers):
            w = l.weight.detach()[sel].float()          # [E, in, out]
            b = l.bias.detach()[sel].float().reshape(len(sel), -1)
            if li == 0:
                w = torch.nn.functional.pad(w, (0, 0, 0, _IN_PAD - w.shape[1]))
            if li == len(layers) - 1:
                w = torch.nn.functional.pad(w, (0, _OUT_PAD - w.shape[2]))
                b = torch.nn.functional.pad(b, (0, _OUT_PAD - b.shape[1]))
            parts += [w.reshape(len(sel), -1), b]
        packed = torch.cat(parts, dim=1).contiguous()
        mf = _lib.load().s2p_transition_member_floats(self.obs_dim, self.action_dim, self.hidden_layers)
        assert packed.shape[1] == mf, (packed.shape, mf)
        self._packed, self._packed_key = packed, key
        return packed

    # ---- kernels ---------------------------------------------------------------------------
    def _launch(self, obs, action, obs_mean, obs_std, want_dist, member_idx, norm):
        _require_cuda(obs, action, self.max_logstd)
        T = obs.shape[0]
        dev = obs.device
        od = self.obs_dim + 1
        w = self.packed_weights()
        E = w.shape[0]
        mean_out = std_out = None
        if want_dist:
            mean_out = torch.empty(E, T, od, dtype=torch.float32, device=dev)
            std_out = torch.empty(E, T, od, dtype=torch.float32, device=dev)
        sel_out = (None, None, None, None)
        nm = ns = None
        rmean = rstd = 0.0
        if member_idx is not None:
            nm, ns, rmean, rstd = norm
            sel_out = (torch.empty(T, self.obs_dim, dtype=torch.float32, device=dev),
                       torch.empty(T, dtype=torch.float32, device=dev),
                       torch.empty(T, 1, dtype=torch.float32, device=dev),
                       torch.empty(T, 1, dtype=torch.float32, device=dev))
        if T == 0:                      # empty batch: nothing to launch (the reference returns empty tensors)
            return mean_out, std_out, sel_out
        lo, hi = self.min_logstd.detach().float().contiguous(), self.max_logstd.detach().float().contiguous()
        rc = _lib.load().s2p_ensemble_transition_fwd(
            obs.data_ptr(), obs.stride(0), action.data_ptr(), action.stride(0), _ptr(obs_mean), _ptr(obs_std),
            w.data_ptr(), E, lo.data_ptr(), hi.data_ptr(), _ptr(mean_out), _ptr(std_out), _ptr(member_idx),
            _ptr(nm), _ptr(ns), float(rmean), float(rstd), *[_ptr(t) for t in sel_out], T, self.obs_dim,
            self.action_dim, self.hidden_layers, _stream())
        _lib.check(rc, "s2p_ensemble_transition_fwd")
        return mean_out, std_out, sel_out

    @torch.no_grad()
    def forward(self, obs_action):
        """obs_action [T, obs_dim+action_dim] (normalised obs | action) -> Normal(mu, std), each
        [len(select), T, obs_dim+1]  (gaussian_ensemble.py:83-96).  Inference only."""
        if obs_action.dim() != 2 or obs_action.shape[1] != self.obs_dim + self.action_dim:
            raise RuntimeError("EnsembleTransitionB200.forward expects [T, obs_dim+action_dim] (the rollout's 2-D call)")
        x = obs_action.float()
        if x.stride(1) != 1:
            x = x.contiguous()
        mu, std, _ = self._launch(x[:, :self.obs_dim], x[:, self.obs_dim:], None, None, True, None, None)
        return torch.distributions.Normal(mu, std, validate_args=False)

    @torch.no_grad()
    def rollout_rows(self, observations, actions, member_idx, normalize_configs, want_dist=False):
        """Fused per-row step of the rollout script for ALL rows at once: raw observations [T,obs_dim],
        actions [T,A], member_idx [T] int32 -> (next_obs [T,obs_dim], reward [T], disagreement [T,1],
        aleatoric [T,1]) de-normalised as state_transition_rollout.py:194-204."""
        dev = observations.device
        f = lambda k: torch.as_tensor(np.asarray(normalize_configs[k], dtype=np.float32)).to(dev).contiguous()
        om, osd, nm, ns = f("obs_mean"), f("obs_std"), f("next_obs_mean"), f("next_obs_std")
        mu, std, sel = self._launch(observations.float().contiguous(), actions.float().contiguous(), om, osd, want_dist,
                                    member_idx.to(torch.int32).contiguous(),
                                    (nm, ns, float(normalize_configs["reward_mean"]), float(normalize_configs["reward_std"])))
        return sel if not want_dist else (sel, mu, std)


def slac_indices(traj_start: torch.Tensor, num_seq: int = SLAC_NUM_SEQUENCES):
    """generate_obs_act_indices (state_transition_rollout.py:105-132) for every row at once.
    traj_start [N] int64 (index of the first row of the row's trajectory) ->
    (slac_observation_indices [N,num_seq+1], slac_action_indices [N,num_seq]) int64."""
    _require_cuda(traj_start)
    ts = traj_start.to(torch.int64).contiguous()
    n = ts.shape[0]
    obs = torch.empty(n, num_seq + 1, dtype=torch.int64, device=ts.device)
    act = torch.empty(n, num_seq, dtype=torch.int64, device=ts.device)
    _lib.check(_lib.load().s2p_slac_indices(ts.data_ptr(), obs.data_ptr(), act.data_ptr(), n, num_seq, INTEGER_INF,
                                            _stream()), "s2p_slac_indices")
    return obs, act


def generate_rollout(model: EnsembleTransitionB200, observations, timeouts, normalize_configs, act_low, act_high,
                     rng=np.random):
    """The ``all_state_1step_random_action`` pass (state_transition_rollout.py:134-229).

    observations [N,obs_dim] float32 and timeouts [N] are host arrays (the HDF5 columns, :62,73);
    returns the dict of dataset keys the script replaces / adds (:222-229) as host numpy arrays.
    Rows after the last timeout are dropped, as in the reference.  np.random is consumed exactly
    as the reference does (per trajectory: uniform actions :175, then randint member :192)."""
    dev = next(model.parameters()).device
    if dev.type != "cuda":
        raise RuntimeError("generate_rollout needs the model on a CUDA device (no CPU fallback)")
    observations = np.ascontiguousarray(observations, dtype=np.float32)
    timeout_idx = np.sort(np.where(np.asarray(timeouts) == 1)[0])  # :74
    if len(timeout_idx) == 0:
        raise NotImplementedError("no timeout in the dataset (state_transition_rollout.py:75-76)")
    n = int(timeout_idx[-1]) + 1
    starts = np.concatenate([[0], timeout_idx[:-1] + 1])
    lengths = timeout_idx - starts + 1
    assert (lengths > SLAC_NUM_SEQUENCES).all(), "traj length must exceed slac num seq (:106)"
    A = len(act_low)
    E = len(model.select)
    actions = np.empty((n, A), dtype=np.float32)
    member = np.empty(n, dtype=np.int32)
    for s, m in zip(starts, lengths):  # host RNG only; same call order as the reference loop
        actions[s:s + m] = rng.uniform(low=act_low, high=act_high, size=(m, A)).astype(np.float32)
        member[s:s + m] = rng.randint(0, E, size=m)
    traj_start = torch.from_numpy(np.repeat(starts, lengths).astype(np.int64)).to(dev)
    nxt, rew, dis, ale = model.rollout_rows(torch.from_numpy(observations[:n]).to(dev), torch.from_numpy(actions).to(dev),
                                            torch.from_numpy(member).to(dev), normalize_configs)
    obs_idx, act_idx = slac_indices(traj_start)
    return {
        "actions": actions, "rewards": rew.cpu().numpy(), "next_observations": nxt.cpu().numpy(),
        "disagreement_uncertainty": dis.cpu().numpy(), "aleatoric_uncertainty": ale.cpu().numpy(),
        "slac_action_indices": act_idx.cpu().numpy(), "slac_observation_indices": obs_idx.cpu().numpy(),
    }
