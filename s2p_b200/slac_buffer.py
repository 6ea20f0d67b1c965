"""SLAC window buffer: the consumer side of the generated frames (SURVEY.md section 8 f, rank 2).

Mirror of the reference's ``ReplayBuffer`` (/root/reference/rlkit/torch/slac/buffer.py:71-147) as it is
filled by ``SlacAlgorithm.load_data_in_buffer`` (rlkit/torch/slac/algo.py:154-416), B200-first:

* the reference walks the dataset row by row in Python (algo.py:278-353 / :368-405), pushing frames
  through a deque and keeping every 9-frame window as a list of host arrays; here the dataset's uint8
  HWC frames are uploaded ONCE as stored on disk and stay in HBM, and a window is 9 int64 codes
  (2*row + {0: image_observations, 1: image_observations_tp1}) computed for all rows at once with
  index arithmetic on the device -- no per-row loop, no 9x copy of the frames;
* ``sample_latent`` / ``sample_sac`` / ``random_batch`` return what the reference returns (float
  windows in [0,1] ``[B, 9, C, H, W]``, actions, rewards, dones) through ONE launch of
  ``window_gather_kernel`` (gather + HWC->CHW + /255).  Sample indices come from ``np.random`` exactly
  as in the reference (buffer.py:131,142,157), so a seeded run draws the same windows.

No CPU path: the buffer lives on a CUDA device.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .ops import _stream

INTEGER_INF = int(1e9)  # algo.py:276


class SlacReplayBufferB200:
    def __init__(self, buffer_size, num_sequences, state_shape, action_shape, device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("SlacReplayBufferB200 needs a CUDA device (no CPU fallback)")
        _lib.load()
        self.buffer_size, self.num_sequences = int(buffer_size), int(num_sequences)
        self.state_shape, self.action_shape = tuple(state_shape), tuple(action_shape)
        self._n = 0
        self._p = 0
        self._real_n = 0
        self.frames = self.frames_tp1 = None      # uint8 [N,H,W,C] in HBM, disk layout
        self.codes = None                          # int64 [buffer_size, num_sequences+1]
        self.action_ = torch.empty(self.buffer_size, self.num_sequences, *self.action_shape, device=self.device)
        self.reward_ = torch.empty(self.buffer_size, self.num_sequences, 1, device=self.device)
        self.done_ = torch.empty(self.buffer_size, self.num_sequences, 1, device=self.device)

    def __len__(self):
        return self._n

    # ------------------------------------------------------------------ loading
    def _upload_frames(self, dataset):
        c, h, w = self.state_shape
        out = []
        for k in ("image_observations", "image_observations_tp1"):
            a = dataset[k]
            t = torch.from_numpy(np.ascontiguousarray(a)) if isinstance(a, np.ndarray) else a
            if t.dtype != torch.uint8 or t.dim() != 4 or tuple(t.shape[1:]) != (h, w, c):
                raise RuntimeError(f"{k}: expected uint8 [N,{h},{w},{c}] (HWC, as stored on disk), got {tuple(t.shape)} {t.dtype}")
            out.append(t.to(self.device).contiguous())
        base = 0
        if self.frames is not None:   # a second dataset (generated after real, algo.py:407-412): rows are appended
            base = self.frames.shape[0]
            out = [torch.cat([self.frames, out[0]]), torch.cat([self.frames_tp1, out[1]])]
        self.frames, self.frames_tp1 = out
        return base

    def _dev(self, a, dtype):
        t = torch.from_numpy(np.ascontiguousarray(a)) if isinstance(a, np.ndarray) else a
        return t.to(self.device).to(dtype)

    def _store(self, codes, action, reward):
        """ring semantics of ReplayBuffer._append (buffer.py:118-125) for K windows appended in order"""
        K, S = codes.shape[0], self.buffer_size
        if self.codes is None:
            self.codes = torch.zeros(S, self.num_sequences + 1, dtype=torch.int64, device=self.device)
        k = torch.arange(K, device=self.device)
        keep = k >= K - S                     # older windows are overwritten
        pos = (self._p + k[keep]) % S
        self.codes[pos] = codes[keep]
        self.action_[pos] = action[keep]
        self.reward_[pos] = reward[keep].unsqueeze(-1)
        self.done_[pos] = 0.0                 # mask = False for every step (algo.py:343, :388)
        self._n = min(self._n + K, S)
        self._p = (self._p + K) % S

    def load_generated(self, dataset, uncertainty_type=None, uncertainty_penalty_lambda=None):
        """algo.py:264-353 (``generated_for_slac`` and ``all_state_1step_random_action``).  ``dataset``: dict of host
        arrays in the on-disk schema (frames uint8 HWC, uncertainties [N,1])."""
        ns = self.num_sequences
        base = self._upload_frames(dataset)
        n = dataset["actions"].shape[0]
        obs_idx = self._dev(dataset["slac_observation_indices"], torch.int64)
        act_idx = self._dev(dataset["slac_action_indices"], torch.int64)
        timeouts = self._dev(dataset["timeouts"], torch.float32).reshape(-1)
        if not torch.equal(act_idx, obs_idx[:, :-1]):
            raise AssertionError("slac_action_indices != slac_observation_indices[:, :-1] (algo.py:286)")
        inf_any, inf_all = (obs_idx >= INTEGER_INF).any(1), (obs_idx >= INTEGER_INF).all(1)
        if not torch.equal(inf_any, inf_all):
            raise AssertionError("partially masked slac_observation_indices row (algo.py:289)")
        keep = ~inf_any
        if bool(timeouts[n - 1] == 1):
            keep[n - 1] = False                                    # :293-297
        rows = torch.nonzero(keep).squeeze(1)                      # dataset row i of each window
        if (timeouts[act_idx[rows]] != 0).any():
            raise NotImplementedError("timeout inside a window (algo.py:313-315)")
        codes = torch.cat([2 * (obs_idx[rows, :ns] + base), (2 * (rows - 1 + base) + 1).unsqueeze(1)], dim=1)  # :301,:335
        actions = self._dev(dataset["actions"], torch.float32)
        orig_a = self._dev(dataset["original_actions"], torch.float32)
        orig_r = self._dev(dataset["original_rewards"], torch.float32).reshape(-1)
        action = torch.cat([orig_a[act_idx[rows, :ns - 1]], actions[rows - 1].unsqueeze(1)], dim=1)
        reward = torch.cat([orig_r[act_idx[rows, :ns - 1]], self._penalised(dataset, uncertainty_type,
                                                                           uncertainty_penalty_lambda)[rows - 1].unsqueeze(1)], 1)
        self._store(codes, action, reward)

    def _penalised(self, dataset, uncertainty_type, lam):
        """algo.py:319-333 for every row at once"""
        r = self._dev(dataset["rewards"], torch.float32).reshape(-1)
        if uncertainty_type is None:
            return r
        assert lam is not None
        a = self._dev(dataset["aleatoric_uncertainty"], torch.float32).reshape(-1) if "aleatoric_uncertainty" in dataset else None
        d = self._dev(dataset["disagreement_uncertainty"], torch.float32).reshape(-1) if "disagreement_uncertainty" in dataset else None
        if uncertainty_type == "aleatoric":
            return r - lam * a
        if uncertainty_type == "disagreement":
            return r - lam * d
        if uncertainty_type == "max_of_both":
            return r - lam * torch.maximum(a, d)
        if uncertainty_type == "min_of_both":
            return r - lam * torch.minimum(a, d)
        if uncertainty_type == "average_both":
            return r - lam * 0.5 * (a + d)
        raise NotImplementedError(uncertainty_type)

    def load_plain(self, dataset, real=True):
        """algo.py:367-405: plain trajectories -- sliding 9-frame windows [img[start], tp1[start], ...] per
        trajectory, one per row whose trajectory already holds num_sequences transitions."""
        ns = self.num_sequences
        base = self._upload_frames(dataset)
        n = dataset["actions"].shape[0]
        timeouts = self._dev(dataset["timeouts"], torch.float32).reshape(-1)
        idx = torch.arange(n, device=self.device)
        is_end = timeouts == 1
        start_mark = torch.zeros(n, dtype=torch.int64, device=self.device)
        start_mark[1:] = torch.where(is_end[:-1], idx[1:], torch.zeros_like(idx[1:]))
        traj_start = torch.cummax(start_mark, dim=0).values
        keep = (idx - traj_start) >= ns - 1
        if bool(is_end[n - 1]):
            keep[n - 1] = False                                    # :374-378
        rows = torch.nonzero(keep).squeeze(1)
        src = rows.unsqueeze(1) - ns + torch.arange(ns + 1, device=self.device).unsqueeze(0)   # i-8 .. i
        ts = traj_start[rows].unsqueeze(1)
        codes = torch.where(src < ts, 2 * (ts + base), 2 * (src + base) + 1)   # first frame of a trajectory is image_observations
        actions = self._dev(dataset["actions"], torch.float32)
        rewards = self._dev(dataset["rewards"], torch.float32).reshape(-1)
        win = rows.unsqueeze(1) - (ns - 1) + torch.arange(ns, device=self.device).unsqueeze(0)
        self._store(codes, actions[win], rewards[win])
        if real:
            self._real_n = self._n                                 # :410-412

    def load_data(self, dataset, generated_for_slac=False, data_mix_type=None, uncertainty_type=None,
                  uncertainty_penalty_lambda=None):
        """branch selection of SlacAlgorithm.load_data_in_buffer (rlkit/torch/slac/algo.py:267, :355-371) on a dataset
        dict: the 1-step generated mode -> load_generated; the multi-step generated modes (random_state_*) and real
        data -> the plain-trajectory loop (the reference resets its sequence buffer first for generated data, :370:
        windows never span two load calls here, which is the same thing)."""
        if generated_for_slac and data_mix_type == "all_state_1step_random_action":
            return self.load_generated(dataset, uncertainty_type, uncertainty_penalty_lambda)
        if generated_for_slac and data_mix_type not in ("random_state_5step_random_action", "random_state_1step_random_action",
                                                        "random_state_5step_offRL_action"):
            raise NotImplementedError(f"data_mix_type {data_mix_type} (algo.py:183-236 reads no such generated dataset)")
        return self.load_plain(dataset, real=not generated_for_slac)

    # ------------------------------------------------------------------ sampling
    def gather_states(self, idxes) -> torch.Tensor:
        """float windows [B, num_seq+1, C, H, W] in [0,1] (buffer.py:132-135)"""
        idx = torch.as_tensor(np.asarray(idxes), dtype=torch.int64, device=self.device)
        codes = self.codes[idx].contiguous()
        c, h, w = self.state_shape
        b, s = codes.shape
        out = torch.empty(b, s, c, h, w, dtype=torch.float32, device=self.device)
        _lib.check(_lib.load().s2p_slac_window_gather(self.frames.data_ptr(), self.frames_tp1.data_ptr(), self.frames.shape[0], codes.data_ptr(),
                                                      out.data_ptr(), b * s, h, w, c, _stream()), "s2p_slac_window_gather")
        return out

    def sample_latent(self, batch_size):
        idxes = np.random.randint(low=0, high=self._n, size=batch_size)
        return self.gather_states(idxes), self.action_[idxes], self.reward_[idxes], self.done_[idxes]

    def sample_sac(self, batch_size):
        idxes = np.random.randint(low=0, high=self._n, size=batch_size)
        return self.gather_states(idxes), self.action_[idxes], self.reward_[idxes, -1], self.done_[idxes, -1]

    def random_batch(self, batch_size):
        s, a, r, d = self.sample_sac(batch_size)
        return dict(observations=s, actions=a, rewards=r, terminals=d)
