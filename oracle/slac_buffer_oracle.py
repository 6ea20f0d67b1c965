"""CPU oracle of the dataset loader / window assembly on the consumer side of the generator
(SURVEY.md section 8 f, rank 2).

TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py's cpu_baseline leg); s2p_b200/ never imports it.

Restates, citations relative to /root/reference:
* ``SequenceBuffer`` / ``ReplayBuffer``  rlkit/torch/slac/buffer.py:22-147 (numpy, no torch);
* the two fill loops of ``SlacAlgorithm.load_data_in_buffer``  rlkit/torch/slac/algo.py:278-353
  (generated ``all_state_1step_random_action`` data: 8 real frames + the generated frame t+1 of row
  i-1, generated action / penalised reward last) and :368-405 (plain trajectories: sliding windows
  over image_observations_tp1);
* the HWC -> CHW transpose of the frames at load time  algo.py:174-175,189-190.

PARITY PINNED: tests/golden/make_slac_buffer_golden.py runs the REFERENCE'S OWN
``rlkit/torch/slac/buffer.py`` ReplayBuffer (imported by file path; it needs only numpy + torch)
under these same fill loops and stores what it holds and samples; tests hold this restatement and
the CUDA path to those vectors.  (``load_data_in_buffer`` itself needs h5py and a constructed
SlacAlgorithm, so its loops are restated around the real buffer class.)
"""
from __future__ import annotations

from collections import deque

import numpy as np

INTEGER_INF = int(1e9)  # algo.py:276


class SequenceBuffer:  # buffer.py:22-68
    def __init__(self, num_sequences=8):
        self.num_sequences = num_sequences
        self._reset_episode = False
        self.state_ = deque(maxlen=num_sequences + 1)
        self.action_ = deque(maxlen=num_sequences)
        self.reward_ = deque(maxlen=num_sequences)
        self.done_ = deque(maxlen=num_sequences)

    def reset(self):
        self._reset_episode = False
        self.state_.clear(); self.action_.clear(); self.reward_.clear(); self.done_.clear()

    def reset_episode(self, state):
        assert not self._reset_episode
        self._reset_episode = True
        self.state_.append(state)

    def append(self, action, reward, done, next_state):
        assert self._reset_episode
        self.action_.append(action); self.reward_.append([reward]); self.done_.append([done])
        self.state_.append(next_state)

    def is_full(self):
        return len(self.reward_) == self.num_sequences


class ReplayBuffer:  # buffer.py:71-147
    def __init__(self, buffer_size, num_sequences, state_shape, action_shape):
        self._n = 0
        self._p = 0
        self.buffer_size, self.num_sequences = buffer_size, num_sequences
        self.state_shape, self.action_shape = state_shape, action_shape
        self.state_ = [None] * buffer_size
        self.action_ = np.empty((buffer_size, num_sequences, *action_shape), dtype=np.float32)
        self.reward_ = np.empty((buffer_size, num_sequences, 1), dtype=np.float32)
        self.done_ = np.empty((buffer_size, num_sequences, 1), dtype=np.float32)
        self.buff = SequenceBuffer(num_sequences)

    def reset_episode(self, state):
        self.buff.reset_episode(state)

    def append(self, action, reward, done, next_state, episode_done):
        self.buff.append(action, reward, done, next_state)
        if self.buff.is_full():
            self.state_[self._p] = list(self.buff.state_)
            self.action_[self._p] = np.array(self.buff.action_, dtype=np.float32)
            self.reward_[self._p] = np.array(self.buff.reward_, dtype=np.float32)
            self.done_[self._p] = np.array(self.buff.done_, dtype=np.float32)
            self._n = min(self._n + 1, self.buffer_size)
            self._p = (self._p + 1) % self.buffer_size
        if episode_done:
            self.buff.reset()

    def gather_states(self, idxes):
        """the state half of sample_latent / sample_sac (buffer.py:131-135): uint8 windows -> float / 255"""
        state_ = np.empty((len(idxes), self.num_sequences + 1, *self.state_shape), dtype=np.uint8)
        for i, idx in enumerate(idxes):
            state_[i, ...] = np.array(self.state_[idx], dtype=np.uint8)
        return state_.astype(np.float32) / np.float32(255.0)

    def sample_latent(self, batch_size, rng=np.random):
        idxes = rng.randint(low=0, high=self._n, size=batch_size)
        return self.gather_states(idxes), self.action_[idxes], self.reward_[idxes], self.done_[idxes]

    def sample_sac(self, batch_size, rng=np.random):
        idxes = rng.randint(low=0, high=self._n, size=batch_size)
        return self.gather_states(idxes), self.action_[idxes], self.reward_[idxes, -1], self.done_[idxes, -1]


def to_chw(frames_hwc):
    """algo.py:174-175: [N,H,W,C] -> [N,C,H,W]"""
    return np.transpose(frames_hwc, (0, 3, 1, 2))


def penalised_reward(dataset, k, uncertainty_type, lam):
    """algo.py:319-333"""
    r = dataset["rewards"][k]
    if uncertainty_type is None:
        return r
    a = dataset.get("aleatoric_uncertainty")
    d = dataset.get("disagreement_uncertainty")
    if uncertainty_type == "aleatoric":
        return r - lam * a[k]
    if uncertainty_type == "disagreement":
        return r - lam * d[k]
    if uncertainty_type == "max_of_both":
        return r - lam * max(a[k], d[k])
    if uncertainty_type == "min_of_both":
        return r - lam * min(a[k], d[k])
    if uncertainty_type == "average_both":
        return r - lam * 0.5 * (a[k] + d[k])
    raise NotImplementedError(uncertainty_type)


def fill_generated(buffer, dataset, uncertainty_type=None, lam=None, num_sequences=8):
    """algo.py:264-353: ``dataset`` holds CHW frames (after to_chw) and 1-D uncertainties (squeezed,
    :198-203).  `buffer` is any object with the reference ReplayBuffer's reset_episode / append."""
    n = dataset["actions"].shape[0]
    buffer.buff.reset()  # :270
    for i in range(n):
        obs_idx = dataset["slac_observation_indices"][i]
        act_idx = dataset["slac_action_indices"][i]
        assert (act_idx == obs_idx[:-1]).all()  # :286
        if (obs_idx >= INTEGER_INF).any():  # :288-290
            assert (obs_idx >= INTEGER_INF).all()
            continue
        if i == n - 1 and dataset["timeouts"][i]:  # :293-297
            break
        previous_state = dataset["image_observations"][obs_idx]  # :301
        previous_act = dataset["original_actions"][act_idx]
        previous_rew = dataset["original_rewards"][act_idx]
        previous_timeout = dataset["timeouts"][act_idx]
        buffer.reset_episode(previous_state[0])  # :306-310
        for j in range(num_sequences):
            if previous_timeout[j]:
                raise NotImplementedError  # :313-315
            if j == num_sequences - 1:  # :317-336
                action = dataset["actions"][i - 1]
                reward = penalised_reward(dataset, i - 1, uncertainty_type, lam)
                state = dataset["image_observations_tp1"][i - 1]
                done = True
            else:  # :337-341
                action, reward, state, done = previous_act[j], previous_rew[j], previous_state[j + 1], previous_timeout[j]
            buffer.append(action, reward, False, state, done)  # mask = False :343-348
            if done:
                break


def fill_plain(buffer, dataset, reset_buff=False):
    """algo.py:367-405: plain trajectories (real data and the multi-step generated modes)."""
    n = dataset["actions"].shape[0]
    state = dataset["image_observations"][0]
    if reset_buff:
        buffer.buff.reset()
    buffer.reset_episode(state)
    for i in range(n):
        if i == n - 1 and dataset["timeouts"][i]:  # :374-378
            break
        timeout = dataset["timeouts"][i]
        done = timeout
        state = dataset["image_observations_tp1"][i]
        buffer.append(dataset["actions"][i], dataset["rewards"][i], False, state, done)  # :393
        if done:
            if i == n - 1:
                break
            buffer.reset_episode(dataset["image_observations"][i + 1])  # :400-403


def multistep_rows(dataset, starts, history, gen):
    """Row layout of the multi-step generated datasets (random_state_5step_*), written as the per-trajectory loop a
    producer script would run; s2p_b200.synthesis.assemble_multistep_rows is the vectorised product form.  The
    reference's own producer for these modes is absent; what IS the reference's is how the
    result is consumed: the plain-trajectory branch of load_data_in_buffer (algo.py:355-405, fill_plain above with
    reset_buff=True) -- a generated trajectory must therefore hold num_sequences rows to yield a window."""
    keys = ("observations", "actions", "rewards", "next_observations", "image_observations", "image_observations_tp1",
            "disagreement_uncertainty", "aleatoric_uncertainty", "timeouts")
    rows = {k: [] for k in keys}
    steps = gen["actions"].shape[1]
    for j, i in enumerate(starts):
        for r in range(i - history, i):          # the real prefix
            rows["observations"].append(dataset["observations"][r])
            rows["actions"].append(dataset["actions"][r])
            rows["rewards"].append(np.float32(np.asarray(dataset["rewards"]).reshape(-1)[r]))
            rows["next_observations"].append(dataset["next_observations"][r] if "next_observations" in dataset else dataset["observations"][r + 1])
            rows["image_observations"].append(dataset["image_observations"][r])
            rows["image_observations_tp1"].append(dataset["image_observations_tp1"][r] if "image_observations_tp1" in dataset
                                                  else dataset["image_observations"][r + 1])
            rows["disagreement_uncertainty"].append(np.zeros(1, np.float32))
            rows["aleatoric_uncertainty"].append(np.zeros(1, np.float32))
            rows["timeouts"].append(np.float32(0))
        for k in range(steps):                   # the generated steps
            for key in keys[:-1]:
                rows[key].append(gen[key][j, k])
            rows["timeouts"].append(np.float32(1 if k == steps - 1 else 0))
    # sentinel: the consumer never appends a LAST row that is a timeout (algo.py:374-378); a trailing copy with
    # timeout 0 keeps the last generated trajectory's window and cannot complete one itself
    for key in keys[:-1]:
        rows[key].append(rows[key][-1])
    rows["timeouts"].append(np.float32(0))
    return {k: np.stack(v) for k, v in rows.items()}
