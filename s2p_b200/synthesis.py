"""Bulk augmented-dataset synthesis: the reference's unpublished "generation code"
(/root/reference/README.md:54, TODO in the reference) between the state-transition
rollout and the SLAC loader:

    image_observations_tp1[i] = netG(image_observations[i], next_observations[i])

Row schema (pinned by the reference on both sides): frames are uint8 HWC
[N,100,100,3] (rlkit/torch/slac/algo.py:174-175,189-190), states are float32 [N,S]
(state_transition_rollout.py:62-64,222-224).  Rows are independent, so N GPUs take
contiguous index ranges with no data-path collective (SURVEY.md section 8 e).

Host side of each rank: pinned staging buffers, a copy-in stream, the compute stream
and a copy-out stream, double buffered so H2D of batch i+1 and D2H of batch i-1
overlap the kernels of batch i.
"""
from __future__ import annotations

import numpy as np
import torch


def shard_range(n_rows: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced partition [lo, hi) of n_rows over `world` ranks (the first
    n_rows % world ranks take one extra row)."""
    if world <= 0 or not (0 <= rank < world) or n_rows < 0:
        raise ValueError(f"bad shard request n_rows={n_rows} rank={rank} world={world}")
    base, rem = divmod(n_rows, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class BulkSynthesizer:
    def __init__(self, net, batch: int = 1024):
        self.net = net
        self.batch = int(batch)
        self.device = next(net.parameters()).device
        if self.device.type != "cuda":
            raise RuntimeError("BulkSynthesizer needs the generator on a CUDA device (no CPU fallback)")
        self._copy_in = torch.cuda.Stream(self.device)
        self._copy_out = torch.cuda.Stream(self.device)
        self._slots = None
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def _make_slots(self, h, w, s):
        b = self.batch
        slots = []
        for _ in range(2):
            slots.append(dict(
                frames=torch.empty(b, h, w, 3, dtype=torch.uint8, device=self.device),
                states=torch.empty(b, s, dtype=torch.float32, device=self.device),
                out=torch.empty(b, h, w, 3, dtype=torch.uint8, device=self.device),
                loaded=torch.cuda.Event(), computed=torch.cuda.Event(), drained=torch.cuda.Event(), used=False))
        self._slots = (slots, (h, w, s))
        self._tick = 0

    @staticmethod
    def _as_host_tensor(a, dtype):
        t = torch.from_numpy(a) if isinstance(a, np.ndarray) else a
        if t.dtype != dtype or t.is_cuda or not t.is_contiguous():
            raise RuntimeError(f"expected a contiguous host {dtype} array")
        return t

    @torch.no_grad()
    def synthesize(self, frames, states, out=None):
        """frames: host uint8 [N,H,W,3]; states: host float32 [N,S]; returns host uint8
        [N,H,W,3] (written into `out` when given; pinned memory makes the copies async)."""
        frames = self._as_host_tensor(frames, torch.uint8)
        states = self._as_host_tensor(states, torch.float32)
        n, h, w, _ = frames.shape
        if states.shape[0] != n:
            raise RuntimeError("frames and states disagree on the number of rows")
        if out is None:
            out = torch.empty_like(frames, pin_memory=True)
        out = self._as_host_tensor(out, torch.uint8)
        if self._slots is None or self._slots[1] != (h, w, states.shape[1]):
            self._make_slots(h, w, states.shape[1])
        slots = self._slots[0]
        compute = torch.cuda.current_stream(self.device)
        nb = (n + self.batch - 1) // self.batch
        for i in range(nb):
            lo, hi = i * self.batch, min(n, (i + 1) * self.batch)
            m = hi - lo
            # slots alternate across calls too: back-to-back synthesize() calls (no host sync in between) keep the
            # copy / compute overlap and never reuse a slot whose previous batch is still in flight
            sl = slots[self._tick % 2]
            self._tick += 1
            with torch.cuda.stream(self._copy_in):
                if sl["used"]:
                    self._copy_in.wait_event(sl["computed"])      # slot's inputs consumed
                sl["frames"][:m].copy_(frames[lo:hi], non_blocking=True)
                sl["states"][:m].copy_(states[lo:hi], non_blocking=True)
                sl["loaded"].record(self._copy_in)
            compute.wait_event(sl["loaded"])
            if sl["used"]:
                compute.wait_event(sl["drained"])                  # slot's output copied out
            sl["used"] = True
            res = self.net.forward_u8(sl["frames"][:m], sl["states"][:m])
            sl["out"][:m].copy_(res)
            sl["computed"].record(compute)
            with torch.cuda.stream(self._copy_out):
                self._copy_out.wait_event(sl["computed"])
                out[lo:hi].copy_(sl["out"][:m], non_blocking=True)
                sl["drained"].record(self._copy_out)
            self.h2d_bytes += m * (h * w * 3 + states.shape[1] * 4)
            self.d2h_bytes += m * h * w * 3
        compute.wait_stream(self._copy_out)
        return out

    def synthesize_shard(self, frames, states, rank: int, world: int, out=None):
        """This rank's contiguous slice of a dataset every rank can see (e.g. a memory
        map): returns (lo, hi, frames_tp1[lo:hi])."""
        lo, hi = shard_range(frames.shape[0], rank, world)
        res = self.synthesize(frames[lo:hi], states[lo:hi], out)
        return lo, hi, res


def synthesize_augmented_dataset(netG, transition_model, dataset, normalize_configs, act_low, act_high, rng=np.random,
                                 batch: int = 1024):
    """The whole "generation code" the reference leaves as a TODO (/root/reference/README.md:54), end to end on the
    GPU: state-transition rollout (state_transition_rollout.py:134-229, ``transition.generate_rollout``) -> S2P
    generator on (image_observations[i], predicted next_observations[i]) -> the ``..._dataset_augment.hdf5``
    key set the SLAC loader reads (rlkit/torch/slac/algo.py:183-204).

    ``dataset``: dict of host arrays in the real dataset's schema (observations, actions, rewards, timeouts,
    image_observations uint8 HWC, ...).  Returns a dict of host arrays; rows after the last timeout are dropped
    as in the reference's rollout script."""
    from .transition import generate_rollout
    roll = generate_rollout(transition_model, dataset["observations"], dataset["timeouts"], normalize_configs, act_low,
                            act_high, rng)
    n = roll["actions"].shape[0]
    out = {k: np.asarray(v)[:n] for k, v in dataset.items()}
    out["original_actions"] = np.asarray(dataset["actions"])[:n]       # state_transition_rollout.py:70-71
    out["original_rewards"] = np.asarray(dataset["rewards"])[:n]
    out.update(roll)                                                    # :222-229
    frames = torch.from_numpy(np.ascontiguousarray(out["image_observations"]))
    states = torch.from_numpy(np.ascontiguousarray(roll["next_observations"], dtype=np.float32))
    tp1 = BulkSynthesizer(netG, batch=min(batch, n)).synthesize(frames, states)
    torch.cuda.current_stream().synchronize()
    out["image_observations_tp1"] = tp1.numpy()                         # what algo.py:175,190,336 consume
    return out


MULTISTEP_MODES = ("random_state_5step_random_action", "random_state_5step_offRL_action", "random_state_1step_random_action")


def multistep_start_rows(timeouts, n_starts, history, rng=np.random):
    """rows a multi-step rollout may start from: `history` earlier rows in the same trajectory exist (they become the
    real prefix of the generated trajectory).  Drawn with ONE rng.randint call over the eligible rows."""
    t = np.asarray(timeouts).reshape(-1)
    n = t.shape[0]
    idx = np.arange(n)
    start_mark = np.zeros(n, dtype=np.int64)
    start_mark[1:] = np.where(t[:-1] == 1, idx[1:], 0)
    traj_start = np.maximum.accumulate(start_mark)
    eligible = idx[(idx - traj_start) >= history]
    if len(eligible) == 0:
        raise RuntimeError(f"no row has {history} earlier rows in its trajectory")
    return eligible[rng.randint(0, len(eligible), size=n_starts)]


def assemble_multistep_rows(dataset, starts, history, gen):
    """row assembly of a multi-step generated dataset (host, numpy; index bookkeeping only).  Per start row i:
    `history` REAL rows i-history .. i-1 (their own actions / rewards / frames) followed by `steps` GENERATED rows
    (gen[...][j, k]: observations s_k, actions a_k, rewards r_k, next_observations s_{k+1}, frames frame_k ->
    frame_{k+1}, uncertainties), timeouts = 1 on the last generated row; one sentinel row closes the file (see the
    end of this function).  Keys and layouts are the ones
    load_data_in_buffer reads for the random_state_* modes (rlkit/torch/slac/algo.py:212-236)."""
    starts = np.asarray(starts, dtype=np.int64)
    J, steps = gen["actions"].shape[:2]
    L = history + steps
    real = (starts[:, None] - history + np.arange(history)[None, :]).reshape(-1)       # [J*history] dataset rows
    out = {}

    def put(key, real_vals, gen_vals):
        gv = np.asarray(gen_vals)
        a = np.empty((J, L) + gv.shape[2:], dtype=gv.dtype)
        a[:, :history] = np.asarray(real_vals).reshape((J, history) + gv.shape[2:])
        a[:, history:] = gv
        out[key] = a.reshape((J * L,) + gv.shape[2:])

    img = np.asarray(dataset["image_observations"])
    tp1 = np.asarray(dataset["image_observations_tp1"]) if "image_observations_tp1" in dataset else None
    put("observations", np.asarray(dataset["observations"])[real], gen["observations"])
    put("actions", np.asarray(dataset["actions"])[real], gen["actions"])
    put("rewards", np.asarray(dataset["rewards"]).reshape(-1)[real], gen["rewards"])
    put("next_observations", np.asarray(dataset["next_observations"])[real] if "next_observations" in dataset
        else np.asarray(dataset["observations"])[real + 1], gen["next_observations"])
    put("image_observations", img[real], gen["image_observations"])
    put("image_observations_tp1", tp1[real] if tp1 is not None else img[real + 1], gen["image_observations_tp1"])
    zeros = np.zeros((J, history, 1), dtype=np.float32)
    put("disagreement_uncertainty", zeros, gen["disagreement_uncertainty"])
    put("aleatoric_uncertainty", zeros, gen["aleatoric_uncertainty"])
    t = np.zeros((J, L), dtype=np.float32)
    t[:, -1] = 1.0
    out["timeouts"] = t.reshape(-1)
    out["terminals"] = np.zeros(J * L, dtype=np.float32)
    step = np.zeros((J, L), dtype=np.int32)
    step[:, history:] = np.arange(1, steps + 1)
    out["generated_step"] = step.reshape(-1)          # 0 = real row, k = k-th generated step (not read by the consumer)
    out["source_row"] = np.repeat(starts, L)
    # the consumer's loop drops the dataset's LAST row when it is a timeout (rlkit/torch/slac/algo.py:374-378), which
    # would cost the last generated trajectory its window: end the file with one sentinel row (a copy of the last
    # row, timeout 0, generated_step -1) that can never complete a window of its own
    for k in list(out):
        out[k] = np.concatenate([out[k], out[k][-1:]])
    out["timeouts"][-1] = 0.0
    out["generated_step"][-1] = -1
    return out


def synthesize_multistep_dataset(netG, transition_model, dataset, normalize_configs, act_low, act_high, n_starts, steps=5,
                                 num_sequences=8, policy=None, rng=np.random, batch: int = 1024):
    """The multi-step generated-dataset modes the consumer recognises (rlkit/torch/slac/algo.py:212-236,
    finetune_rl.py:524): ``random_state_5step_random_action`` (policy=None: uniform random actions, as the 1-step
    script draws them, state_transition_rollout.py:175) and ``random_state_5step_offRL_action`` (policy = callable
    raw observation batch [T, obs_dim] torch tensor -> action batch [T, A]).  BASELINE configs[2]'s workload (5-step
    autoregressive rollout) turned into the dataset the SLAC loader's plain-trajectory branch reads
    (algo.py:355-405 with generated_for_slac=True):

      for each of n_starts random rows i of the real dataset:
        s_0 = observations[i], frame_0 = image_observations[i]
        for k in 0..steps-1:  a_k ~ U(act_low, act_high) | policy(s_k);  member ~ randint
                              (s_{k+1}, r_k, uncertainties) = ensemble step (EnsembleTransitionB200.rollout_rows)
        frames_{1..steps} = netG.rollout(frame_0, s_{1..steps})          (fp32 frames fed back unquantised, README.md:30-33)
        rows: num_sequences - steps real rows before i, then the generated rows, timeout on the last one

    so every generated trajectory holds exactly num_sequences transitions = one 9-frame SLAC window (real prefix +
    generated frames).  The reference's own producer for these modes is not in the snapshot (its launcher never even
    defines their file name, finetune_rl.py:536-548): the row layout above is this repo's, the key schema and the
    consumer's semantics are the reference's.  Returns a dict of host arrays."""
    dev = next(netG.parameters()).device
    history = num_sequences - steps
    if history < 0:
        raise ValueError("steps must not exceed num_sequences")
    obs_all = np.ascontiguousarray(dataset["observations"], dtype=np.float32)
    starts = multistep_start_rows(dataset["timeouts"], n_starts, history, rng)
    A, E = len(act_low), len(transition_model.select)
    J = len(starts)
    gen = {k: [] for k in ("observations", "actions", "rewards", "next_observations", "disagreement_uncertainty", "aleatoric_uncertainty")}
    s = torch.from_numpy(obs_all[starts]).to(dev)
    for k in range(steps):                       # host RNG in a fixed order: actions then member, per step
        if policy is None:
            a = torch.from_numpy(rng.uniform(low=act_low, high=act_high, size=(J, A)).astype(np.float32)).to(dev)
        else:
            a = policy(s).to(dev).float()
        member = torch.from_numpy(rng.randint(0, E, size=J).astype(np.int32)).to(dev)
        nxt, rew, dis, ale = transition_model.rollout_rows(s, a, member, normalize_configs)
        for key, v in (("observations", s), ("actions", a), ("rewards", rew), ("next_observations", nxt),
                       ("disagreement_uncertainty", dis), ("aleatoric_uncertainty", ale)):
            gen[key].append(v)
        s = nxt
    gen = {k: torch.stack(v, dim=1).cpu().numpy() for k, v in gen.items()}          # [J, steps, ...]
    # frames: the generator's N-step rollout from the start frames, in batches
    f0 = np.ascontiguousarray(np.asarray(dataset["image_observations"])[starts])     # uint8 [J,H,W,3]
    frames = np.empty((J, steps + 1) + f0.shape[1:], dtype=np.uint8)
    frames[:, 0] = f0
    states = torch.from_numpy(gen["next_observations"]).to(dev)
    with torch.no_grad():
        for lo in range(0, J, batch):
            hi = min(J, lo + batch)
            img = torch.from_numpy(f0[lo:hi]).to(dev).permute(0, 3, 1, 2).float().mul_(1.0 / 127.5).sub_(1.0)
            out = netG.rollout(img, states[lo:hi])                                  # [b, steps, 3, H, W] in [-1, 1]
            q = torch.floor((out + 1.0) * 127.5 + 0.5).clamp_(0, 255).to(torch.uint8).permute(0, 1, 3, 4, 2)
            frames[lo:hi, 1:] = q.cpu().numpy()
    gen["image_observations"], gen["image_observations_tp1"] = frames[:, :-1], frames[:, 1:]
    out = assemble_multistep_rows(dataset, starts, history, gen)
    out["data_mix_type"] = np.array(f"random_state_{steps}step_{'random' if policy is None else 'offRL'}_action")
    return out


# ---------------------------------------------------------------------------------------- flat-file twin of the HDF5 schema
def save_dataset_dir(path: str, dataset: dict) -> None:
    """Write every key of a dataset dict as `<path>/<key>.npy` -- a flat binary twin of the reference's HDF5 files
    (state_transition_rollout.py:232-243 writes the same keys with h5py, which this image does not have).  The
    arrays keep the on-disk schema of SURVEY.md Appendix B (frames uint8 HWC)."""
    import os
    os.makedirs(path, exist_ok=True)
    for k, v in dataset.items():
        np.save(os.path.join(path, f"{k}.npy"), np.ascontiguousarray(v))


def load_dataset_dir(path: str, keys=None, data_num=None, mmap: bool = True) -> dict:
    """Read a dataset written by save_dataset_dir; `data_num` is the row limit of load_data_in_buffer
    (rlkit/torch/slac/algo.py:183-204), `mmap` maps the big frame arrays instead of reading them, so every rank of a
    sharded synthesis run can open the same files and touch only its own rows (BulkSynthesizer.synthesize_shard)."""
    import os
    out = {}
    names = sorted(f[:-4] for f in os.listdir(path) if f.endswith(".npy")) if keys is None else list(keys)
    for k in names:
        a = np.load(os.path.join(path, f"{k}.npy"), mmap_mode="r" if mmap else None)
        out[k] = a if data_num is None else a[:data_num]
    return out
