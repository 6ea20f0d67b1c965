"""Golden vectors of the SLAC window assembly, produced by the REFERENCE'S OWN ReplayBuffer
(/root/reference/rlkit/torch/slac/buffer.py, imported by file path) filled by the loops of
SlacAlgorithm.load_data_in_buffer (rlkit/torch/slac/algo.py:264-353 and :367-405, restated in
oracle/slac_buffer_oracle.py because the method itself needs h5py + a constructed SlacAlgorithm).

Run in the build container:  python tests/golden/make_slac_buffer_golden.py
"""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import slac_buffer_oracle as SO  # noqa: E402
import transition_oracle as TO  # noqa: E402

spec = importlib.util.spec_from_file_location("ref_slac_buffer", "/root/reference/rlkit/torch/slac/buffer.py")
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

sys.path.insert(0, os.path.join(ROOT, "tests"))
from slac_dataset import A, C, H, W, loaded, make_dataset  # noqa: E402


def dump(buf):
    n = buf._n
    states = np.stack([np.array(buf.state_[k], dtype=np.uint8) for k in range(n)])
    return dict(n=n, p=buf._p, states=states, action=buf.action_[:n].numpy(), reward=buf.reward_[:n].numpy(),
                done=buf.done_[:n].numpy())


def main():
    out = {}
    # ---- generated data, all_state_1step_random_action (algo.py:264-353), three penalty flavours
    lengths = [20, 12, 25]
    d = make_dataset(21, lengths, 0)
    for tag, (ut, lam) in {"gen_none": (None, None), "gen_max": ("max_of_both", 0.7), "gen_dis": ("disagreement", 1.5)}.items():
        buf = ref.ReplayBuffer(1000, 8, (C, H, W), (A,), torch.device("cpu"))
        SO.fill_generated(buf, loaded(d), ut, lam)
        for k, v in dump(buf).items():
            out[f"{tag}_{k}"] = v
    # sampling through the reference's sample_latent / sample_sac (global np.random, buffer.py:131,142)
    np.random.seed(77)
    st, ac, rw, dn = buf.sample_latent(5)
    out.update(lat_state=st.numpy(), lat_action=ac.numpy(), lat_reward=rw.numpy(), lat_done=dn.numpy())
    st, ac, rw, dn = buf.sample_sac(4)
    out.update(sac_state=st.numpy(), sac_action=ac.numpy(), sac_reward=rw.numpy(), sac_done=dn.numpy())
    # ---- plain trajectories (algo.py:367-405), with a trailing partial trajectory and a small ring
    d2 = make_dataset(22, lengths, 11)
    for tag, size in {"plain": 1000, "plain_ring": 17}.items():
        buf = ref.ReplayBuffer(size, 8, (C, H, W), (A,), torch.device("cpu"))
        SO.fill_plain(buf, loaded(d2))
        for k, v in dump(buf).items():
            out[f"{tag}_{k}"] = v
    # ---- multi-step generated mode (random_state_5step_*): a real dataset, then a generated one through the same
    # plain branch with generated_for_slac=True (algo.py:355-405: buff.reset() first), in ONE reference buffer
    from slac_dataset import make_multistep  # noqa: E402
    real, ms = make_multistep(23, lengths, 11, n_starts=9, steps=5)
    buf = ref.ReplayBuffer(1000, 8, (C, H, W), (A,), torch.device("cpu"))
    SO.fill_plain(buf, loaded(real))
    SO.fill_plain(buf, loaded(ms), reset_buff=True)
    for k, v in dump(buf).items():
        out[f"multistep_{k}"] = v
    path = os.path.join(HERE, "slac_buffer.npz")
    np.savez_compressed(path, seed_gen=21, seed_plain=22, seed_multistep=23, multistep_starts=9, lengths=np.array(lengths), tail_plain=11, sample_seed=77, **out)
    print("wrote", path, os.path.getsize(path), {k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items() if k.endswith("_n") or k.endswith("states")})


if __name__ == "__main__":
    main()
