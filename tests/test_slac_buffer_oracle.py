"""CPU: the numpy restatement of the SLAC buffer + loader loops (oracle/slac_buffer_oracle.py) against
what the REFERENCE'S OWN ReplayBuffer held and sampled (tests/golden/slac_buffer.npz, made by
tests/golden/make_slac_buffer_golden.py from /root/reference/rlkit/torch/slac/buffer.py)."""
import numpy as np
import pytest

import slac_buffer_oracle as SO
from common import load_golden
from slac_dataset import H, W, C, A, loaded, make_dataset


def _dump(buf):
    n = buf._n
    return dict(n=n, p=buf._p, states=np.stack([np.array(buf.state_[k], dtype=np.uint8) for k in range(n)]),
                action=buf.action_[:n], reward=buf.reward_[:n], done=buf.done_[:n])


@pytest.mark.parametrize("tag,ut,lam", [("gen_none", None, None), ("gen_max", "max_of_both", 0.7),
                                        ("gen_dis", "disagreement", 1.5)])
def test_generated_fill_matches_reference_buffer(tag, ut, lam):
    g = load_golden("slac_buffer")
    d = make_dataset(int(g["seed_gen"]), list(g["lengths"]), 0)
    buf = SO.ReplayBuffer(1000, 8, (C, H, W), (A,))
    SO.fill_generated(buf, loaded(d), ut, lam)
    for k, v in _dump(buf).items():
        assert np.array_equal(np.asarray(v), g[f"{tag}_{k}"]), (tag, k)


@pytest.mark.parametrize("tag,size", [("plain", 1000), ("plain_ring", 17)])
def test_plain_fill_matches_reference_buffer(tag, size):
    g = load_golden("slac_buffer")
    d = make_dataset(int(g["seed_plain"]), list(g["lengths"]), int(g["tail_plain"]))
    buf = SO.ReplayBuffer(size, 8, (C, H, W), (A,))
    SO.fill_plain(buf, loaded(d))
    for k, v in _dump(buf).items():
        assert np.array_equal(np.asarray(v), g[f"{tag}_{k}"]), (tag, k)


def test_sampling_matches_reference_buffer():
    g = load_golden("slac_buffer")
    d = make_dataset(int(g["seed_gen"]), list(g["lengths"]), 0)
    buf = SO.ReplayBuffer(1000, 8, (C, H, W), (A,))
    SO.fill_generated(buf, loaded(d), "disagreement", 1.5)
    rng = np.random.RandomState(int(g["sample_seed"]))
    for tag, fn, b in (("lat", buf.sample_latent, 5), ("sac", buf.sample_sac, 4)):
        st, ac, rw, dn = fn(b, rng)
        assert np.array_equal(st, g[f"{tag}_state"]) and st.dtype == np.float32
        assert np.array_equal(ac, g[f"{tag}_action"]) and np.array_equal(rw, g[f"{tag}_reward"])
        assert np.array_equal(dn, g[f"{tag}_done"])
