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


def test_multistep_generated_mode_matches_reference_buffer():
    """random_state_5step_*: a real dataset then the generated one through the plain branch with the sequence
    buffer reset first (algo.py:355-405, generated_for_slac=True), one buffer: what the REFERENCE'S ReplayBuffer held"""
    from slac_dataset import make_multistep
    g = load_golden("slac_buffer")
    real, ms = make_multistep(int(g["seed_multistep"]), list(g["lengths"]), int(g["tail_plain"]), int(g["multistep_starts"]), 5)
    buf = SO.ReplayBuffer(1000, 8, (C, H, W), (A,))
    SO.fill_plain(buf, loaded(real))
    n_real = buf._n
    SO.fill_plain(buf, loaded(ms), reset_buff=True)
    assert buf._n - n_real == int(g["multistep_starts"])          # one 9-frame window per generated trajectory
    for k, v in _dump(buf).items():
        assert np.array_equal(np.asarray(v), g[f"multistep_{k}"]), k
    # a generated window = 4 real frames (the prefix) + the 5 generated frames, actions / rewards likewise
    starts, gen = ms["_starts"], ms["_gen"]
    w0 = np.array(buf.state_[n_real], dtype=np.uint8)
    i = starts[0]
    assert np.array_equal(w0[0], SO.to_chw(real["image_observations"][i - 3:i - 2])[0])
    assert np.array_equal(w0[4:], SO.to_chw(gen["image_observations_tp1"][0]))
    assert np.array_equal(buf.action_[n_real][3:], gen["actions"][0])


def test_product_row_assembly_equals_the_oracle_loop():
    """s2p_b200.synthesis.assemble_multistep_rows (vectorised) == oracle multistep_rows (per-trajectory loop)"""
    from s2p_b200.synthesis import assemble_multistep_rows, multistep_start_rows
    from slac_dataset import make_multistep
    real, ms = make_multistep(5, [20, 12, 25], 6, n_starts=7, steps=5)
    got = assemble_multistep_rows(real, ms["_starts"], 3, ms["_gen"])
    for k, v in ms.items():
        if not k.startswith("_"):
            assert got[k].dtype == v.dtype and np.array_equal(got[k], v), k
    assert got["generated_step"].tolist()[:8] == [0, 0, 0, 1, 2, 3, 4, 5] and got["generated_step"][-1] == -1
    assert got["timeouts"][-2] == 1 and got["timeouts"][-1] == 0 and len(got["timeouts"]) == 7 * 8 + 1
    # start rows: always at least `history` rows into their trajectory, drawn with one randint call
    st = multistep_start_rows(real["timeouts"], 50, 3, np.random.RandomState(1))
    t = real["timeouts"]
    for i in st:
        assert t[i - 3:i].sum() == 0
    rs = np.random.RandomState(1)
    elig = [i for i in range(len(t)) if i >= 3 and t[max(0, i - 3):i].sum() == 0 and (i - 3 == 0 or True)]
    assert len(st) == 50
