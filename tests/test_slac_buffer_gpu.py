"""GPU: SlacReplayBufferB200 (device-side window table + window_gather_kernel through the C ABI) against
(1) what the REFERENCE'S OWN ReplayBuffer held / sampled (tests/golden/slac_buffer.npz) and
(2) the numpy oracle at the real frame size (100x100x3), ring wrap-around and real+generated mixing.
uint8 -> float/255 is bit-exact, so every comparison is array_equal."""
import numpy as np
import pytest
import torch

import slac_buffer_oracle as SO
from common import load_golden
from slac_dataset import A, C, H, W, loaded, make_dataset

pytestmark = pytest.mark.gpu


def _buf(size=1000, h=H, w=W):
    from s2p_b200.slac_buffer import SlacReplayBufferB200
    return SlacReplayBufferB200(size, 8, (C, h, w), (A,), "cuda")


def _check_against(buf, g, tag):
    n = int(g[f"{tag}_n"])
    assert buf._n == n and buf._p == int(g[f"{tag}_p"])
    st = buf.gather_states(np.arange(n)).cpu().numpy()
    assert np.array_equal(st, g[f"{tag}_states"].astype(np.float32) / np.float32(255.0)), tag
    assert np.array_equal(buf.action_[:n].cpu().numpy(), g[f"{tag}_action"]), tag
    assert np.array_equal(buf.reward_[:n].cpu().numpy(), g[f"{tag}_reward"]), tag
    assert np.array_equal(buf.done_[:n].cpu().numpy(), g[f"{tag}_done"]), tag


@pytest.mark.parametrize("tag,ut,lam", [("gen_none", None, None), ("gen_max", "max_of_both", 0.7),
                                        ("gen_dis", "disagreement", 1.5)])
def test_generated_vs_reference_buffer(tag, ut, lam):
    g = load_golden("slac_buffer")
    buf = _buf()
    buf.load_generated(make_dataset(int(g["seed_gen"]), list(g["lengths"]), 0), ut, lam)
    _check_against(buf, g, tag)


@pytest.mark.parametrize("tag,size", [("plain", 1000), ("plain_ring", 17)])
def test_plain_vs_reference_buffer(tag, size):
    g = load_golden("slac_buffer")
    buf = _buf(size)
    buf.load_plain(make_dataset(int(g["seed_plain"]), list(g["lengths"]), int(g["tail_plain"])))
    _check_against(buf, g, tag)


def test_sampling_vs_reference_buffer():
    g = load_golden("slac_buffer")
    buf = _buf()
    buf.load_generated(make_dataset(int(g["seed_gen"]), list(g["lengths"]), 0), "disagreement", 1.5)
    np.random.seed(int(g["sample_seed"]))
    for tag, fn, b in (("lat", buf.sample_latent, 5), ("sac", buf.sample_sac, 4)):
        st, ac, rw, dn = fn(b)
        assert st.dtype == torch.float32 and np.array_equal(st.cpu().numpy(), g[f"{tag}_state"])
        assert np.array_equal(ac.cpu().numpy(), g[f"{tag}_action"])
        assert np.array_equal(rw.cpu().numpy(), g[f"{tag}_reward"])
        assert np.array_equal(dn.cpu().numpy(), g[f"{tag}_done"])


def test_full_size_frames_real_then_generated_vs_oracle():
    """100x100x3 frames; a real dataset loaded first, a generated one appended (one buffer, algo.py:407-412)."""
    real = make_dataset(31, [30, 14], 9, h=100, w=100)
    gen = make_dataset(32, [20, 11], 0, h=100, w=100)
    ora = SO.ReplayBuffer(500, 8, (C, 100, 100), (A,))
    SO.fill_plain(ora, loaded(real))
    SO.fill_generated(ora, loaded(gen), "average_both", 0.3)
    buf = _buf(500, 100, 100)
    buf.load_plain(real)
    buf.load_generated(gen, "average_both", 0.3)
    assert buf._n == ora._n and buf._p == ora._p and buf._real_n == 23 + 7 + 2
    idx = np.random.RandomState(0).randint(0, ora._n, size=16)
    assert np.array_equal(buf.gather_states(idx).cpu().numpy(), ora.gather_states(idx))
    n = ora._n
    assert np.array_equal(buf.action_[:n].cpu().numpy(), ora.action_[:n])
    assert np.array_equal(buf.reward_[:n].cpu().numpy(), ora.reward_[:n])
    assert float(buf.done_[:n].abs().sum()) == 0.0


def test_odd_frame_size_unaligned_path():
    """H*W*C not a multiple of 16 and H*W not a multiple of 4: scalar load / store paths."""
    d = make_dataset(41, [15], 0, h=7, w=9)
    ora = SO.ReplayBuffer(100, 8, (C, 7, 9), (A,))
    SO.fill_plain(ora, loaded(d))
    buf = _buf(100, 7, 9)
    buf.load_plain(d)
    idx = np.arange(ora._n)
    assert np.array_equal(buf.gather_states(idx).cpu().numpy(), ora.gather_states(idx))


def test_rejects_cpu_and_bad_layout():
    from s2p_b200.slac_buffer import SlacReplayBufferB200
    with pytest.raises(RuntimeError):
        SlacReplayBufferB200(10, 8, (3, 12, 10), (6,), "cpu")
    buf = _buf()
    d = make_dataset(1, [12], 0)
    d["image_observations"] = SO.to_chw(d["image_observations"]).copy()   # CHW instead of the disk layout
    with pytest.raises(RuntimeError):
        buf.load_plain(d)


def test_multistep_generated_mode_vs_reference_buffer():
    """random_state_5step_random_action through load_data (the consumer's branch selection, algo.py:355-371): real
    data first, the generated trajectories appended: bit-exact against what the reference's ReplayBuffer held"""
    from slac_dataset import make_multistep
    g = load_golden("slac_buffer")
    real, ms = make_multistep(int(g["seed_multistep"]), list(g["lengths"]), int(g["tail_plain"]), int(g["multistep_starts"]), 5)
    buf = _buf()
    buf.load_data({k: v for k, v in real.items()}, generated_for_slac=False)
    n_real = buf._n
    buf.load_data({k: v for k, v in ms.items() if not k.startswith("_")}, generated_for_slac=True,
                  data_mix_type="random_state_5step_random_action")
    assert buf._real_n == n_real and buf._n - n_real == int(g["multistep_starts"])
    _check_against(buf, g, "multistep")
    with pytest.raises(NotImplementedError):
        buf.load_data(ms, generated_for_slac=True, data_mix_type="no_such_mode")
