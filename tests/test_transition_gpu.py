"""GPU: ensemble_transition_kernel / slac_indices_kernel through the C ABI against
(1) vectors produced by the reference's own gaussian_ensemble.py (tests/golden/transition_cheetah.npz)
and (2) the numpy oracle on seeded inputs (ragged tiles, walker dims, member subsets)."""
import numpy as np
import pytest
import torch

import transition_oracle as TO
from common import load_golden

pytestmark = pytest.mark.gpu


def _model(sd, obs_dim=17, act_dim=6, hidden_layers=3):
    from s2p_b200.transition import EnsembleTransitionB200
    m = EnsembleTransitionB200(obs_dim, act_dim, 256, hidden_layers)
    full = dict(m.state_dict())
    for k, v in sd.items():
        full[k] = torch.from_numpy(v)
    for name in TO.layer_names(hidden_layers):
        full[f"{name}.saved_weight"] = full[f"{name}.weight"].clone()
        full[f"{name}.saved_bias"] = full[f"{name}.bias"].clone()
    m.load_state_dict(full)
    return m.cuda()


def _golden():
    g = load_golden("transition_cheetah")
    norm = {k[5:]: g[k] for k in g.files if k.startswith("norm_")}
    sd = TO.make_weights(17, 6, seed=int(g["weight_seed"]))
    return g, norm, sd


def test_forward_vs_reference_vectors():
    g, norm, sd = _golden()
    m = _model(sd)
    xin = np.concatenate([(g["observations"] - norm["obs_mean"]) / norm["obs_std"], g["fwd_actions"]], -1)
    dist = m(torch.from_numpy(xin.astype(np.float32)).cuda())
    assert isinstance(dist, torch.distributions.Normal)
    np.testing.assert_allclose(dist.mean.cpu().numpy(), g["fwd_mean"], rtol=0, atol=1e-5)
    np.testing.assert_allclose(dist.stddev.cpu().numpy(), g["fwd_std"], rtol=5e-6, atol=1e-6)


def test_generate_rollout_vs_reference_vectors():
    from s2p_b200.transition import generate_rollout
    g, norm, sd = _golden()
    m = _model(sd)
    out = generate_rollout(m, g["observations"], g["timeouts"], norm, g["act_low"], g["act_high"],
                           rng=np.random.RandomState(int(g["np_seed"])))
    for k in ("actions", "slac_action_indices", "slac_observation_indices"):
        assert out[k].dtype == g["roll_" + k].dtype and np.array_equal(out[k], g["roll_" + k]), k
    for k in ("rewards", "next_observations", "disagreement_uncertainty", "aleatoric_uncertainty"):
        assert out[k].shape == g["roll_" + k].shape, k
        np.testing.assert_allclose(out[k], g["roll_" + k], rtol=5e-6, atol=2e-5, err_msg=k)


@pytest.mark.parametrize("T", [1, 63, 64, 65, 1000])
def test_ragged_tiles_vs_oracle(T):
    _, norm, sd = _golden()
    m = _model(sd)
    rs = np.random.RandomState(T)
    x = rs.randn(T, 23).astype(np.float32)
    mu, std = TO.ensemble_forward(sd, x)
    dist = m(torch.from_numpy(x).cuda())
    np.testing.assert_allclose(dist.mean.cpu().numpy(), mu, atol=1e-5)
    np.testing.assert_allclose(dist.stddev.cpu().numpy(), std, rtol=5e-6, atol=1e-6)


def test_walker_dims_two_hidden_layers_member_subset():
    sd = TO.make_weights(24, 6, hidden_layers=2, seed=3)
    m = _model(sd, 24, 6, 2)
    m.set_select([6, 0, 3])
    x = np.random.RandomState(5).randn(200, 30).astype(np.float32)
    mu, std = TO.ensemble_forward(sd, x, obs_dim=24, hidden_layers=2, select=[6, 0, 3])
    dist = m(torch.from_numpy(x).cuda())
    assert dist.mean.shape == (3, 200, 25)
    np.testing.assert_allclose(dist.mean.cpu().numpy(), mu, atol=1e-5)
    np.testing.assert_allclose(dist.stddev.cpu().numpy(), std, rtol=5e-6, atol=1e-6)


def test_rollout_rows_large_and_bit_stable():
    """100k rows: selection / uncertainty against the oracle on a sample, run-to-run identical,
    and independent of how the rows are split into launches (rows are independent)."""
    _, norm, sd = _golden()
    m = _model(sd)
    T = 100_000
    rs = np.random.RandomState(9)
    obs = (norm["obs_mean"] + norm["obs_std"] * rs.randn(T, 17)).astype(np.float32)
    act = rs.uniform(-1, 1, size=(T, 6)).astype(np.float32)
    idx = rs.randint(0, 7, size=T).astype(np.int32)
    o, a, i = (torch.from_numpy(v).cuda() for v in (obs, act, idx))
    r1 = m.rollout_rows(o, a, i, norm)
    r2 = m.rollout_rows(o, a, i, norm)
    for x, y in zip(r1, r2):
        assert torch.equal(x, y)
    lo = 12_345
    r3 = m.rollout_rows(o[lo:lo + 777], a[lo:lo + 777], i[lo:lo + 777], norm)
    for x, y in zip(r1, r3):
        assert torch.equal(x[lo:lo + 777], y)
    sl = slice(50_000, 50_300)
    xin = np.concatenate([(obs[sl] - norm["obs_mean"]) / norm["obs_std"], act[sl]], -1).astype(np.float32)
    mu, std = TO.ensemble_forward(sd, xin)
    nxt, rew, dis, ale = TO.select_and_uncertainty(mu, std, idx[sl], norm)
    for got, want in zip(r1, (nxt, rew, dis, ale)):
        np.testing.assert_allclose(got[sl].cpu().numpy(), want, rtol=5e-6, atol=2e-5)


def test_empty_batch():
    _, _, sd = _golden()
    dist = _model(sd)(torch.zeros(0, 23, device="cuda"))
    assert dist.mean.shape == (7, 0, 18) and dist.stddev.shape == (7, 0, 18)


def test_no_cpu_path():
    _, _, sd = _golden()
    from s2p_b200.transition import EnsembleTransitionB200
    m = EnsembleTransitionB200(17, 6, 256, 3)
    with pytest.raises(RuntimeError):
        m(torch.zeros(4, 23))
    with pytest.raises(RuntimeError):
        EnsembleTransitionB200(17, 6, 128, 3)
