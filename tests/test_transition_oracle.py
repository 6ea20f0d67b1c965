"""CPU: the numpy restatement of the state-transition producer (oracle/transition_oracle.py) is held
to vectors produced by the REFERENCE'S OWN code (tests/golden/make_transition_golden.py imports
/root/reference/gaussian_ensemble.py) -- parity pinned for this component."""
import numpy as np
import pytest

import transition_oracle as TO
from common import load_golden


def _golden():
    g = load_golden("transition_cheetah")
    norm = {k[5:]: g[k] for k in g.files if k.startswith("norm_")}
    sd = TO.make_weights(17, 6, seed=int(g["weight_seed"]))
    return g, norm, sd


def test_forward_matches_reference_vectors():
    g, norm, sd = _golden()
    xin = np.concatenate([(g["observations"] - norm["obs_mean"]) / norm["obs_std"], g["fwd_actions"]], -1)
    mu, std = TO.ensemble_forward(sd, xin.astype(np.float32))
    assert mu.shape == g["fwd_mean"].shape == (7, 304, 18)
    np.testing.assert_allclose(mu, g["fwd_mean"], rtol=0, atol=5e-6)
    np.testing.assert_allclose(std, g["fwd_std"], rtol=2e-6, atol=1e-6)
    # both sides of soft_clamp are exercised by the fixture
    assert g["fwd_std"].min() < 0.01 and g["fwd_std"].max() > 3.0


def test_rollout_matches_reference_vectors():
    g, norm, sd = _golden()
    rng = np.random.RandomState(int(g["np_seed"]))
    out = TO.rollout_1step_random_action(sd, g["observations"], g["timeouts"], norm, g["act_low"], g["act_high"], rng)
    assert out["actions"].shape[0] == 299  # the 5 rows after the last timeout are dropped
    for k in ("actions", "slac_action_indices", "slac_observation_indices"):
        assert np.array_equal(out[k], g["roll_" + k]), k
    for k in ("rewards", "next_observations", "disagreement_uncertainty", "aleatoric_uncertainty"):
        assert out[k].shape == g["roll_" + k].shape
        np.testing.assert_allclose(out[k], g["roll_" + k], rtol=2e-6, atol=1e-5, err_msg=k)


def test_select_subset_is_member_slice():
    _, norm, sd = _golden()
    x = np.random.RandomState(0).randn(10, 23).astype(np.float32)
    mu, std = TO.ensemble_forward(sd, x)
    mu2, std2 = TO.ensemble_forward(sd, x, select=[5, 1])
    np.testing.assert_allclose(mu2, mu[[5, 1]], atol=2e-6)
    np.testing.assert_allclose(std2, std[[5, 1]], rtol=2e-6)


def test_slac_indices_shape_and_sentinel():
    obs, act = TO.obs_act_indices(12, 100)
    assert obs.shape == (12, 9) and act.shape == (12, 8)
    assert (obs[:8] == TO.INTEGER_INF).all() and (act[:8] == TO.INTEGER_INF).all()
    assert obs[8].tolist() == list(range(100, 109)) and act[11].tolist() == list(range(103, 111))
