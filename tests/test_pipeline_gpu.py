"""GPU: the chain either side of the generator, end to end through the product path and checked stage by stage
against the oracles: state-transition rollout -> S2P generator -> augmented dataset -> SLAC window buffer ->
SLAC encoder (README.md:54 "generation code" + rlkit/torch/slac/algo.py loader)."""
import numpy as np
import pytest
import torch

import common
import slac_buffer_oracle as SO
import slac_encoder_oracle as EO
import transition_oracle as TO
from slac_dataset import make_dataset

pytestmark = pytest.mark.gpu


def test_augmented_dataset_pipeline_vs_oracles():
    import s2p_b200
    import s2p_oracle as O
    from s2p_b200.networks.slac_encoder import SlacEncoderB200
    from s2p_b200.slac_buffer import SlacReplayBufferB200
    from s2p_b200.synthesis import synthesize_augmented_dataset
    from s2p_b200.transition import EnsembleTransitionB200
    g = common.load_golden("transition_cheetah")
    norm = {k[5:]: g[k] for k in g.files if k.startswith("norm_")}
    # real dataset: two trajectories of 12 and 10 rows, 100x100x3 frames, cheetah states
    d = make_dataset(61, [12, 10], 3, h=100, w=100)
    rs = np.random.RandomState(62)
    n_all = d["actions"].shape[0]
    d["observations"] = (norm["obs_mean"] + norm["obs_std"] * rs.randn(n_all, 17)).astype(np.float32)
    for k in ("original_actions", "original_rewards", "slac_observation_indices", "slac_action_indices",
              "aleatoric_uncertainty", "disagreement_uncertainty", "image_observations_tp1"):
        d.pop(k)
    sd = TO.make_weights(17, 6, seed=7)
    tm = EnsembleTransitionB200(17, 6, 256, 3)
    full = dict(tm.state_dict())
    full.update({k: torch.from_numpy(v) for k, v in sd.items()})
    for name in TO.layer_names():
        full[f"{name}.saved_weight"], full[f"{name}.saved_bias"] = full[f"{name}.weight"].clone(), full[f"{name}.bias"].clone()
    tm.load_state_dict(full)
    tm = tm.cuda()
    ora_g = common.build_oracle("cheetah", "scaled")
    netG = s2p_b200.define_G(s2p_b200.make_options(gpu_ids="0"))
    netG.load_state_dict(ora_g.state_dict())
    netG.eval()
    act_low, act_high = -np.ones(6, np.float32), np.ones(6, np.float32)
    aug = synthesize_augmented_dataset(netG, tm, d, norm, act_low, act_high, rng=np.random.RandomState(5), batch=16)
    n = 22
    assert aug["image_observations_tp1"].shape == (n, 100, 100, 3) and aug["image_observations_tp1"].dtype == np.uint8
    # stage 1: rollout keys against the numpy oracle (same RNG stream)
    ref = TO.rollout_1step_random_action(sd, d["observations"], d["timeouts"], norm, act_low, act_high, np.random.RandomState(5))
    for k, v in ref.items():
        if v.dtype.kind == "i" or k == "actions":
            assert np.array_equal(aug[k], v), k
        else:
            np.testing.assert_allclose(aug[k], v, rtol=5e-6, atol=2e-5, err_msg=k)
    assert np.array_equal(aug["original_actions"], d["actions"][:n])
    # stage 2: generated frames against the fp32 oracle generator on the same rows (uint8 tolerance of the bf16 path)
    with torch.no_grad():
        want = O.to_uint8_hwc(ora_g(O.from_uint8_hwc(torch.from_numpy(d["image_observations"][:n])),
                                    torch.from_numpy(ref["next_observations"])))
    diff = (aug["image_observations_tp1"].astype(np.int32) - want.numpy().astype(np.int32))
    assert np.abs(diff).max() <= 12 and np.abs(diff).mean() <= 1.0
    # stage 3: the SLAC loader + sampler on the generated dataset, bit-exact against the reference-pinned oracle
    buf = SlacReplayBufferB200(100, 8, (3, 100, 100), (6,), "cuda")
    buf.load_generated(aug, "disagreement", 0.5)
    e = dict(aug)
    e["image_observations"], e["image_observations_tp1"] = SO.to_chw(aug["image_observations"]), SO.to_chw(aug["image_observations_tp1"])
    e["disagreement_uncertainty"] = aug["disagreement_uncertainty"].squeeze(-1)
    ob = SO.ReplayBuffer(100, 8, (3, 100, 100), (6,))
    SO.fill_generated(ob, e, "disagreement", 0.5)
    assert buf._n == ob._n == (12 - 8) + (10 - 8) - 1
    idx = np.arange(ob._n)
    x = buf.gather_states(idx)
    assert np.array_equal(x.cpu().numpy(), ob.gather_states(idx))
    assert np.array_equal(buf.reward_[:ob._n].cpu().numpy(), ob.reward_[:ob._n])
    # stage 4: encoder on the sampled windows, fused uint8 path
    enc = SlacEncoderB200(3, 256, 100)
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in EO.make_weights(100, seed=100).items()})
    enc = enc.cuda()
    y = enc.forward_from_buffer(buf, idx).cpu().numpy()
    yr = EO.encoder_forward(EO.make_weights(100, seed=100), x.cpu().numpy(), 100)
    assert np.abs(y - yr).max() <= 0.02 * np.abs(yr).max()
