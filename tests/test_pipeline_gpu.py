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


def _transition_model():
    from s2p_b200.transition import EnsembleTransitionB200
    sd = TO.make_weights(17, 6, seed=7)
    tm = EnsembleTransitionB200(17, 6, 256, 3)
    full = dict(tm.state_dict())
    full.update({k: torch.from_numpy(v) for k, v in sd.items()})
    for name in TO.layer_names():
        full[f"{name}.saved_weight"], full[f"{name}.saved_bias"] = full[f"{name}.weight"].clone(), full[f"{name}.bias"].clone()
    tm.load_state_dict(full)
    return sd, tm.cuda()


@pytest.mark.parametrize("mode", ["random", "offRL"])
def test_multistep_dataset_pipeline_vs_oracles(mode):
    """random_state_5step_{random,offRL}_action: 5 chained ensemble steps -> 5-step generator rollout -> dataset rows
    -> SLAC loader, checked stage by stage (SURVEY.md section 8 a6 / f, VERDICT r01 missing item 1)"""
    import s2p_b200
    import s2p_oracle as O
    from s2p_b200.slac_buffer import SlacReplayBufferB200
    from s2p_b200.synthesis import synthesize_multistep_dataset
    g = common.load_golden("transition_cheetah")
    norm = {k[5:]: g[k] for k in g.files if k.startswith("norm_")}
    d = make_dataset(71, [14, 11], 4, h=100, w=100)
    rs = np.random.RandomState(72)
    n_all = d["actions"].shape[0]
    d["observations"] = (norm["obs_mean"] + norm["obs_std"] * rs.randn(n_all, 17)).astype(np.float32)
    d["next_observations"] = np.roll(d["observations"], -1, axis=0)
    sd, tm = _transition_model()
    ora_g = common.build_oracle("cheetah", "scaled")
    netG = s2p_b200.define_G(s2p_b200.make_options(gpu_ids="0"))
    netG.load_state_dict(ora_g.state_dict())
    netG.eval()
    act_low, act_high = -np.ones(6, np.float32), np.ones(6, np.float32)
    pol_w = torch.from_numpy(np.random.RandomState(3).randn(17, 6).astype(np.float32) * 0.1)
    policy = None if mode == "random" else (lambda s: torch.tanh(s @ pol_w.to(s.device)))
    J, steps, hist = 6, 5, 3
    ms = synthesize_multistep_dataset(netG, tm, d, norm, act_low, act_high, n_starts=J, steps=steps, policy=policy,
                                      rng=np.random.RandomState(9), batch=4)
    assert str(ms["data_mix_type"]) == f"random_state_5step_{mode}_action"
    assert ms["image_observations_tp1"].shape == (J * 8 + 1, 100, 100, 3) and ms["timeouts"].sum() == J
    full_ms = ms
    ms = {k: (v[:-1] if getattr(v, "ndim", 0) >= 1 else v) for k, v in ms.items()}       # drop the sentinel row for the per-step checks
    # stage 1: the chained transition steps against the numpy ensemble oracle with the same host draws
    rng = np.random.RandomState(9)
    from s2p_b200.synthesis import multistep_start_rows
    starts = multistep_start_rows(d["timeouts"], J, hist, rng)
    assert np.array_equal(ms["source_row"].reshape(J, 8)[:, 0], starts)
    s = d["observations"][starts]
    rows = ms["observations"].reshape(J, 8, 17)
    for k in range(steps):
        a = rng.uniform(low=act_low, high=act_high, size=(J, 6)).astype(np.float32) if policy is None else \
            np.tanh(s @ pol_w.numpy()).astype(np.float32)
        member = rng.randint(0, 7, size=J)
        np.testing.assert_allclose(rows[:, hist + k], s, rtol=1e-5, atol=1e-5)
        got_a = ms["actions"].reshape(J, 8, 6)[:, hist + k]
        if policy is None:
            assert np.array_equal(got_a, a)
        else:
            np.testing.assert_allclose(got_a, a, rtol=1e-5, atol=1e-6)
        x = np.concatenate([(s - norm["obs_mean"]) / norm["obs_std"], got_a], axis=1).astype(np.float32)
        mu, std = TO.ensemble_forward(sd, x)
        sel = mu[member, np.arange(J)]
        nxt = sel[:, :17] * norm["next_obs_std"] + norm["next_obs_mean"]
        np.testing.assert_allclose(ms["next_observations"].reshape(J, 8, 17)[:, hist + k], nxt, rtol=2e-5, atol=5e-5)
        np.testing.assert_allclose(ms["rewards"].reshape(J, 8)[:, hist + k], sel[:, 17] * norm["reward_std"] + norm["reward_mean"],
                                   rtol=2e-5, atol=5e-5)
        s = ms["next_observations"].reshape(J, 8, 17)[:, hist + k]      # chain on the product's own states
    # real prefix rows are the dataset's rows i-3 .. i-1
    assert np.array_equal(ms["image_observations"].reshape(J, 8, 100, 100, 3)[:, :hist],
                          np.stack([d["image_observations"][i - hist:i] for i in starts]))
    # stage 2: the generated frames, teacher-forced per step against the fp32 oracle generator (PSNR >= 50 dB on the
    # fp32 frames is asserted in test_generator_gpu; here the stored uint8 frames: <= a few LSB)
    fr = ms["image_observations"].reshape(J, 8, 100, 100, 3)
    tp1 = ms["image_observations_tp1"].reshape(J, 8, 100, 100, 3)
    assert np.array_equal(fr[:, hist], d["image_observations"][starts])          # frame_0 = the start row's frame
    assert np.array_equal(fr[:, hist + 1:], tp1[:, hist:-1])                      # frame_k of a row = frame_{k} of the previous row's tp1
    st_t = torch.from_numpy(ms["next_observations"].reshape(J, 8, 17)[:, hist:])
    with torch.no_grad():
        want1 = O.to_uint8_hwc(ora_g(O.from_uint8_hwc(torch.from_numpy(fr[:, hist])), st_t[:, 0]))
    diff = np.abs(tp1[:, hist].astype(np.int32) - want1.numpy().astype(np.int32))
    assert diff.max() <= 12 and diff.mean() <= 1.0
    # stage 3: the consumer: real data, then the generated trajectories through the plain branch -- bit-exact vs the oracle loops
    real = dict(d)
    real["image_observations_tp1"] = np.roll(d["image_observations"], -1, axis=0)
    buf = SlacReplayBufferB200(200, 8, (3, 100, 100), (6,), "cuda")
    buf.load_data(real, generated_for_slac=False)
    ms = full_ms
    buf.load_data(ms, generated_for_slac=True, data_mix_type=str(ms["data_mix_type"]))
    ob = SO.ReplayBuffer(200, 8, (3, 100, 100), (6,))
    to_loaded = lambda dd: dict(dd, image_observations=SO.to_chw(dd["image_observations"]),  # noqa: E731
                                image_observations_tp1=SO.to_chw(dd["image_observations_tp1"]))
    SO.fill_plain(ob, to_loaded(real))
    SO.fill_plain(ob, to_loaded(ms), reset_buff=True)
    assert buf._n == ob._n and buf._n - buf._real_n == J
    idx = np.arange(ob._n)
    assert np.array_equal(buf.gather_states(idx).cpu().numpy(), ob.gather_states(idx))
    assert np.array_equal(buf.action_[:ob._n].cpu().numpy(), ob.action_[:ob._n])
    assert np.array_equal(buf.reward_[:ob._n].cpu().numpy(), ob.reward_[:ob._n])
