"""CPU: the frozen oracle reproduces its committed golden vectors (spec-drift guard).
The reference holds no golden vectors for this path: these pin
ORACLE_SPEC v1 itself -- parity with the reference stays 'unpinned'."""
import pytest
import torch

import common
import s2p_oracle as O


@pytest.mark.parametrize("name", ["cheetah_xavier_b2", "cheetah_scaled_b2", "walker_scaled_b2", "cheetah_trained_like_b2"])
def test_oracle_matches_golden_one_step(name):
    case = common.manifest()["cases"][name]
    net = common.build_oracle(case["env"], case["init_type"])
    wsum = float(sum(p.detach().double().abs().sum() for p in net.parameters()))
    assert abs(wsum - case["weight_abs_sum"]) <= 1e-6 * case["weight_abs_sum"], "seeded init drifted"
    image, states = common.make_inputs(case["env"], case["batch"], case["steps"])
    with torch.no_grad():
        out = net(image, states[:, 0])
    gold = torch.from_numpy(common.load_golden(name)["first"])
    # same machine class and torch build -> tiny thread-order noise only
    assert (out - gold).abs().max().item() <= 2e-4
    u8 = O.to_uint8_hwc(out)
    assert u8.shape == (case["batch"], 100, 100, 3) and u8.dtype == torch.uint8
    gold_u8 = torch.from_numpy(common.load_golden(name)["u8_last"]).int()
    assert (u8.int() - gold_u8).abs().max().item() <= 1


def test_trained_like_init_is_non_degenerate():
    """ORACLE_SPEC v1.3 init: tanh unsaturated, output O(0.3), every blend factor away from 0.5"""
    case = common.manifest()["cases"]["cheetah_trained_like_b2"]
    assert case["saturated_frac"] == 0.0 and 0.2 <= case["out_std"] <= 0.5 and case["out_range"] >= 1.0
    net = common.build_oracle("cheetah", "trained_like")
    blends = [torch.sigmoid(p).item() for n, p in net.named_parameters() if "blend_" in n]
    assert len(blends) == 26 and all(abs(b - 0.5) >= 0.2 for b in blends)


@pytest.mark.parametrize("name", ["cheetah_scaled_roll5_b1", "walker_trained_like_roll5_b1"])
def test_oracle_rollout_golden(name):
    case = common.manifest()["cases"][name]
    net = common.build_oracle(case["env"], case["init_type"])
    image, states = common.make_inputs(case["env"], case["batch"], case["steps"])
    out = net.rollout(image, states)
    assert out.shape == (1, 5, 3, 100, 100)
    gold = common.load_golden(name)
    assert (out[:, 0] - torch.from_numpy(gold["first"])).abs().max().item() <= 2e-4
    assert common.psnr01(out[:, -1], torch.from_numpy(gold["last"])) >= 60.0


def test_spec_arithmetic():
    """56.44 GMAC per frame for ORACLE_SPEC v1."""
    m = O.flops_per_frame()
    assert abs(m["total"] / 1e9 - 56.44) < 0.01
    assert abs(m["modulation"] / 1e9 - 38.71) < 0.01
    assert abs(m["trunk"] / 1e9 - 16.59) < 0.01


def test_positional_encoding_layout():
    s = torch.tensor([[0.5, -2.0]])
    pe = O.positional_encoding(s, 2)
    exp = torch.cat([s, torch.sin(s), torch.cos(s), torch.sin(2 * s), torch.cos(2 * s)], dim=-1)
    assert torch.equal(pe, exp) and pe.shape == (1, 10)


def test_u8_roundtrip():
    u = torch.randint(0, 256, (2, 100, 100, 3), dtype=torch.uint8)
    assert torch.equal(O.to_uint8_hwc(O.from_uint8_hwc(u)), u)
