"""CPU: the numpy restatement of the RAD augmentations (oracle/data_augs_oracle.py) against the outputs of the
REFERENCE'S OWN functions (tests/golden/data_augs.npz, made by tests/golden/make_data_augs_golden.py from
/root/reference/examples/iql/data_augs.py).  Bit-exact: index moves and the grayscale's uint8 roundings."""
import numpy as np
import pytest

import data_augs_oracle as DO
from common import load_golden


def golden_inputs(g):
    n, c, h, w = (int(v) for v in g["shape"])
    rng = np.random.RandomState(int(g["seed_img"]))
    u8 = rng.randint(0, 256, size=(n, c, h, w)).astype(np.uint8)
    f32 = (rng.randint(0, 256, size=(n, c, h, w)) / 255.).astype(np.float32)
    return u8, f32


CASES = [("crop", lambda x: DO.random_crop(x, out=32)), ("cutout", lambda x: DO.random_cutout(x, 10, 30)),
         ("cutout_color", lambda x: DO.random_cutout_color(x, 10, 30)), ("translate", lambda x: DO.random_translate(x, 52))]


@pytest.mark.parametrize("name,fn", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("tag", ["u8", "f32"])
def test_index_augs_match_reference(name, fn, tag):
    g = load_golden("data_augs")
    u8, f32 = golden_inputs(g)
    np.random.seed(int(g["seed_draw"]))
    out = fn(u8 if tag == "u8" else f32)
    ref = g[f"{name}_{tag}"]
    assert out.dtype == ref.dtype and out.shape == ref.shape
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("name,fn", [("flip", lambda x: DO.random_flip(x, p=.5)), ("rotation", lambda x: DO.random_rotation(x, p=.7)),
                                     ("grayscale", lambda x: DO.random_grayscale(x, p=.5))], ids=["flip", "rotation", "grayscale"])
def test_tensor_augs_match_reference(name, fn):
    g = load_golden("data_augs")
    _, f32 = golden_inputs(g)
    np.random.seed(int(g["seed_draw"]))
    out = fn(f32)
    assert np.array_equal(out, g[f"{name}_f32"])
    assert not np.array_equal(out, f32)   # the draw changed at least one image


def test_translate_returns_and_reuses_draws():
    u8 = np.arange(2 * 3 * 5 * 5, dtype=np.uint8).reshape(2, 3, 5, 5)
    np.random.seed(3)
    out, idx = DO.random_translate(u8, 9, return_random_idxs=True)
    again = DO.random_translate(u8, 9, **idx)
    assert np.array_equal(out, again)
    assert np.array_equal(DO.crop_at(out, idx["h1s"], idx["w1s"], 5), u8)   # crop undoes translate


def test_grayscale_helper_matches_reference():
    g = load_golden("data_augs")
    u8, f32 = golden_inputs(g)
    assert np.array_equal(DO.grayscale(u8), g["gray_plain_u8"])
    assert np.array_equal(DO.grayscale((f32 * np.float32(255.0))), g["gray_plain_f32"])
    assert g["gray_plain_u8"].shape == (u8.shape[0], u8.shape[1] // 3, 3, u8.shape[2], u8.shape[3])


def test_random_convolution_matches_reference():
    """the reference's own output for the filters IT drew (torch.manual_seed, kept in the fixture)"""
    g = load_golden("data_augs")
    _, f32 = golden_inputs(g)
    w = g["randconv_weights"]
    assert w.shape == (f32.shape[0], 3, 3, 3, 3)
    out = DO.random_convolution_at(f32, w)
    ref = g["randconv_f32"]
    assert out.shape == ref.shape and np.abs(out - ref).max() <= 2e-6 * np.abs(ref).max()
