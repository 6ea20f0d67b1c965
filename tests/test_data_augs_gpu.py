"""GPU: s2p_b200.data_augs (rad augment_kernel / grayscale_kernel through the C ABI) against
(1) the outputs of the REFERENCE'S OWN functions on the same np.random seed (tests/golden/data_augs.npz),
(2) the numpy oracle at the repo's frame shape (9 x 100 x 100 stacks, uint8 and float32) and on ragged sizes that take
    the scalar-store path, and (3) size-independent properties at a large batch (crop undoes translate, flip and a
    full turn are involutions).  Index moves and the grayscale's uint8 roundings are bit-exact: array_equal."""
import numpy as np
import pytest
import torch

import data_augs_oracle as DO
from common import load_golden
from test_data_augs_oracle import golden_inputs

pytestmark = pytest.mark.gpu


def rad():
    import s2p_b200.data_augs as R
    return R


def _np(t):
    return t.cpu().numpy() if isinstance(t, torch.Tensor) else t


GOLD = [("crop", "u8", lambda R, x: R.random_crop(x, out=32)), ("crop", "f32", lambda R, x: R.random_crop(x, out=32)),
        ("cutout", "u8", lambda R, x: R.random_cutout(x, 10, 30)), ("cutout", "f32", lambda R, x: R.random_cutout(x, 10, 30)),
        ("cutout_color", "u8", lambda R, x: R.random_cutout_color(x, 10, 30)),
        ("cutout_color", "f32", lambda R, x: R.random_cutout_color(x, 10, 30)),
        ("translate", "u8", lambda R, x: R.random_translate(x, 52)), ("translate", "f32", lambda R, x: R.random_translate(x, 52)),
        ("flip", "f32", lambda R, x: R.random_flip(x, p=.5)), ("rotation", "f32", lambda R, x: R.random_rotation(x, p=.7)),
        ("grayscale", "f32", lambda R, x: R.random_grayscale(x, p=.5))]


@pytest.mark.parametrize("name,tag,fn", GOLD, ids=[f"{g[0]}_{g[1]}" for g in GOLD])
@pytest.mark.parametrize("host", [False, True], ids=["cuda_tensor", "numpy_in_out"])
def test_matches_reference_outputs(name, tag, fn, host):
    g = load_golden("data_augs")
    u8, f32 = golden_inputs(g)
    x = u8 if tag == "u8" else f32
    np.random.seed(int(g["seed_draw"]))
    out = fn(rad(), x if host else torch.from_numpy(x).cuda())
    assert isinstance(out, np.ndarray) == host
    ref = g[f"{name}_{tag}"]
    out = _np(out)
    assert out.dtype == ref.dtype and out.shape == ref.shape
    assert np.array_equal(out, ref)


def _imgs(shape, dtype, seed):
    rng = np.random.RandomState(seed)
    if dtype == np.uint8:
        return rng.randint(0, 256, size=shape).astype(np.uint8)
    return rng.rand(*shape).astype(np.float32)


@pytest.mark.parametrize("dtype", [np.uint8, np.float32], ids=["u8", "f32"])
@pytest.mark.parametrize("shape,out,size", [((16, 9, 100, 100), 84, 108), ((5, 3, 37, 37), 30, 41), ((3, 1, 7, 7), 7, 7),
                                             ((4, 3, 37, 37), 32, 40), ((6, 2, 52, 52), 36, 64), ((3, 5, 100, 100), 84, 108),
                                             ((2, 2, 35, 36), 28, 44)],
                         ids=["frames100", "ragged37", "tiny7", "unaligned_rows37", "words52", "ragged_plane_groups",
                              "staged_skew_nonsquare"])
def test_every_op_vs_oracle(dtype, shape, out, size):
    R = rad()
    x = _imgs(shape, dtype, 5)
    xc = torch.from_numpy(x).cuda()
    lo, hi = (10, 30) if shape[2] >= 37 else (1, 6)
    calls = [("random_crop", dict(out=out)), ("random_cutout", dict(min_cut=lo, max_cut=hi)),
             ("random_cutout_color", dict(min_cut=lo, max_cut=hi)), ("random_translate", dict(size=size)),
             ("random_flip", dict(p=.5)), ("random_rotation", dict(p=.8))]
    if shape[2] != shape[3]:
        calls = calls[:-1]           # quarter turns need square frames (data_augs.py:164-186 asserts nothing, numpy would raise)
    for k, (name, kw) in enumerate(calls):
        np.random.seed(100 + k)
        got = getattr(R, name)(xc, **kw)
        np.random.seed(100 + k)
        want = getattr(DO, name)(x, **kw)
        assert got.dtype == xc.dtype and tuple(got.shape) == want.shape, name
        assert np.array_equal(_np(got), want), name
    assert np.array_equal(_np(xc), x)   # inputs untouched


def test_grayscale_vs_oracle_and_rounding_edges():
    R = rad()
    # every k/255 (trunc(x*255) sits on an integer edge for some k) and random values, 3 frames of 100 x 100
    rng = np.random.RandomState(9)
    x = rng.rand(8, 9, 100, 100).astype(np.float32)
    x[0] = (rng.randint(0, 256, size=(9, 100, 100)) / 255.).astype(np.float32)
    x[1, :, 0, :] = 1.0
    x[2, :, 0, :] = 0.0
    for p in (0.0, 0.5, 1.0):
        np.random.seed(4)
        got = R.random_grayscale(torch.from_numpy(x).cuda(), p=p)
        np.random.seed(4)
        want = DO.random_grayscale(x, p=p)
        bad = int((_np(got) != want).sum())
        assert bad == 0, (p, bad)
    xr = rng.rand(5, 6, 7, 9).astype(np.float32)     # H*W % 4 != 0: the scalar kernel
    np.random.seed(6)
    got = R.random_grayscale(torch.from_numpy(xr).cuda(), p=.6)
    np.random.seed(6)
    assert np.array_equal(_np(got), DO.random_grayscale(xr, p=.6))
    # and against torch's own device arithmetic for the reference's gray expression (no fused multiply-add); the last
    # division is done on the host: torch's CUDA `x / 255.` multiplies by the reciprocal, the reference's CPU path divides
    xc = torch.from_numpy(x).cuda()
    u8 = (xc * 255.).type(torch.uint8).view(8, 3, 3, 100, 100)
    gray = (u8[:, :, 0] * 0.2989 + u8[:, :, 1] * 0.587 + u8[:, :, 2] * 0.114).type(torch.uint8).float()
    np.random.seed(4)
    got = R.random_grayscale(xc, p=1.0)
    assert torch.equal(got.view(8, 3, 3, 100, 100)[:, :, 1].cpu(), gray.cpu() / 255.)


def test_properties_large_batch():
    R = rad()
    x = torch.randint(0, 256, (2048, 9, 100, 100), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    np.random.seed(8)
    t, idx = R.random_translate(x, 108, return_random_idxs=True)
    assert t.shape == (2048, 9, 108, 108)
    # crop undoes translate (same offsets): feed the translate draws to the crop kernel through the C ABI path
    from s2p_b200 import _lib
    back = R._augment(t, _lib.RAD_CROP, 100, 100, idx["h1s"], idx["w1s"])
    assert torch.equal(back, x)
    assert int(t.sum(dtype=torch.int64)) == int(x.sum(dtype=torch.int64))   # everything outside the paste is zero
    ones = np.ones(2048, dtype=np.int32)
    f = R._augment(x, _lib.RAD_FLIP, 100, 100, ones)
    assert not torch.equal(f, x) and torch.equal(R._augment(f, _lib.RAD_FLIP, 100, 100, ones), x)
    assert torch.equal(f, x.flip([3]))
    r = R._augment(x, _lib.RAD_ROT90, 100, 100, ones)
    assert torch.equal(r, x.rot90(1, [2, 3]))
    assert torch.equal(R._augment(r, _lib.RAD_ROT90, 100, 100, 3 * ones), x)


@pytest.mark.parametrize("dtype", [np.uint8, np.float32], ids=["u8", "f32"])
def test_more_images_than_one_grid(dtype):
    """N > 65535 (gridDim.y): the launch loop's second chunk starts at the right image, draws and fill included."""
    R = rad()
    from s2p_b200 import _lib
    n = 70000
    x = _imgs((n, 2, 8, 8), dtype, 3)
    xc = torch.from_numpy(x).cuda()
    rng = np.random.RandomState(1)
    h1, w1 = rng.randint(0, 5, n), rng.randint(0, 5, n)
    assert np.array_equal(_np(R._augment(xc, _lib.RAD_CROP, 4, 4, h1, w1)), DO.crop_at(x, h1, w1, 4))
    assert np.array_equal(_np(R._augment(xc, _lib.RAD_TRANSLATE, 12, 12, h1, w1)), DO.translate_at(x, h1, w1, 12))
    fill = _imgs((n, 2), dtype, 4)
    got = R._augment(xc, _lib.RAD_CUTOUT, 8, 8, h1, w1, torch.from_numpy(fill).cuda())
    assert np.array_equal(_np(got), DO.cutout_at(x, h1, w1, fill))
    turns = rng.randint(0, 4, n)
    assert np.array_equal(_np(R._augment(xc, _lib.RAD_ROT90, 8, 8, turns)), DO.rot90_at(x, turns))
    assert np.array_equal(_np(R._augment(xc, _lib.RAD_FLIP, 8, 8, turns & 1)), DO.flip_at(x, turns & 1))


def test_errors_and_empty():
    R = rad()
    from s2p_b200 import _lib
    with pytest.raises(RuntimeError):
        R.random_flip(torch.zeros(2, 3, 8, 8))                      # CPU tensor: no fallback
    with pytest.raises(RuntimeError):
        R.random_crop(torch.zeros(2, 3, 8, 8, dtype=torch.float64, device="cuda"), out=4)
    with pytest.raises(ValueError):
        R.random_rotation(torch.zeros(2, 3, 8, 6, device="cuda"))
    x = torch.zeros(2, 3, 8, 8, device="cuda")
    d = torch.zeros(2, dtype=torch.int32, device="cuda")
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    assert lib.s2p_rad_augment(x.data_ptr(), x.data_ptr(), 4, 2, 3, 8, 8, 8, 8, _lib.RAD_FLIP, d.data_ptr(), 0, 0, st) == -1  # aliasing
    y = torch.empty_like(x)
    assert lib.s2p_rad_augment(x.data_ptr(), y.data_ptr(), 2, 2, 3, 8, 8, 8, 8, _lib.RAD_FLIP, d.data_ptr(), 0, 0, st) == -3  # elem size
    assert lib.s2p_rad_augment(x.data_ptr(), y.data_ptr(), 4, 2, 3, 8, 8, 9, 9, _lib.RAD_CROP, d.data_ptr(), d.data_ptr(), 0, st) == -1
    assert lib.s2p_rad_augment(x.data_ptr(), y.data_ptr(), 4, 2, 3, 8, 8, 8, 8, 7, d.data_ptr(), 0, 0, st) == -3
    assert b"rad_augment" in lib.s2p_last_error_string()
    e = R.random_cutout(torch.zeros(0, 3, 8, 8, dtype=torch.uint8, device="cuda"))
    assert e.shape == (0, 3, 8, 8)


@pytest.mark.parametrize("tag", ["u8", "f32"])
def test_grayscale_helper_vs_reference_output(tag):
    g = load_golden("data_augs")
    u8, f32 = golden_inputs(g)
    x = torch.from_numpy(u8).cuda() if tag == "u8" else (torch.from_numpy(f32) * 255.).cuda()
    out = rad().grayscale(x)
    assert out.dtype == torch.float32 and np.array_equal(out.cpu().numpy(), g[f"gray_plain_{tag}"])
    big = torch.randint(0, 256, (64, 9, 100, 100), dtype=torch.uint8, device="cuda")
    assert np.array_equal(rad().grayscale(big).cpu().numpy(), DO.grayscale(big.cpu().numpy()))


def test_random_convolution_vs_reference_output_and_draw_order():
    """(1) the reference's own output for the filters it drew (fixture), <= 2e-6 relative (fp32 summation order);
    (2) at the repo's frame shape against the float64 oracle; (3) the filters are drawn with the reference's calls in
    the reference's order: nn.Conv2d construction (CPU generator), then one xavier_normal_ per image on the device."""
    R = rad()
    g = load_golden("data_augs")
    _, f32 = golden_inputs(g)
    out = R.random_convolution(torch.from_numpy(f32).cuda(), weights=g["randconv_weights"])
    ref = g["randconv_f32"]
    assert np.abs(out.cpu().numpy() - ref).max() <= 2e-6 * np.abs(ref).max()
    x = torch.rand(16, 9, 100, 100, device="cuda")
    torch.manual_seed(5)
    torch.cuda.manual_seed(6)
    y, w = R.random_convolution(x, return_weights=True)
    want = DO.random_convolution_at(x.cpu().numpy(), w.cpu().numpy())
    assert np.abs(y.cpu().numpy() - want).max() <= 2e-6 * np.abs(want).max()
    torch.manual_seed(5)
    torch.cuda.manual_seed(6)
    conv = torch.nn.Conv2d(3, 3, kernel_size=3, bias=False, padding=1).cuda()       # data_augs.py:208
    ws = []
    for _ in range(16):
        torch.nn.init.xavier_normal_(conv.weight.data)                              # :212
        ws.append(conv.weight.data.clone())
    assert torch.equal(w, torch.stack(ws))
    # ragged width (not a multiple of 4) and a single frame
    xs = torch.rand(3, 3, 7, 9, device="cuda")
    ys, wsmall = R.random_convolution(xs, return_weights=True)
    want = DO.random_convolution_at(xs.cpu().numpy(), wsmall.cpu().numpy())
    assert np.abs(ys.cpu().numpy() - want).max() <= 2e-6 * np.abs(want).max() + 1e-7
