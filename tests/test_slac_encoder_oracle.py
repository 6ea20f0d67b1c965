"""CPU: oracle/slac_encoder_oracle.py against outputs of the reference's own Encoder (tests/golden/slac_encoder.npz)."""
import numpy as np
import pytest

import slac_encoder_oracle as EO
from common import load_golden


@pytest.mark.parametrize("size", [100, 64])
def test_encoder_oracle_matches_reference_vectors(size):
    g = load_golden("slac_encoder")
    b, s = (int(v) for v in g[f"shape_{size}"])
    frames = EO.make_frames(size + 1, b, s, size)
    x = np.transpose(frames, (0, 1, 4, 2, 3)).astype(np.float32) / np.float32(255.0)
    y = EO.encoder_forward(EO.make_weights(size, seed=size), x, size)
    assert y.shape == g[f"out_{size}"].shape == (b, s, 256)
    np.testing.assert_allclose(y, g[f"out_{size}"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("size", [100, 64])
def test_decoder_oracle_matches_reference_vectors(size):
    g = load_golden("slac_encoder")
    b, s = (int(v) for v in g[f"dec_shape_{size}"])
    z = np.random.RandomState(300 + size).randn(b, s, 288).astype(np.float32)
    mean, std = EO.decoder_forward(EO.make_decoder_weights(size, seed=200 + size), z, size)
    assert mean.shape == g[f"dec_mean_{size}"].shape == (b, s, 3, size, size) and float(std.min()) == 1.0
    np.testing.assert_allclose(mean, g[f"dec_mean_{size}"], rtol=1e-5, atol=1e-6)
