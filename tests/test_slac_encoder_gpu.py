"""GPU: SlacEncoderB200 (CUDA-core conv1 + tcgen05 implicit-GEMM layers through the C ABI) against outputs of
the REFERENCE'S OWN Encoder (tests/golden/slac_encoder.npz) and the fp32 oracle.

Tolerance (bf16 activations and weights, fp32 accumulation, six layers): max-abs error <= 2 % of the output's
max-abs and relative RMS error <= 1 %; the conv1 kernel alone (fp32 math, bf16 store) is checked tighter."""
import numpy as np
import pytest
import torch

import slac_encoder_oracle as EO
from common import load_golden

pytestmark = pytest.mark.gpu


def _enc(size):
    from s2p_b200.networks.slac_encoder import SlacEncoderB200
    enc = SlacEncoderB200(3, 256, size)
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in EO.make_weights(size, seed=size).items()})
    return enc.cuda()


def _close(y, ref):
    err = np.abs(y - ref)
    rms = np.sqrt((err ** 2).mean()) / np.sqrt((ref ** 2).mean())
    assert err.max() <= 0.02 * np.abs(ref).max() and rms <= 0.01, (err.max(), np.abs(ref).max(), rms)


@pytest.mark.parametrize("size", [100, 64])
def test_forward_vs_reference_vectors(size):
    g = load_golden("slac_encoder")
    b, s = (int(v) for v in g[f"shape_{size}"])
    frames = EO.make_frames(size + 1, b, s, size)
    x = torch.from_numpy(frames).permute(0, 1, 4, 2, 3).float().div_(255.0).cuda()
    y = _enc(size)(x)
    assert y.shape == (b, s, 256) and y.dtype == torch.float32
    _close(y.cpu().numpy(), g[f"out_{size}"])


def test_fused_uint8_path_equals_float_path_and_oracle():
    """forward_from_buffer (gather + /255 fused into conv1) == forward(gather_states(...)) bit for bit."""
    from s2p_b200.slac_buffer import SlacReplayBufferB200
    from slac_dataset import make_dataset
    d = make_dataset(51, [40, 20], 0, h=100, w=100)
    buf = SlacReplayBufferB200(200, 8, (3, 100, 100), (6,), "cuda")
    buf.load_generated(d)
    enc = _enc(100)
    idx = np.random.RandomState(3).randint(0, buf._n, size=7)
    y_fused = enc.forward_from_buffer(buf, idx)
    x = buf.gather_states(idx)
    y_float = enc(x)
    assert torch.equal(y_fused, y_float)
    ref = EO.encoder_forward(EO.make_weights(100, seed=100), x.cpu().numpy(), 100)
    _close(y_fused.cpu().numpy(), ref)


def test_conv1_kernel_alone():
    import torch.nn.functional as F
    enc = _enc(100)
    x = torch.rand(5, 3, 100, 100, device="cuda")
    y = enc._conv1(5, 100, 100, x=x)
    sd = EO.make_weights(100, seed=100)
    ref = F.leaky_relu(F.conv2d(x.cpu(), torch.from_numpy(sd["net.0.weight"]), torch.from_numpy(sd["net.0.bias"]), 2, 2), 0.2)
    got = y.float().cpu().permute(0, 3, 1, 2)
    assert got.shape == (5, 64, 50, 50) and float(got[:, 32:].abs().max()) == 0.0
    assert float((got[:, :32] - ref).abs().max()) <= 4e-3 * float(ref.abs().max()) + 1e-6


def test_no_cpu_path():
    from s2p_b200.networks.slac_encoder import SlacEncoderB200
    with pytest.raises(RuntimeError):
        SlacEncoderB200(3, 256, 100)(torch.zeros(1, 1, 3, 100, 100))
    with pytest.raises(RuntimeError):
        SlacEncoderB200(3, 256, 84)


@pytest.mark.parametrize("size", [100, 64])
def test_decoder_vs_reference_vectors(size):
    """ConvTranspose stack on the implicit-GEMM kernel's data-gradient form vs the reference's own Decoder."""
    from s2p_b200.networks.slac_encoder import SlacDecoderB200
    g = load_golden("slac_encoder")
    b, s = (int(v) for v in g[f"dec_shape_{size}"])
    dec = SlacDecoderB200(288, 3, 1.0, size)
    dec.load_state_dict({k: torch.from_numpy(v) for k, v in EO.make_decoder_weights(size, seed=200 + size).items()})
    dec = dec.cuda()
    z = torch.from_numpy(np.random.RandomState(300 + size).randn(b, s, 288).astype(np.float32)).cuda()
    mean, std = dec(z)
    assert mean.shape == (b, s, 3, size, size) and float(std.min()) == float(std.max()) == 1.0
    _close(mean.cpu().numpy(), g[f"dec_mean_{size}"])
