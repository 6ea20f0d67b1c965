"""Golden vectors of the SLAC encoder produced by the REFERENCE'S OWN module
(/root/reference/rlkit/torch/slac/network/latent.py Encoder, image_size 100 and 64) on seeded uint8 frames.
Run in the build container:  python tests/golden/make_slac_encoder_golden.py"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, "/root/reference")
import slac_encoder_oracle as EO  # noqa: E402
from rlkit.torch.slac.network.latent import Decoder, Encoder  # noqa: E402  (the reference's modules)

out = {}
for size, (b, s) in {100: (2, 9), 64: (3, 2)}.items():
    sd = EO.make_weights(size, seed=size)
    enc = Encoder(3, 256, size)
    enc.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    frames = EO.make_frames(size + 1, b, s, size)                      # uint8 [B,S,H,W,3]
    x = torch.from_numpy(frames).permute(0, 1, 4, 2, 3).float().div_(255.0)   # buffer.py:135
    with torch.no_grad():
        y = enc(x).numpy()
    out[f"out_{size}"] = y
    out[f"shape_{size}"] = np.array([b, s])
    print(size, y.shape, float(np.abs(y).mean()), float(y.std()))
# ---- Decoder (latent.py:55-115) on seeded latents
for size, (b, s) in {100: (1, 2), 64: (1, 2)}.items():
    sd = EO.make_decoder_weights(size, seed=200 + size)
    dec = Decoder(288, 3, 1.0, size) if size == 64 else Decoder(input_dim=288, output_dim=3, std=1.0, image_size=size)
    dec.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    z = np.random.RandomState(300 + size).randn(b, s, 288).astype(np.float32)
    with torch.no_grad():
        mean, std = dec(torch.from_numpy(z))
    out[f"dec_mean_{size}"], out[f"dec_shape_{size}"] = mean.numpy(), np.array([b, s])
    assert float(std.min()) == float(std.max()) == 1.0
    print("decoder", size, mean.shape, float(mean.abs().mean()), float(mean.std()))
path = os.path.join(HERE, "slac_encoder.npz")
np.savez_compressed(path, **out)
print("wrote", path, os.path.getsize(path))
