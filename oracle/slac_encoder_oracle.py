"""CPU oracle of the SLAC image encoder (SURVEY.md section 8 f, rank 3).  TEST INFRASTRUCTURE ONLY.

fp32 restatement of ``Encoder`` (/root/reference/rlkit/torch/slac/network/latent.py:117-171): a stack of
Conv2d + LeakyReLU(0.2) (image_size 100: 5x5 s2 p2, four 3x3 s2 p1, 4x4 valid; image_size 64: one 3x3 fewer),
``forward`` folds [B,S] into the batch and returns [B,S,output_dim].

PARITY PINNED: tests/golden/make_slac_encoder_golden.py imports the reference's own Encoder, loads the seeded
weights of ``make_weights`` into it and stores its outputs; tests hold this restatement and the CUDA path to them.
"""
import numpy as np
import torch
import torch.nn.functional as F


def specs(image_size=100, input_dim=3, output_dim=256):
    """(cin, cout, k, stride, pad), latent.py:123-160"""
    s = [(input_dim, 32, 5, 2, 2), (32, 64, 3, 2, 1), (64, 128, 3, 2, 1), (128, 256, 3, 2, 1)]
    if image_size == 100:
        s.append((256, 256, 3, 2, 1))
    return s + [(256, output_dim, 4, 1, 0)]


def make_weights(image_size=100, seed=0, output_dim=256):
    """seeded numpy weights under the reference's state_dict keys (net.0.weight, net.2.weight, ...); xavier-uniform
    bounds as initialize_weight (network/initializer.py:4-8) but NON-zero biases so the bias path is exercised."""
    rs = np.random.RandomState(seed)
    sd = {}
    for i, (cin, cout, k, _, _) in enumerate(specs(image_size, 3, output_dim)):
        bound = np.sqrt(6.0 / ((cin + cout) * k * k))
        sd[f"net.{2 * i}.weight"] = rs.uniform(-bound, bound, size=(cout, cin, k, k)).astype(np.float32)
        sd[f"net.{2 * i}.bias"] = (rs.randn(cout) * 0.05).astype(np.float32)
    return sd


def make_frames(seed, b, s, image_size=100):
    """smooth-ish uint8 HWC frames (blocks + noise) so activations are not pure noise"""
    rs = np.random.RandomState(seed)
    base = rs.randint(0, 256, size=(b, s, image_size // 10 + 1, image_size // 10 + 1, 3))
    up = np.repeat(np.repeat(base, 10, axis=2), 10, axis=3)[:, :, :image_size, :image_size]
    return np.clip(up + rs.randint(-20, 21, size=up.shape), 0, 255).astype(np.uint8)


def encoder_forward(sd, x, image_size=100):
    """x: float [B,S,3,H,W] in [0,1] -> [B,S,output_dim]  (latent.py:162-170)"""
    x = torch.as_tensor(x, dtype=torch.float32)
    B, S, C, H, W = x.shape
    y = x.reshape(B * S, C, H, W)
    for i, (_, _, _, st, pad) in enumerate(specs(image_size)):
        y = F.conv2d(y, torch.as_tensor(sd[f"net.{2 * i}.weight"]), torch.as_tensor(sd[f"net.{2 * i}.bias"]), st, pad)
        y = F.leaky_relu(y, 0.2)
    return y.reshape(B, S, -1).numpy()


# ---------------------------------------------------------------------------------------------------
# Decoder (latent.py:55-115): ConvTranspose2d + LeakyReLU(0.2) stack, [B,S,latent] -> ([B,S,3,H,W], std * ones)
def decoder_specs(image_size=100, input_dim=288, output_dim=3):
    """(cin, cout, k, stride, pad, output_padding), latent.py:63-101"""
    if image_size == 64:
        return [(input_dim, 256, 4, 1, 0, 0), (256, 128, 3, 2, 1, 1), (128, 64, 3, 2, 1, 1), (64, 32, 3, 2, 1, 1),
                (32, output_dim, 5, 2, 2, 1)]
    return [(input_dim, 256, 4, 1, 0, 0), (256, 256, 3, 2, 1, 0), (256, 128, 3, 2, 1, 0), (128, 64, 3, 2, 1, 0),
            (64, 32, 3, 2, 1, 1), (32, output_dim, 5, 2, 2, 1)]


def make_decoder_weights(image_size=100, seed=0, input_dim=288):
    rs = np.random.RandomState(seed)
    sd = {}
    for i, (cin, cout, k, _, _, _) in enumerate(decoder_specs(image_size, input_dim)):
        bound = np.sqrt(6.0 / ((cin + cout) * k * k))
        sd[f"net.{2 * i}.weight"] = rs.uniform(-bound, bound, size=(cin, cout, k, k)).astype(np.float32)  # ConvTranspose2d layout
        sd[f"net.{2 * i}.bias"] = (rs.randn(cout) * 0.05).astype(np.float32)
    return sd


def decoder_forward(sd, z, image_size=100, std=1.0):
    """z: [B,S,latent] -> (mean [B,S,3,H,W], std [B,S,3,H,W])  (latent.py:105-115)"""
    z = torch.as_tensor(z, dtype=torch.float32)
    B, S, L = z.shape
    y = z.reshape(B * S, L, 1, 1)
    for i, (_, _, _, st, pad, op) in enumerate(decoder_specs(image_size, L)):
        y = F.conv_transpose2d(y, torch.as_tensor(sd[f"net.{2 * i}.weight"]), torch.as_tensor(sd[f"net.{2 * i}.bias"]), st, pad, op)
        y = F.leaky_relu(y, 0.2)
    y = y.reshape(B, S, *y.shape[1:])
    return y.numpy(), np.full(y.shape, std, dtype=np.float32)
