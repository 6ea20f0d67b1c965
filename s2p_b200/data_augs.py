"""RAD image augmentations on the B200 (SURVEY.md section 8 f, rank 4): same function names, arguments and
np.random draw order as the reference module (/root/reference/examples/iql/data_augs.py, used as
``from examples.iql import data_augs as rad``, rlkit/launchers/experiments/awac/finetune_rl.py:34), so a seeded
caller gets the same pixels -- but every function is ONE kernel launch over a batch that stays in HBM
(csrc/rad_augs.cu) instead of a per-image host loop (crop / cutout / translate) or several full-size tensor passes
(flip / rotation / grayscale).

Input: a CUDA tensor ``[B, C, H, W]`` of uint8 or float32 (returned: a new CUDA tensor), or a numpy array of those
dtypes (uploaded to ``device``, result returned as numpy: the reference's host-array contract for crop / cutout /
translate).  The per-image draws are made on the host and uploaded as int32 (a few bytes per image).

No CPU path: without a CUDA device or libs2p_sm100.so every function raises.
Not provided: random_color_jitter (its ColorJitterLayer import is commented out in the reference, data_augs.py:5,
so it raises NameError there too).
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .ops import _stream

_ELEM = {torch.uint8: 1, torch.float32: 4}


def _enter(imgs, device):
    """-> (contiguous CUDA tensor, was_numpy)"""
    was_np = isinstance(imgs, np.ndarray)
    t = torch.from_numpy(np.ascontiguousarray(imgs)) if was_np else imgs
    if was_np:
        dev = torch.device(device if device is not None else "cuda")
        if dev.type != "cuda" or not torch.cuda.is_available():
            raise RuntimeError("s2p_b200.data_augs needs a CUDA device (no CPU fallback)")
        t = t.to(dev)
    if t.device.type != "cuda":
        raise RuntimeError("s2p_b200.data_augs needs a CUDA tensor (no CPU fallback)")
    if t.dim() != 4 or t.dtype not in _ELEM:
        raise RuntimeError(f"expected uint8 or float32 [B,C,H,W], got {tuple(t.shape)} {t.dtype}")
    return t.contiguous(), was_np


def _leave(t, was_np):
    return t.cpu().numpy() if was_np else t


def _draws(dev, *arrays):
    """One host->device copy for all per-image draws of a call: int32 [len(arrays), n], row k = arrays[k]."""
    host = np.empty((len(arrays), len(arrays[0])), dtype=np.int32)
    for k, a in enumerate(arrays):
        host[k] = a
    # pageable source: cudaMemcpyAsync returns once the bytes are staged, so `host` may be dropped at once, and the call
    # does not wait for the stream to drain as the blocking copy did
    return torch.from_numpy(host).to(dev, non_blocking=True)


def _augment(x, op, ho, wo, p0, p1=None, fill=None):
    n, c, h, w = x.shape
    out = torch.empty((n, c, ho, wo), dtype=x.dtype, device=x.device)
    if n == 0:
        return out
    def launch():
        d = _draws(x.device, p0) if p1 is None else _draws(x.device, p0, p1)
        base = d.data_ptr()
        _lib.check(_lib.load().s2p_rad_augment(x.data_ptr(), out.data_ptr(), _ELEM[x.dtype], n, c, h, w, ho, wo, op,
                                               base, 0 if p1 is None else base + 4 * n,
                                               0 if fill is None else fill.data_ptr(), _stream()), "s2p_rad_augment")
    if torch.cuda.current_device() == x.device.index:     # the usual case: skip the device-guard context manager
        launch()
    else:
        with torch.cuda.device(x.device):
            launch()
    return out


def _np_dtype(t):
    return np.uint8 if t.dtype == torch.uint8 else np.float32


def random_crop(imgs, out=84, device=None):
    x, was_np = _enter(imgs, device)
    n, _, h, w = x.shape
    if out > h or out > w:
        raise ValueError(f"crop size {out} larger than the image {h}x{w}")
    span = h - out + 1
    w1 = np.random.randint(0, span, n)   # the reference draws the column offsets first, both from the height's span
    h1 = np.random.randint(0, span, n)
    if out + int(w1.max(initial=0)) > w:
        raise ValueError("crop window leaves the image (non-square input)")
    return _leave(_augment(x, _lib.RAD_CROP, out, out, h1, w1), was_np)


def _cut(imgs, min_cut, max_cut, color, device):
    x, was_np = _enter(imgs, device)
    n, c, h, w = x.shape
    w1 = np.random.randint(min_cut, max_cut, n)
    h1 = np.random.randint(min_cut, max_cut, n)
    fill = None
    if color:
        box = np.random.randint(0, 255, size=(n, c)) / 255.
        fill = torch.from_numpy(np.ascontiguousarray(box.astype(_np_dtype(x)))).to(x.device)  # numpy's assignment cast
    return _leave(_augment(x, _lib.RAD_CUTOUT, h, w, h1, w1, fill), was_np)


def random_cutout(imgs, min_cut=10, max_cut=30, device=None):
    return _cut(imgs, min_cut, max_cut, False, device)


def random_cutout_color(imgs, min_cut=10, max_cut=30, device=None):
    return _cut(imgs, min_cut, max_cut, True, device)


def random_flip(images, p=.2, device=None):
    x, was_np = _enter(images, device)
    n, _, h, w = x.shape
    flags = np.random.uniform(0., 1., size=(n,)) <= p
    return _leave(_augment(x, _lib.RAD_FLIP, h, w, flags), was_np)


def random_rotation(images, p=.3, device=None):
    x, was_np = _enter(images, device)
    n, _, h, w = x.shape
    if h != w:
        raise ValueError("random_rotation needs square images (the reference's rot90 blend does too)")
    rnd = np.random.uniform(0., 1., size=(n,))
    turns = np.random.randint(1, 4, size=(n,)) * (rnd <= p)
    return _leave(_augment(x, _lib.RAD_ROT90, h, w, turns), was_np)


def random_grayscale(images, p=.3, device=None):
    x, was_np = _enter(images, device)
    n, c, h, w = x.shape
    if x.dtype != torch.float32 or c % 3:
        raise RuntimeError("random_grayscale: float32 images in [0,1] with 3 channels per frame")
    flags = np.random.uniform(0., 1., size=(n,)) <= p
    out = torch.empty_like(x)
    if n:
        with torch.cuda.device(x.device):
            m = _draws(x.device, flags)
            _lib.check(_lib.load().s2p_rad_grayscale(x.data_ptr(), out.data_ptr(), n, c // 3, h, w, m.data_ptr(), _stream()),
                       "s2p_rad_grayscale")
    return _leave(out, was_np)


def grayscale(imgs):
    """data_augs.py:26-38: [B, frames*3, H, W] (uint8 or float32, values as they are) -> float32 [B, frames, 3, H, W],
    every frame replaced by its truncated luma, tiled over the 3 channels."""
    x, _ = _enter(imgs, None)
    n, c, h, w = x.shape
    if c % 3:
        raise RuntimeError("grayscale: 3 channels per frame")
    out = torch.empty(n, c // 3, 3, h, w, dtype=torch.float32, device=x.device)
    if n:
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().s2p_rad_grayscale_plain(x.data_ptr(), _ELEM[x.dtype], out.data_ptr(), n, c // 3, h, w, _stream()),
                       "s2p_rad_grayscale_plain")
    return out


def random_convolution(imgs, weights=None, return_weights=False):
    """data_augs.py:193-223 ("network randomization"): every image gets its own random 3x3 conv 3 -> 3 channels
    (xavier-normal weights, padding 1, no bias) applied to each of its frames.  imgs: float32 CUDA [B, frames*3, H, W].

    The weights come from torch's RNG exactly as in the reference: it constructs ``nn.Conv2d(3, 3, 3, bias=False,
    padding=1)`` (whose default init consumes the CPU generator), moves it to the images' device and calls
    ``xavier_normal_`` on its weight once per image (the DEVICE generator) -- the same calls are made here, so a
    caller that seeds torch gets the reference's filters.  What changes is the arithmetic's shape: one launch over
    the batch instead of B conv launches and B torch.cat copies of a growing result.  ``weights`` ([B,3,3,3,3]
    float32) overrides the draw (tests, replay)."""
    x, _ = _enter(imgs, None)
    if x.dtype != torch.float32 or x.shape[1] % 3:
        raise RuntimeError("random_convolution: float32 images with 3 channels per frame")
    n, c, h, w = x.shape
    if weights is None:
        rand_conv = torch.nn.Conv2d(3, 3, kernel_size=3, bias=False, padding=1).to(x.device)      # same RNG consumption as :208
        rand_conv.weight.requires_grad = False
        weights = torch.empty(n, 3, 3, 3, 3, dtype=torch.float32, device=x.device)
        for i in range(n):                                                                       # :211-212, one draw per image
            torch.nn.init.xavier_normal_(rand_conv.weight.data)
            weights[i].copy_(rand_conv.weight.data)
    else:
        weights = torch.as_tensor(weights, dtype=torch.float32).to(x.device).contiguous()
        if tuple(weights.shape) != (n, 3, 3, 3, 3):
            raise RuntimeError(f"random_convolution: weights must be [{n}, 3, 3, 3, 3]")
    out = torch.empty_like(x)
    if n:
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().s2p_rad_random_convolution(x.data_ptr(), out.data_ptr(), weights.data_ptr(), n, c // 3, h, w,
                                                              _stream()), "s2p_rad_random_convolution")
    return (out, weights) if return_weights else out


def random_translate(imgs, size, return_random_idxs=False, h1s=None, w1s=None, device=None):
    x, was_np = _enter(imgs, device)
    n, _, h, w = x.shape
    assert size >= h and size >= w
    h1s = np.random.randint(0, size - h + 1, n) if h1s is None else h1s
    w1s = np.random.randint(0, size - w + 1, n) if w1s is None else w1s
    a, b = np.asarray(h1s), np.asarray(w1s)
    if n and (a.min() < 0 or b.min() < 0 or a.max() + h > size or b.max() + w > size):
        raise ValueError("random_translate: offsets place the image outside the canvas")
    outs = _leave(_augment(x, _lib.RAD_TRANSLATE, size, size, a, b), was_np)
    if return_random_idxs:
        return outs, dict(h1s=h1s, w1s=w1s)
    return outs


def no_aug(x):
    return x
