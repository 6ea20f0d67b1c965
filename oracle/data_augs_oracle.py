"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the RAD augmentations of the reference
(/root/reference/examples/iql/data_augs.py), SURVEY.md section 8 (f) rank 4.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import this; the product path (s2p_b200/data_augs.py -> s2p_rad_augment /
s2p_rad_grayscale in libs2p_sm100.so) never does.

Parity pinned: tests/golden/data_augs.npz holds the outputs of the REFERENCE'S OWN functions (module imported by
file path with a stub for its unused matplotlib import, tests/golden/make_data_augs_golden.py) on seeded inputs;
tests/test_data_augs_oracle.py checks every function here against them bit for bit.

Two layers per augmentation:
  * `<name>_at(imgs, draws...)`: the pixel movement for given per-image draws (what the CUDA kernel is compared to);
  * `<name>(imgs, ...)`: the reference's signature, drawing from np.random in the reference's order.
Not restated: random_color_jitter (data_augs.py:226-241) needs `ColorJitterLayer`, whose import is commented out in
the reference (:5) -- the function cannot run there; random_convolution (:193-223) is a per-sample random-weight 3x3
conv drawn from torch's RNG inside the call (no index-stable oracle).
"""
import numpy as np


# ------------------------------------------------------------------ pixel movement for given draws
def crop_at(imgs, h1, w1, out):
    """data_augs.py:18-22: cropped[i] = img[:, h11:h11+out, w11:w11+out]"""
    n, c = imgs.shape[:2]
    res = np.empty((n, c, out, out), dtype=imgs.dtype)
    for i in range(n):
        res[i] = imgs[i, :, h1[i]:h1[i] + out, w1[i]:w1[i] + out]
    return res


def translate_at(imgs, h1s, w1s, size):
    """data_augs.py:247-251: zeros of [size, size], image pasted at (h1, w1)"""
    n, c, h, w = imgs.shape
    outs = np.zeros((n, c, size, size), dtype=imgs.dtype)
    for i in range(n):
        outs[i, :, h1s[i]:h1s[i] + h, w1s[i]:w1s[i] + w] = imgs[i]
    return outs


def cutout_at(imgs, h1, w1, fill=None):
    """data_augs.py:84-87 / :105-111: the box is rows [h1, h1+h1) x columns [w1, w1+w1) -- the offset doubles as the
    size in the reference -- clipped by numpy slicing; zero, or fill[i][c] cast to the image dtype."""
    res = imgs.copy()
    for i in range(imgs.shape[0]):
        if fill is None:
            res[i, :, h1[i]:h1[i] + h1[i], w1[i]:w1[i] + w1[i]] = 0
        else:
            res[i, :, h1[i]:h1[i] + h1[i], w1[i]:w1[i] + w1[i]] = fill[i].reshape(-1, 1, 1)
    return res


def flip_at(imgs, flags):
    """data_augs.py:131-145: mask * images.flip([3]) + (1 - mask) * images with a 0/1 mask per image"""
    res = imgs.copy()
    for i in range(imgs.shape[0]):
        if flags[i]:
            res[i] = imgs[i, :, :, ::-1]
    return res


def rot90_at(imgs, turns):
    """data_augs.py:164-186: sum over k of mask_k * images.rot90(k, [2, 3]), one mask is 1 per image"""
    res = np.empty_like(imgs)
    for i in range(imgs.shape[0]):
        res[i] = np.rot90(imgs[i], int(turns[i]), axes=(1, 2))
    return res


def grayscale_at(images, flags):
    """data_augs.py:26-66 with its dtypes: u8 = (images*255).type(uint8); gray = (u8_r*0.2989 + u8_g*0.587 +
    u8_b*0.114) in float32 (uint8 tensor * python float), .type(uint8).float(); out = (flag ? gray : u8) / 255."""
    images = np.asarray(images, dtype=np.float32)
    b, c, h, w = images.shape
    u8 = (images * np.float32(255.0)).astype(np.uint8).reshape(b, c // 3, 3, h, w)
    f = u8.astype(np.float32)
    gray = (f[:, :, 0] * np.float32(0.2989) + f[:, :, 1] * np.float32(0.587)) + f[:, :, 2] * np.float32(0.114)
    gray = gray.astype(np.uint8).astype(np.float32)[:, :, None] * np.ones((1, 1, 3, 1, 1), dtype=np.float32)
    m = np.asarray(flags).astype(bool)[:, None, None, None, None]
    out = np.where(m, gray, f)
    return (out.reshape(b, c, h, w) / np.float32(255.0)).astype(np.float32)


# ------------------------------------------------------------------ the reference's signatures and draw order
def random_crop(imgs, out=84):
    n, c, h, w = imgs.shape
    crop_max = h - out + 1                       # data_augs.py:15 (h for both axes)
    w1 = np.random.randint(0, crop_max, n)       # :16 -- w1 is drawn first
    h1 = np.random.randint(0, crop_max, n)       # :17
    return crop_at(imgs, h1, w1, out)


def random_cutout(imgs, min_cut=10, max_cut=30):
    n = imgs.shape[0]
    w1 = np.random.randint(min_cut, max_cut, n)  # :80
    h1 = np.random.randint(min_cut, max_cut, n)  # :81
    return cutout_at(imgs, h1, w1)


def random_cutout_color(imgs, min_cut=10, max_cut=30):
    n, c = imgs.shape[:2]
    w1 = np.random.randint(min_cut, max_cut, n)  # :99
    h1 = np.random.randint(min_cut, max_cut, n)  # :100
    rand_box = np.random.randint(0, 255, size=(n, c)) / 255.  # :103 (float64; cast by the assignment)
    return cutout_at(imgs, h1, w1, rand_box.astype(imgs.dtype))


def random_flip(images, p=.2):
    rnd = np.random.uniform(0., 1., size=(images.shape[0],))  # :132
    return flip_at(images, rnd <= p)


def random_rotation(images, p=.3):
    rnd = np.random.uniform(0., 1., size=(images.shape[0],))   # :168
    rnd_rot = np.random.randint(1, 4, size=(images.shape[0],))  # :169
    return rot90_at(images, rnd_rot * (rnd <= p))


def random_grayscale(images, p=.3):
    rnd = np.random.uniform(0., 1., size=(images.shape[0],))  # :56
    return grayscale_at(images, rnd <= p)


def random_translate(imgs, size, return_random_idxs=False, h1s=None, w1s=None):
    n, c, h, w = imgs.shape
    assert size >= h and size >= w
    h1s = np.random.randint(0, size - h + 1, n) if h1s is None else h1s  # :248
    w1s = np.random.randint(0, size - w + 1, n) if w1s is None else w1s  # :249
    outs = translate_at(imgs, h1s, w1s, size)
    if return_random_idxs:
        return outs, dict(h1s=h1s, w1s=w1s)
    return outs


def no_aug(x):
    return x


def grayscale(imgs):
    """data_augs.py:26-38 on a numpy array [B, frames*3, H, W] (uint8 or float32, values as they are): float32 luma
    with separately rounded fp32 products and sums, truncated to uint8, tiled over a new 3-channel axis."""
    b, c, h, w = imgs.shape
    x = imgs.reshape(b, c // 3, 3, h, w).astype(np.float32)
    g = (x[:, :, 0] * np.float32(0.2989) + x[:, :, 1] * np.float32(0.587)) + x[:, :, 2] * np.float32(0.114)
    g = g.astype(np.uint8).astype(np.float32)
    return np.repeat(g[:, :, None], 3, axis=2)


def random_convolution_at(imgs, weights):
    """data_augs.py:193-223 with the per-image filters given ([B,3,3,3,3] = nn.Conv2d.weight of each transformation):
    every frame (3 channels) of image b is convolved 3x3, padding 1, no bias, with weights[b]; float32 accumulation in
    float64 here (the checker), compared with a tolerance."""
    b, c, h, w = imgs.shape
    x = imgs.reshape(b, c // 3, 3, h, w).astype(np.float64)
    xp = np.pad(x, ((0, 0), (0, 0), (0, 0), (1, 1), (1, 1)))
    out = np.zeros((b, c // 3, 3, h, w), dtype=np.float64)
    for r in range(3):
        for s in range(3):
            patch = xp[:, :, :, r:r + h, s:s + w]                      # [b, f, ci, h, w]
            out += np.einsum("bfchw,boc->bfohw", patch, weights[:, :, :, r, s].astype(np.float64))
    return out.reshape(b, c, h, w).astype(np.float32)
