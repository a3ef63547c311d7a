"""Shared helpers of the parity tests."""
import numpy as np

from stereo_vo_b200 import synth

_SEQ_CACHE = {}


def frames(seed, n, cam=synth.KITTI00, motion=synth.KITTI_MOTION, start=0):
    key = (seed, n, cam, motion, start)
    if key not in _SEQ_CACHE:
        _SEQ_CACHE[key] = synth.sequence(seed, n, cam, motion, start)
    return _SEQ_CACHE[key]


def small_camera(w, h):
    """A reduced-size camera with KITTI-like geometry (for fast oracle runs)."""
    k = synth.KITTI00
    return synth.Camera(k.focal * w / k.width, w / 2.0 - 3.3, h / 2.0 + 1.7, k.baseline, w, h)


def stress_image(h, w, seed):
    """Texture with flat patches, a periodic stripe band and blur: hits the StereoBM texture
    and uniqueness rejections and gives plateau ties for the NMS."""
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (h, w)).astype(np.float64)
    # blur (3x3 box twice) for realistic gradients
    for _ in range(2):
        p = np.pad(img, 1, mode="edge")
        img = sum(p[dy:dy + h, dx:dx + w] for dy in range(3) for dx in range(3)) / 9.0
    img = (img - img.mean()) * 6.0 + 128.0
    img[h // 4:h // 4 + 40, w // 5:w // 5 + 120] = 77.0                  # flat patch
    xs = np.arange(w)
    stripe = 128 + 100 * np.sin(2 * np.pi * xs / 12.0)
    img[h // 2:h // 2 + 50, :] = stripe[None, :]                           # ambiguous stripes
    img[-30:, : w // 3] = 255.0
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def shifted(img, d):
    """right(x) = left(x + d) for an integer disparity d (edge replicated)."""
    out = np.empty_like(img)
    out[:, :img.shape[1] - d] = img[:, d:]
    out[:, img.shape[1] - d:] = img[:, -1:]
    return out
