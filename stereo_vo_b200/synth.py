"""Synthetic KITTI-shaped stereo sequences.

One fronto-parallel textured plane at depth Z_t seen by a rectified stereo pair, so the
sequence is 3D-consistent (PnP on the triangulated points recovers the camera motion):

    left_t (x, y) = tex((x - cx) * s_t + k * tx_t + ox, (y - cy) * s_t + oy)
    right_t(x, y) = left_t(x + d_t, y)
    s_t = Z_t / Z_ref,  k = f / Z_ref,  d_t = f * b / Z_t

Everything is plain numpy float64 arithmetic (add / mul / floor only on the pixel path), so a
given (seed, t) gives the same bytes on every machine; tests pin that with sha256 digests.
There is no network and no dataset in this repo: these frames stand in for KITTI image_0/image_1
(reference: src/kitti_node.cpp:56-68 reads the PNG pair that this replaces).
"""
from __future__ import annotations

import hashlib
import math
from dataclasses import dataclass

import numpy as np

# texture canvas; sampled with wrap-around so any camera motion stays inside it
TEX_W = 4096
TEX_H = 1024


@dataclass(frozen=True)
class Camera:
    """Intrinsics as the reference reads them (src/vo_node.cpp:82-98: narrowed to float)."""
    focal: float
    cx: float
    cy: float
    baseline: float
    width: int
    height: int

    def f32(self):
        return (np.float32(self.focal), np.float32(self.cx), np.float32(self.cy),
                np.float32(self.baseline))


# config/kitti00.yaml:1-4 and config/d435i.yaml:1-4 of the reference
KITTI00 = Camera(7.18856e+02, 6.071928e+02, 1.852157e+02, 0.537165718864418, 1241, 376)
D435I = Camera(385.7544860839844, 323.1204833984375, 236.7432098388672, 0.05, 640, 480)
UHD4K = Camera(7.18856e+02 * 3.0, 1920.0, 1080.0, 0.537165718864418, 3840, 2160)


def _upsample_bilinear(g: np.ndarray, cell: int, h: int, w: int) -> np.ndarray:
    """Separable bilinear upsampling of a coarse grid by an integer factor (numpy only)."""
    ys = np.arange(h, dtype=np.float64) / cell
    xs = np.arange(w, dtype=np.float64) / cell
    y0 = np.floor(ys).astype(np.int64)
    x0 = np.floor(xs).astype(np.int64)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    rows = g[y0, :] * (1.0 - fy) + g[y0 + 1, :] * fy
    return rows[:, x0] * (1.0 - fx) + rows[:, x0 + 1] * fx


_TEX_CACHE: dict = {}


def make_texture(seed: int) -> np.ndarray:
    """Six octaves of seeded Gaussian value noise + 200 random rectangles, contrast-stretched
    so that image sigma is about 60-70 (LK's minEigThreshold=1e-2 in the reference,
    src/feature_tracker.cpp:26, rejects most tracks on low-contrast input)."""
    if seed in _TEX_CACHE:
        return _TEX_CACHE[seed]
    rng = np.random.default_rng(seed)
    tex = np.zeros((TEX_H, TEX_W), dtype=np.float64)
    for cell, amp in zip((64, 32, 16, 8, 4, 2), (60.0, 40.0, 30.0, 20.0, 12.0, 8.0)):
        gh = TEX_H // cell + 2
        gw = TEX_W // cell + 2
        g = rng.standard_normal((gh, gw))
        # make the coarse grid periodic so that wrap-around sampling is seamless
        g[-2:, :] = g[:2, :]
        g[:, -2:] = g[:, :2]
        tex += amp * _upsample_bilinear(g, cell, TEX_H, TEX_W)
    for _ in range(200):
        x0 = int(rng.integers(0, TEX_W - 8))
        y0 = int(rng.integers(0, TEX_H - 8))
        rw = int(rng.integers(8, 120))
        rh = int(rng.integers(8, 90))
        val = float(rng.integers(-80, 81))
        tex[y0:y0 + rh, x0:x0 + rw] += val
    tex = (tex - tex.mean()) * 2.5 + 128.0
    if len(_TEX_CACHE) > 8:
        _TEX_CACHE.clear()
    _TEX_CACHE[seed] = tex
    return tex


def _sample_separable(tex: np.ndarray, u: np.ndarray, v: np.ndarray) -> np.ndarray:
    """Bilinear sampling of `tex` at the outer product grid (v[y], u[x]) with wrap-around.
    The mapping of a fronto-parallel plane is separable, so two 1-D gathers suffice."""
    v0 = np.floor(v)
    u0 = np.floor(u)
    fv = (v - v0)[:, None]
    fu = (u - u0)[None, :]
    v0i = v0.astype(np.int64) % TEX_H
    v1i = (v0i + 1) % TEX_H
    u0i = u0.astype(np.int64) % TEX_W
    u1i = (u0i + 1) % TEX_W
    # blend the two texture rows only over the column span that is sampled (the same
    # arithmetic per element as blending whole rows; a wrapped span takes the whole row)
    lo, hi = int(u0i.min()), int(u1i.max())
    if hi - lo + 1 > 2 * len(u) + 8:
        lo, hi = 0, TEX_W - 1
    rows = tex[v0i, lo:hi + 1] * (1.0 - fv) + tex[v1i, lo:hi + 1] * fv
    return rows[:, u0i - lo] * (1.0 - fu) + rows[:, u1i - lo] * fu


@dataclass(frozen=True)
class Motion:
    """Scene / trajectory parameters of one synthetic sequence."""
    z_ref: float = 15.0
    z_amp: float = 4.0
    z_period: float = 80.0
    tx_per_frame: float = 0.12
    noise_sigma: float = 1.5


KITTI_MOTION = Motion()
# d435i: f*b = 19.3 px*m, so Z_ref = 1.0 m gives d ~ 19 px (< 48, the StereoBM range)
D435I_MOTION = Motion(z_ref=1.0, z_amp=0.2, z_period=80.0, tx_per_frame=0.004, noise_sigma=1.5)
# 4K: f*b = 1158 px*m; d < 48 needs Z > 24 m
UHD4K_MOTION = Motion(z_ref=40.0, z_amp=8.0, z_period=80.0, tx_per_frame=0.12, noise_sigma=1.5)


def depth_at(t: int, motion: Motion = KITTI_MOTION) -> float:
    return motion.z_ref + motion.z_amp * math.sin(2.0 * math.pi * t / motion.z_period)


def ground_truth(t: int, motion: Motion = KITTI_MOTION):
    """Camera position in the world frame of frame 0: (tx_t, 0, Z_0 - Z_t)."""
    return (motion.tx_per_frame * t, 0.0, depth_at(0, motion) - depth_at(t, motion))


def stereo_frame(seed: int, t: int, cam: Camera = KITTI00, motion: Motion = KITTI_MOTION):
    """Return (left, right) uint8 images of frame t of sequence `seed`."""
    tex = make_texture(seed)
    z = depth_at(t, motion)
    s = z / motion.z_ref
    k = cam.focal / motion.z_ref
    d = cam.focal * cam.baseline / z
    tx = motion.tx_per_frame * t
    ox = TEX_W / 2.0
    oy = TEX_H / 2.0
    xs = np.arange(cam.width, dtype=np.float64)
    ys = np.arange(cam.height, dtype=np.float64)
    v = (ys - cam.cy) * s + oy
    u_left = (xs - cam.cx) * s + k * tx + ox
    u_right = (xs + d - cam.cx) * s + k * tx + ox
    out = []
    for which, u in ((0, u_left), (1, u_right)):
        img = _sample_separable(tex, u, v)
        rng = np.random.default_rng(1000 + t + 100003 * which + 7919 * seed)
        img = img + motion.noise_sigma * rng.standard_normal(img.shape)
        out.append(np.clip(np.floor(img + 0.5), 0.0, 255.0).astype(np.uint8))
    return out[0], out[1]


def sequence(seed: int, n_frames: int, cam: Camera = KITTI00, motion: Motion = KITTI_MOTION,
             start: int = 0):
    """List of (left, right) pairs for frames start..start+n_frames-1."""
    return [stereo_frame(seed, t, cam, motion) for t in range(start, start + n_frames)]


def digest(*arrays: np.ndarray) -> str:
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def read_camera_yaml(path: str):
    """The seven flat keys of the reference's camera files (config/kitti00.yaml:1-7), read the
    way src/vo_node.cpp:82-98 does: four floats narrowed to f32, two topic strings and an int
    frame_rate that are parsed and then unused."""
    vals = {}
    with open(path, "r", encoding="utf-8") as fh:
        for line in fh:
            line = line.split("#", 1)[0].strip()
            if not line or ":" not in line:
                continue
            key, val = line.split(":", 1)
            vals[key.strip()] = val.strip().strip('"')
    out = {
        "focal_length": float(np.float32(float(vals["focal_length"]))),
        "cx": float(np.float32(float(vals["cx"]))),
        "cy": float(np.float32(float(vals["cy"]))),
        "baseline": float(np.float32(float(vals["baseline"]))),
        "left_topic": vals.get("left_topic", ""),
        "right_topic": vals.get("right_topic", ""),
        "frame_rate": int(float(vals.get("frame_rate", "0"))),
    }
    return out
