"""TEST INFRASTRUCTURE: ctypes binding of oracle/_ref/libref_ba.so - the reference's own
BundleAdjuster and ReprojectionFactor compiled by oracle/build_ref.py (see there for what is
real and what is a stand-in). Exists only in the build container (/root/reference is needed to
build it); tests that want it skip elsewhere and use the committed fixture
tests/golden/ba_reference.npz instead.

Only tests/ and tests/golden/make_golden_ba.py may import this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build_ref

_lib = None


def available() -> bool:
    return build_ref.available() or os.path.exists(build_ref.LIB)


def load():
    global _lib
    if _lib is not None:
        return _lib
    path = build_ref.build() if build_ref.available() else build_ref.LIB
    lib = C.CDLL(path)
    vp, i64, u64, f64 = C.c_void_p, C.c_int64, C.c_uint64, C.c_double
    lib.ref_ba_create.restype = vp
    lib.ref_ba_create.argtypes = [u64, f64, f64, f64, f64]
    lib.ref_ba_destroy.argtypes = [vp]
    lib.ref_ba_destroy.restype = None
    lib.ref_ba_max_features.restype = u64
    lib.ref_ba_add_keyframe.restype = i64
    lib.ref_ba_add_keyframe.argtypes = [vp, vp, vp, vp, i64, vp, vp, i64, vp, i64, C.POINTER(i64)]
    lib.ref_ba_bundle_adjust.argtypes = [vp]
    lib.ref_ba_bundle_adjust.restype = None
    lib.ref_ba_last_pose.argtypes = [vp, vp]
    lib.ref_ba_world_points.argtypes = [vp, vp, i64, vp]
    lib.ref_ba_world_points.restype = None
    lib.ref_ba_problem_stats.argtypes = [vp, vp]
    lib.ref_ba_problem_stats.restype = None
    lib.ref_ba_evaluate_window.argtypes = [vp, vp, i64]
    lib.ref_ba_evaluate_window.restype = i64
    lib.ref_factor_evaluate.argtypes = [f64, f64, f64, f64, f64, vp, vp, vp, vp, vp]
    lib.ref_factor_evaluate.restype = None
    _lib = lib
    return lib


class RefBundleAdjuster:
    """BundleAdjuster(window_size, CameraInfo{focal, cx, cy, 0, 0, 0, 0, baseline}) of the reference"""

    def __init__(self, window: int, focal: float, cx: float, cy: float, baseline: float):
        self.lib = load()
        self.h = self.lib.ref_ba_create(window, focal, cx, cy, baseline)
        self.max_features = int(self.lib.ref_ba_max_features())

    def add_keyframe(self, pose7, tracked_ids, tracked_2d, new_2d, new_3d):
        """-> (new_ids exactly as the reference leaves Keyframe::new_ids, new_features_2d.size()
        after the call)"""
        pose = np.ascontiguousarray(pose7, np.float32).reshape(7)
        tid = np.ascontiguousarray(tracked_ids, np.uint64).reshape(-1)
        t2 = np.ascontiguousarray(tracked_2d, np.float32).reshape(-1, 2)
        n2 = np.ascontiguousarray(new_2d, np.float32).reshape(-1, 2)
        n3 = np.ascontiguousarray(new_3d, np.float32).reshape(-1, 3)
        assert t2.shape[0] == tid.shape[0] and n3.shape[0] == n2.shape[0]
        cap = 2 * (n2.shape[0] + self.max_features) + 8
        out = np.zeros(cap, np.uint64)
        after = C.c_int64(0)
        n = self.lib.ref_ba_add_keyframe(self.h, pose.ctypes.data, tid.ctypes.data, t2.ctypes.data,
                                         tid.shape[0], n2.ctypes.data, n3.ctypes.data, n2.shape[0],
                                         out.ctypes.data, cap, C.byref(after))
        assert n <= cap
        return out[:n].copy(), int(after.value)

    def bundle_adjust(self):
        self.lib.ref_ba_bundle_adjust(self.h)

    def last_pose(self):
        p = np.zeros(7, np.float32)
        return p if self.lib.ref_ba_last_pose(self.h, p.ctypes.data) else None

    def world_points(self, ids):
        ids = np.ascontiguousarray(ids, np.uint64).reshape(-1)
        out = np.zeros((ids.shape[0], 3), np.float32)
        self.lib.ref_ba_world_points(self.h, ids.ctypes.data, ids.shape[0], out.ctypes.data)
        return out

    def stats(self):
        s = np.zeros(6, np.int64)
        self.lib.ref_ba_problem_stats(self.h, s.ctypes.data)
        return dict(residual_blocks=int(s[0]), parameter_blocks=int(s[1]), api_errors=int(s[2]),
                    dropped_with_parameter=int(s[3]), poses=int(s[4]), constant_poses=int(s[5]))

    def evaluate_window(self):
        """every live residual block through the reference's Evaluate: dict of arrays
        pose (n, 7), point (n, 3), resid (n, 2), jac_pose (n, 2, 7), jac_point (n, 2, 3), obs (n, 2)"""
        n = self.stats()["residual_blocks"]
        buf = np.zeros((n, 34), np.float64)
        got = self.lib.ref_ba_evaluate_window(self.h, buf.ctypes.data, n)
        assert got == n
        return dict(pose=buf[:, :7].copy(), point=buf[:, 7:10].copy(), resid=buf[:, 10:12].copy(),
                    jac_pose=buf[:, 12:26].reshape(n, 2, 7).copy(),
                    jac_point=buf[:, 26:32].reshape(n, 2, 3).copy(), obs=buf[:, 32:34].copy())

    def close(self):
        if self.h:
            self.lib.ref_ba_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def factor_evaluate(obs, focal, cx, cy, pose7, point3):
    """ReprojectionFactor(obs.x, obs.y, info).Evaluate of the reference -> resid (2,), (2, 7), (2, 3)"""
    lib = load()
    pose = np.ascontiguousarray(pose7, np.float64).reshape(7)
    pt = np.ascontiguousarray(point3, np.float64).reshape(3)
    r, jp, jx = np.zeros(2), np.zeros(14), np.zeros(6)
    lib.ref_factor_evaluate(float(obs[0]), float(obs[1]), focal, cx, cy, pose.ctypes.data, pt.ctypes.data,
                            r.ctypes.data, jp.ctypes.data, jx.ctypes.data)
    return r, jp.reshape(2, 7), jx.reshape(2, 3)
