"""TEST INFRASTRUCTURE: drives oracle/_ref/libref_frontend.so - the reference's OWN
ImageProcessor::process / FeatureTracker::track_features / triangulate_stereo /
BundleAdjuster::add_keyframe, compiled where they lie by oracle/build_ref.py - and serves the
OpenCV algorithms they call with the real OpenCV (cv2), through the callback table of
oracle/ref_stubs/opencv2/opencv.hpp. Every callback also records the arguments the reference's
code passed, so the tests can hold the literal parameters of the restatement and of the product
(300 / 0.1 / 30, (21, 21) / 3 / {COUNT+EPS, 30, 0.01} / 0 / 1e-2, (48, 21), 100 / 8.0 / 0.99)
against the ones the compiled reference really uses.

Only tests/ and bench.py's CPU-baseline leg may import this module; it needs /root/reference to
build the library (build container) and runs wherever the built library is.
"""
from __future__ import annotations

import ctypes as C

import cv2
import numpy as np

from . import build_ref

_u8p, _f32p, _f64p, _i32p, _i16p = (C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_double),
                                    C.POINTER(C.c_int), C.POINTER(C.c_short))
_GFTT = C.CFUNCTYPE(C.c_int, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, _f32p, C.c_int)
_LK = C.CFUNCTYPE(None, _u8p, C.c_int, _u8p, C.c_int, C.c_int, C.c_int, _f32p, C.c_int, _f32p, _u8p, _f32p,
                  C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_double)
_SBM = C.CFUNCTYPE(None, _u8p, C.c_int, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _i16p)
_PNP = C.CFUNCTYPE(C.c_int, _f32p, _f32p, C.c_int, _f32p, _f64p, _f64p, C.c_int, C.c_int, C.c_int,
                   C.c_float, C.c_double, _i32p, C.c_int)
_ROD = C.CFUNCTYPE(None, _f64p, C.c_int, _f64p)
_GEMM = C.CFUNCTYPE(None, _f64p, _f64p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, _f64p)
_DEPTH = {5: np.float32, 6: np.float64}


class _Callbacks(C.Structure):
    _fields_ = [("gftt", _GFTT), ("lk", _LK), ("stereo_bm", _SBM), ("pnp", _PNP), ("rodrigues", _ROD),
                ("gemm", _GEMM)]


def _img(ptr, rows, cols, step):
    """a view of the caller's pixels (the stand-in Mats are continuous: no copy needed)"""
    a = np.ctypeslib.as_array(ptr, (rows, step))
    return a if step == cols else np.ascontiguousarray(a[:, :cols])


class RefPipeline:
    """vo_node's construction (src/vo_node.cpp:104-119) + one process() per stereo pair, on the
    reference's compiled classes. `calls` holds, per frame, what every OpenCV call was given and
    what it answered."""

    def __init__(self, focal, cx, cy, baseline, min_feature_distance=30.0, parallax_thresh=20.0, window=5,
                 guess="carried", record=True):
        """record: keep, per frame, the arguments and results of every OpenCV call in `calls`
        (the tests read them; the timing legs of bench.py switch it off).
        guess: how the extrinsic guess reaches cv2.solvePnPRansac. "as_stored": the values the
        reference's CV_32F rvec / tvec Mats hold (OpenCV 3.x keeps them in that type between
        keyframes). "carried": the float64 vectors the previous call returned whenever the stored
        ones are exactly their float32 rounding - how oracle/cv2_frontend.py carries them (cv2
        hands back float64). RANSAC's hypotheses start from the guess, so the two can differ in a
        borderline inlier; the control flow under test is the same."""
        assert guess in ("carried", "as_stored")
        self.guess = guess
        self.record = record
        self._last_pnp = None
        # built from the reference tree when it is here; elsewhere (the GPU box) the library
        # built in the container travels with the repository snapshot
        if build_ref.available():
            build_ref.build()
        self.lib = C.CDLL(build_ref.FRONTEND_LIB)
        lib = self.lib
        lib.ref_vo_create.restype = C.c_void_p
        lib.ref_vo_create.argtypes = [C.c_float] * 6 + [C.c_uint64]
        lib.ref_vo_destroy.argtypes = [C.c_void_p]
        lib.ref_vo_destroy.restype = None
        lib.ref_vo_process.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double]
        lib.ref_vo_last_error.restype = C.c_char_p
        lib.ref_vo_tracker.restype = C.c_int64
        lib.ref_vo_tracker.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]
        lib.ref_vo_keyframe_sizes.argtypes = [C.c_void_p, C.c_void_p]
        lib.ref_vo_keyframe.argtypes = [C.c_void_p] * 7
        lib.ref_vo_keyframe.restype = None
        lib.ref_cv_set_callbacks.argtypes = [C.POINTER(_Callbacks)]
        lib.ref_cv_set_callbacks.restype = None
        self.calls = []
        self._cb = _Callbacks(_GFTT(self._gftt), _LK(self._lk), _SBM(self._sbm), _PNP(self._pnp),
                              _ROD(self._rodrigues), _GEMM(self._gemm))
        lib.ref_cv_set_callbacks(C.byref(self._cb))
        self.h = lib.ref_vo_create(focal, cx, cy, baseline, min_feature_distance, parallax_thresh, window)
        self.log = []

    # ---- the OpenCV algorithms, real ones -------------------------------------------------
    def _gftt(self, img, rows, cols, step, max_corners, quality, min_distance, out, cap):
        image = _img(img, rows, cols, step)
        pts = cv2.goodFeaturesToTrack(image, max_corners, quality, min_distance)
        pts = np.zeros((0, 2), np.float32) if pts is None else pts.reshape(-1, 2)
        n = min(pts.shape[0], cap)
        np.ctypeslib.as_array(out, (2 * max(cap, 1),))[:2 * n] = pts[:n].reshape(-1)
        if self.record:
            self.calls.append(("goodFeaturesToTrack", dict(max_corners=max_corners, quality=quality,
                                                           min_distance=min_distance, corners=pts[:n].copy())))
        return n

    def _lk(self, prev, pstep, nxt, nstep, rows, cols, pts, n, out_pts, status, err, win_w, win_h,
            max_level, crit_type, crit_count, crit_eps, flags, min_eig):
        a, b = _img(prev, rows, cols, pstep), _img(nxt, rows, cols, nstep)
        p0 = np.ctypeslib.as_array(pts, (n, 2)).copy().reshape(-1, 1, 2)
        p1, st, er = cv2.calcOpticalFlowPyrLK(a, b, p0, None, winSize=(win_w, win_h), maxLevel=max_level,
                                              criteria=(crit_type, crit_count, crit_eps), flags=flags,
                                              minEigThreshold=min_eig)
        np.ctypeslib.as_array(out_pts, (n, 2))[:] = p1.reshape(-1, 2)
        np.ctypeslib.as_array(status, (n,))[:] = st.reshape(-1)
        np.ctypeslib.as_array(err, (n,))[:] = er.reshape(-1)
        if self.record:
            self.calls.append(("calcOpticalFlowPyrLK", dict(win=(win_w, win_h), max_level=max_level,
                                                            criteria=(crit_type, crit_count, crit_eps), flags=flags,
                                                            min_eig=min_eig, pts=p0.reshape(-1, 2).copy(),
                                                            out=p1.reshape(-1, 2).copy(), status=st.reshape(-1).copy())))

    def _sbm(self, left, lstep, right, rstep, rows, cols, ndisp, block, out):
        disp = cv2.StereoBM_create(ndisp, block).compute(_img(left, rows, cols, lstep), _img(right, rows, cols, rstep))
        np.ctypeslib.as_array(out, (rows, cols))[:] = disp
        if self.record:
            self.calls.append(("StereoBM", dict(ndisp=ndisp, block=block, disp16=disp.copy())))

    def _pnp(self, world, image, n, k9, rvec, tvec, depth, use_guess, iterations, reproj, confidence, inliers, cap):
        w = np.ctypeslib.as_array(world, (n, 3)).copy()
        im = np.ctypeslib.as_array(image, (n, 2)).copy()
        K = np.ctypeslib.as_array(k9, (9,)).copy().reshape(3, 3)
        rv = np.ctypeslib.as_array(rvec, (3,))
        tv = np.ctypeslib.as_array(tvec, (3,))
        # (float64 copies of the guess: cv2 4.13's binding misreads float32 in/out vectors, see
        # oracle/cv2_frontend.py:solve_pnp; the VALUES are the reference's, already rounded to
        # the type its Mats hold)
        r0, t0 = rv.copy().reshape(3, 1), tv.copy().reshape(3, 1)
        if self.guess == "carried" and use_guess and self._last_pnp is not None:
            lr, lt = self._last_pnp
            if depth == 5 and np.array_equal(lr.astype(np.float32), r0.astype(np.float32).reshape(3)) and \
                    np.array_equal(lt.astype(np.float32), t0.astype(np.float32).reshape(3)):
                r0, t0 = lr.copy().reshape(3, 1), lt.copy().reshape(3, 1)
        ok, r1, t1, inl = cv2.solvePnPRansac(w.reshape(-1, 1, 3), im.reshape(-1, 1, 2), K,
                                             np.zeros((4, 1), np.float32), r0, t0, bool(use_guess),
                                             iterations, reproj, confidence)
        rv[:] = np.asarray(r1, np.float64).reshape(3)
        tv[:] = np.asarray(t1, np.float64).reshape(3)
        self._last_pnp = (rv.copy(), tv.copy())
        inl = np.zeros(0, np.int32) if inl is None else inl.reshape(-1).astype(np.int32)
        m = min(inl.shape[0], cap)
        np.ctypeslib.as_array(inliers, (max(cap, 1),))[:m] = inl[:m]
        if self.record:
            self.calls.append(("solvePnPRansac", dict(n=n, K=K, depth=depth, use_guess=use_guess,
                                                  iterations=iterations, reproj=reproj, confidence=confidence,
                                                  world=w, image=im, inliers=inl[:m].copy(),
                                                  rvec=rv.copy(), tvec=tv.copy())))
        return m

    def _rodrigues(self, rvec, depth, rmat):
        rv = np.ctypeslib.as_array(rvec, (3,)).astype(_DEPTH[depth]).reshape(3, 1)
        R, _ = cv2.Rodrigues(rv)
        np.ctypeslib.as_array(rmat, (9,))[:] = np.asarray(R, np.float64).reshape(9)
        if self.record:
            self.calls.append(("Rodrigues", dict(depth=depth)))

    def _gemm(self, a, b, depth, m, k, n, alpha, out):
        A = np.ctypeslib.as_array(a, (m, k)).astype(_DEPTH[depth])
        B = np.ctypeslib.as_array(b, (k, n)).astype(_DEPTH[depth])
        np.ctypeslib.as_array(out, (m, n))[:] = cv2.gemm(A, B, alpha, None, 0.0)

    # ---- driving it -----------------------------------------------------------------------
    def process(self, left, right, t=0.0):
        left = np.ascontiguousarray(left, np.uint8)
        right = np.ascontiguousarray(right, np.uint8)
        self.calls = []
        rc = self.lib.ref_vo_process(self.h, left.ctypes.data, right.ctypes.data, left.shape[1], left.shape[0],
                                     left.strides[0], t)
        if rc < 0:
            raise RuntimeError("the reference threw: " + self.lib.ref_vo_last_error().decode(errors="replace"))
        rec = dict(made_keyframe=bool(rc), calls=self.calls)
        if not self.record:                      # timing runs: no read-back, no log
            return rec
        cap = 4096
        xy, ids = np.zeros((cap, 2), np.float32), np.zeros(cap, np.uint64)
        n = self.lib.ref_vo_tracker(self.h, xy.ctypes.data, ids.ctypes.data, cap)
        rec["tracker_pts"], rec["tracker_ids"] = xy[:n].copy(), ids[:n].copy()
        if rc == 1:
            sz = np.zeros(4, np.int64)
            self.lib.ref_vo_keyframe_sizes(self.h, sz.ctypes.data)
            pose = np.zeros(7, np.float32)
            tid, t2d = np.zeros(max(1, sz[0]), np.uint64), np.zeros((max(1, sz[0]), 2), np.float32)
            n2d, n3d = np.zeros((max(1, sz[1]), 2), np.float32), np.zeros((max(1, sz[2]), 3), np.float32)
            nid = np.zeros(max(1, sz[3]), np.uint64)
            self.lib.ref_vo_keyframe(self.h, pose.ctypes.data, tid.ctypes.data, t2d.ctypes.data,
                                     n2d.ctypes.data, n3d.ctypes.data, nid.ctypes.data)
            rec.update(pose7=pose, kf_tracked_ids=tid[:sz[0]], kf_tracked_2d=t2d[:sz[0]], new2d=n2d[:sz[1]],
                       new3d=n3d[:sz[2]], new_ids=nid[:sz[3]])
        self.log.append(rec)
        return rec

    def close(self):
        if self.h:
            self.lib.ref_vo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
