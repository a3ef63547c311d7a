"""ctypes binding of include/svo_c.h (libsvo.so). Test / bench harness plumbing only: numpy
arrays in, numpy arrays out. There is no fallback: if the CUDA library is missing or no GPU is
usable, loading or engine creation raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsvo.so")

SVO_OK = 0
ERR_NAMES = {0: "SVO_OK", -1: "SVO_ERR_INVALID", -2: "SVO_ERR_CUDA", -3: "SVO_ERR_CAPACITY",
             -4: "SVO_ERR_STATE", -5: "SVO_ERR_NOMEM"}


class SvoParams(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32),
        ("max_corners", C.c_int32), ("quality_level", C.c_double), ("min_distance", C.c_float),
        ("max_candidates", C.c_int32),
        ("num_disparities", C.c_int32), ("sad_window", C.c_int32), ("prefilter_cap", C.c_int32),
        ("texture_threshold", C.c_int32), ("uniqueness_ratio", C.c_int32),
        ("lk_max_level", C.c_int32), ("lk_max_iters", C.c_int32), ("lk_epsilon", C.c_double),
        ("lk_min_eig_threshold", C.c_double),
        ("fb_max_error", C.c_float), ("max_parallax", C.c_float), ("max_features", C.c_int32),
        ("keep_eig_map", C.c_int32),
    ]


class SvoTrackStats(C.Structure):
    _fields_ = [("av_parallax", C.c_float), ("percent_lost", C.c_float),
                ("n_kept", C.c_int32), ("n_attempted", C.c_int32)]


PNP_DTYPE = np.dtype([("rvec", np.float64, 3), ("tvec", np.float64, 3), ("rmat", np.float64, 9),
                      ("n_inliers", np.int32), ("ok", np.int32), ("best_hypothesis", np.int32),
                      ("hypotheses_tried", np.int32)])

STATS_DTYPE = np.dtype([("av_parallax", np.float32), ("percent_lost", np.float32),
                        ("n_kept", np.int32), ("n_attempted", np.int32)])

# every symbol include/svo_c.h declares (tests check that the library exports all of them)
EXPORTS = [
    "svo_default_params", "svo_create", "svo_destroy", "svo_last_error", "svo_get_params",
    "svo_num_levels", "svo_kernel_launches", "svo_last_gpu_ms", "svo_last_timing",
    "svo_good_features_to_track", "svo_triangulate_stereo", "svo_stereo_bm_dense",
    "svo_stereo_bm_dense_resident", "svo_tracker_init",
    "svo_tracker_track", "svo_tracker_get", "svo_dedupe", "svo_tracker_debug",
    "svo_debug_min_eigen_map", "svo_debug_pyramid_level", "svo_debug_candidates",
    "svo_step", "svo_prefetch", "svo_step_device", "svo_set_right", "svo_device_can_read", "svo_keyframe", "svo_keyframe_tracked", "svo_step_keyframe", "svo_tracker_set",
    "svo_device_alloc", "svo_device_free", "svo_device_upload", "svo_host_alloc", "svo_host_free",
    "svo_set_profiling", "svo_profile_slots", "svo_profile_name", "svo_get_profile",
    "svo_timer_begin", "svo_timer_end", "svo_pnp_ransac", "svo_reprojection_evaluate", "svo_draw_track", "svo_set_blocking_sync",
    "svo_dedupe_static", "svo_detector_spills", "svo_lk_handover",
]

_lib = None


class SvoError(RuntimeError):
    pass


def load_library(path: Optional[str] = None):
    """dlopen libsvo.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise SvoError(f"{p} is missing: build it with `python -m stereo_vo_b200.build` "
                       "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(p)
    vp, i32, f32 = C.c_void_p, C.c_int32, C.c_float
    lib.svo_default_params.argtypes = [C.POINTER(SvoParams), i32, i32]
    lib.svo_default_params.restype = None
    lib.svo_create.argtypes = [C.POINTER(SvoParams), i32, i32, C.POINTER(vp)]
    lib.svo_destroy.argtypes = [vp]
    lib.svo_destroy.restype = None
    lib.svo_last_error.argtypes = [vp]
    lib.svo_last_error.restype = C.c_char_p
    lib.svo_get_params.argtypes = [vp, C.POINTER(SvoParams)]
    lib.svo_num_levels.argtypes = [vp]
    lib.svo_kernel_launches.argtypes = [vp]
    lib.svo_kernel_launches.restype = C.c_int64
    lib.svo_last_gpu_ms.argtypes = [vp]
    lib.svo_last_gpu_ms.restype = f32
    lib.svo_last_timing.argtypes = [vp, C.POINTER(f32)]
    lib.svo_good_features_to_track.argtypes = [vp, i32, vp, i32, vp, C.POINTER(i32)]
    lib.svo_triangulate_stereo.argtypes = [vp, i32, vp, vp, i32, vp, i32, vp, f32, f32, f32, f32,
                                           vp, vp, vp, vp, C.POINTER(i32)]
    lib.svo_stereo_bm_dense.argtypes = [vp, i32, vp, vp, i32, vp, i32]
    lib.svo_stereo_bm_dense_resident.argtypes = [vp, i32, vp, vp, i32]
    lib.svo_tracker_init.argtypes = [vp, i32, vp, i32, vp, vp, i32]
    lib.svo_tracker_track.argtypes = [vp, i32, vp, i32, i32, C.POINTER(f32), C.POINTER(f32)]
    lib.svo_tracker_get.argtypes = [vp, i32, vp, vp, C.POINTER(i32)]
    lib.svo_dedupe.argtypes = [vp, i32, vp, i32, vp, i32, f32, vp, vp, C.POINTER(i32)]
    lib.svo_tracker_debug.argtypes = [vp, i32, vp, vp, vp, vp, vp, C.POINTER(i32)]
    lib.svo_debug_min_eigen_map.argtypes = [vp, i32, vp]
    lib.svo_debug_pyramid_level.argtypes = [vp, i32, i32, vp, C.POINTER(i32), C.POINTER(i32)]
    lib.svo_debug_candidates.argtypes = [vp, i32, vp, i32, C.POINTER(i32)]
    lib.svo_step.argtypes = [vp, i32, vp, vp, vp, i32, vp, vp, vp, vp, vp]
    lib.svo_step_device.argtypes = [vp, i32, vp, vp, vp, i32, vp, vp, vp, vp, vp]
    lib.svo_prefetch.argtypes = [vp, i32, vp, vp, vp, i32]
    lib.svo_set_right.argtypes = [vp, i32, vp, vp, i32, i32]
    lib.svo_device_can_read.argtypes = [vp]
    lib.svo_keyframe.argtypes = [vp, i32, vp, vp, vp, vp, f32, f32, f32, f32, vp, vp, vp, vp, vp,
                                 vp, vp]
    lib.svo_keyframe_tracked.argtypes = [vp, i32, vp, vp, vp, f32, f32, f32, f32, vp, vp, vp, vp,
                                         vp, vp, vp]
    lib.svo_step_keyframe.argtypes = [vp, i32, vp, vp, vp, i32, i32, vp, vp, f32, f32, f32, f32, f32,
                                      vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.svo_tracker_set.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.svo_reprojection_evaluate.argtypes = [vp, i32, vp, vp, vp, i32, vp, i32, vp, C.c_double,
                                              C.c_double, C.c_double, vp, vp, vp]
    lib.svo_set_blocking_sync.argtypes = [vp, i32]
    lib.svo_draw_track.argtypes = [vp, vp, i32, vp, vp, i32, vp, i32]
    lib.svo_pnp_ransac.argtypes = [vp, i32, vp, vp, vp, f32, f32, f32, i32, f32, C.c_double, vp, vp, vp]
    lib.svo_detector_spills.argtypes = [vp]
    lib.svo_detector_spills.restype = C.c_int64
    lib.svo_lk_handover.argtypes = [vp, i32, i32, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    lib.svo_dedupe_static.argtypes = [i32, vp, i32, vp, i32, f32, vp, vp, vp, vp, i32, C.c_char_p,
                                      C.c_size_t]
    lib.svo_device_alloc.argtypes = [vp, C.c_size_t, C.POINTER(vp)]
    lib.svo_device_free.argtypes = [vp, vp]
    lib.svo_device_upload.argtypes = [vp, vp, vp, C.c_size_t]
    lib.svo_host_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
    lib.svo_host_free.argtypes = [vp]
    lib.svo_set_profiling.argtypes = [vp, i32]
    lib.svo_profile_slots.argtypes = []
    lib.svo_profile_name.argtypes = [i32]
    lib.svo_profile_name.restype = C.c_char_p
    lib.svo_get_profile.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    lib.svo_timer_begin.argtypes = [vp]
    lib.svo_timer_end.argtypes = [vp, C.POINTER(f32)]
    if path is None:
        _lib = lib
    return lib


def default_params(width: int, height: int) -> SvoParams:
    p = SvoParams()
    load_library().svo_default_params(C.byref(p), width, height)
    return p


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _img(a: np.ndarray) -> np.ndarray:
    if a.dtype != np.uint8 or a.ndim != 2:
        raise ValueError("image must be a 2-D uint8 array")
    if a.strides[1] != 1:
        a = np.ascontiguousarray(a)
    return a


def pinned_empty(shape, dtype=np.uint8) -> np.ndarray:
    """numpy view on page-locked host memory (freed when the array is garbage collected)."""
    lib = load_library()
    nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = C.c_void_p()
    if lib.svo_host_alloc(max(nbytes, 1), C.byref(p)) != 0:
        raise SvoError("svo_host_alloc failed")
    buf = (C.c_uint8 * max(nbytes, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[id(buf)] = (buf, p)
    return arr


_PINNED: dict = {}

DEDUPE_STATIC_MAX_DET = 16384      # SVO_DEDUPE_STATIC_MAX_DET of include/svo_c.h


def dedupe_static(det, n_det: int, tracked, n_tracked: int, min_distance: float, out_xy, out_keep,
                  out_n, device: int = 0, stream: int = 0, wait: bool = True) -> None:
    """svo_dedupe_static: the allocation-free variant of svo_dedupe. Every array argument is a
    numpy view on PINNED memory (pinned_empty) or an integer device address; nothing is staged
    or allocated by the call. Raises SvoError with the library's message on failure."""
    lib = load_library()

    def addr(a):
        if a is None:
            return None
        return C.c_void_p(a) if isinstance(a, int) else _ptr(a)
    err = C.create_string_buffer(256)
    rc = lib.svo_dedupe_static(device, addr(det), n_det, addr(tracked), n_tracked,
                               float(min_distance), addr(out_xy), addr(out_keep), addr(out_n),
                               C.c_void_p(stream) if stream else None, 1 if wait else 0, err, 256)
    if rc != 0:
        e = SvoError(err.value.decode(errors="replace") or f"svo_dedupe_static failed ({rc})")
        e.code = rc
        raise e


class Engine:
    """One svo_engine: `n_seq` independent stereo sequences on one device / stream."""

    def __init__(self, width: int, height: int, n_seq: int = 1, device: int = 0,
                 params: Optional[SvoParams] = None, **overrides):
        self.lib = load_library()
        p = params or default_params(width, height)
        p.width, p.height = width, height
        for k, v in overrides.items():
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
        h = C.c_void_p()
        rc = self.lib.svo_create(C.byref(p), device, n_seq, C.byref(h))
        if rc != 0:
            msg = self.lib.svo_last_error(None)
            raise SvoError(f"svo_create failed ({ERR_NAMES.get(rc, rc)}): "
                           f"{msg.decode() if msg else ''}")
        self.h = h
        self.params = SvoParams()
        self.lib.svo_get_params(self.h, C.byref(self.params))
        self.n_seq = n_seq
        self.width, self.height = width, height
        self.max_corners = self.params.max_corners
        self.max_features = self.params.max_features
        self.cap = max(self.max_corners, self.max_features)

    def close(self):
        if getattr(self, "h", None):
            self.lib.svo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc: int, what: str):
        if rc != 0:
            msg = self.lib.svo_last_error(self.h)
            raise SvoError(f"{what} failed ({ERR_NAMES.get(rc, rc)}): {msg.decode() if msg else ''}")

    # -- introspection ------------------------------------------------------------------
    @property
    def num_levels(self) -> int:
        return self.lib.svo_num_levels(self.h)

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.svo_kernel_launches(self.h))

    @property
    def detector_spills(self) -> int:
        """detections repeated on the spill lists (plateau images), svo_detector_spills"""
        return int(self.lib.svo_detector_spills(self.h))

    def lk_handover(self, enable: int = -1, count: int = -1):
        """svo_lk_handover: switch the backward -> forward hand-over of lk_track / its counters;
        returns (look-ups, hits) of the forward passes since counting was switched on"""
        a, b = C.c_int64(0), C.c_int64(0)
        self._check(self.lib.svo_lk_handover(self.h, enable, count, C.byref(a), C.byref(b)),
                    "svo_lk_handover")
        return int(a.value), int(b.value)

    def last_timing(self):
        ms = (C.c_float * 3)()
        self.lib.svo_last_timing(self.h, ms)
        return float(ms[0]), float(ms[1]), float(ms[2])

    def set_profiling(self, on: bool):
        self._check(self.lib.svo_set_profiling(self.h, 1 if on else 0), "svo_set_profiling")

    def get_profile(self) -> dict:
        n = self.lib.svo_profile_slots()
        ms = (C.c_double * n)()
        cnt = (C.c_int64 * n)()
        self._check(self.lib.svo_get_profile(self.h, ms, cnt), "svo_get_profile")
        return {self.lib.svo_profile_name(i).decode(): (float(ms[i]), int(cnt[i]))
                for i in range(n)}

    def timer_begin(self):
        self._check(self.lib.svo_timer_begin(self.h), "svo_timer_begin")

    def timer_end(self) -> float:
        ms = C.c_float()
        self._check(self.lib.svo_timer_end(self.h, C.byref(ms)), "svo_timer_end")
        return float(ms.value)

    # -- stage calls ----------------------------------------------------------------------
    def good_features_to_track(self, img: np.ndarray, seq: int = 0) -> np.ndarray:
        img = _img(img)
        out = np.zeros((self.max_corners, 2), np.float32)
        n = C.c_int32()
        self._check(self.lib.svo_good_features_to_track(self.h, seq, _ptr(img), img.strides[0],
                                                        _ptr(out), C.byref(n)),
                    "svo_good_features_to_track")
        return out[:n.value].copy()

    def triangulate_stereo(self, left, right, feats, pose, focal, cx, cy, baseline, seq: int = 0):
        left, right = _img(left), _img(right)
        if left.strides[0] != right.strides[0]:
            right = np.ascontiguousarray(right)
            left = np.ascontiguousarray(left)
        feats = np.ascontiguousarray(feats, np.float32).reshape(-1, 2)
        n = feats.shape[0]
        pose = np.ascontiguousarray(pose, np.float32).reshape(16)
        o3 = np.zeros((max(n, 1), 3), np.float32)
        o2 = np.zeros((max(n, 1), 2), np.float32)
        oi = np.zeros(max(n, 1), np.int32)
        od = np.zeros(max(n, 1), np.int16)
        nv = C.c_int32()
        self._check(self.lib.svo_triangulate_stereo(
            self.h, seq, _ptr(left), _ptr(right), left.strides[0], _ptr(feats), n, _ptr(pose),
            float(focal), float(cx), float(cy), float(baseline), _ptr(o3), _ptr(o2), _ptr(oi),
            _ptr(od), C.byref(nv)), "svo_triangulate_stereo")
        k = nv.value
        return o3[:k].copy(), o2[:k].copy(), oi[:k].copy(), od[:n].copy()

    def stereo_bm_dense(self, left, right, seq: int = 0) -> np.ndarray:
        """svo_stereo_bm_dense: the full CV_16S map of StereoBM(48, 21)->compute(left, right)."""
        left, right = _img(left), _img(right)
        if left.strides[0] != right.strides[0]:
            right = np.ascontiguousarray(right)
            left = np.ascontiguousarray(left)
        out = np.empty(left.shape, np.int16)
        self._check(self.lib.svo_stereo_bm_dense(self.h, seq, _ptr(left), _ptr(right),
                                                 left.strides[0], _ptr(out), out.shape[1]),
                    "svo_stereo_bm_dense")
        return out

    def stereo_bm_dense_resident(self, seqs, shape=None, fetch: bool = True):
        """svo_stereo_bm_dense_resident on the stereo pairs already on the device; `shape` =
        (height, width) of the engine's images. fetch=False leaves the maps on the device."""
        seqs_a = np.ascontiguousarray(seqs, np.int32)
        n = len(seqs_a)
        if not fetch:
            self._check(self.lib.svo_stereo_bm_dense_resident(self.h, n, _ptr(seqs_a), None, 0),
                        "svo_stereo_bm_dense_resident")
            return None
        outs = [np.empty(shape, np.int16) for _ in range(n)]
        self._check(self.lib.svo_stereo_bm_dense_resident(self.h, n, _ptr(seqs_a),
                                                          self._ptr_array(outs), shape[1]),
                    "svo_stereo_bm_dense_resident")
        return outs

    def tracker_init(self, img, feats, ids, seq: int = 0):
        img = _img(img)
        feats = np.ascontiguousarray(feats, np.float32).reshape(-1, 2)
        ids = np.ascontiguousarray(ids, np.uint64).reshape(-1)
        self._check(self.lib.svo_tracker_init(self.h, seq, _ptr(img), img.strides[0], _ptr(feats),
                                              _ptr(ids), feats.shape[0]), "svo_tracker_init")

    def tracker_track(self, img, flow_back: bool = True, seq: int = 0):
        img = _img(img)
        av, lost = C.c_float(), C.c_float()
        self._check(self.lib.svo_tracker_track(self.h, seq, _ptr(img), img.strides[0],
                                               1 if flow_back else 0, C.byref(av), C.byref(lost)),
                    "svo_tracker_track")
        return np.float32(av.value), np.float32(lost.value)

    def tracker_get(self, seq: int = 0):
        feats = np.zeros((self.max_features, 2), np.float32)
        ids = np.zeros(self.max_features, np.uint64)
        n = C.c_int32()
        self._check(self.lib.svo_tracker_get(self.h, seq, _ptr(feats), _ptr(ids), C.byref(n)),
                    "svo_tracker_get")
        return feats[:n.value].copy(), ids[:n.value].copy()

    def tracker_debug(self, seq: int = 0):
        mf = self.max_features
        tracked = np.zeros((mf, 2), np.float32)
        reverse = np.zeros((mf, 2), np.float32)
        s1 = np.zeros(mf, np.uint8)
        s2 = np.zeros(mf, np.uint8)
        kept = np.zeros(mf, np.int32)
        n = C.c_int32()
        self._check(self.lib.svo_tracker_debug(self.h, seq, _ptr(tracked), _ptr(s1), _ptr(reverse),
                                               _ptr(s2), _ptr(kept), C.byref(n)),
                    "svo_tracker_debug")
        k = n.value
        return dict(tracked=tracked[:k], status1=s1[:k], reverse=reverse[:k], status2=s2[:k],
                    kept_idx=kept)

    def dedupe(self, det, tracked, min_distance: float, seq: int = 0):
        det = np.ascontiguousarray(det, np.float32).reshape(-1, 2)
        tracked = np.ascontiguousarray(tracked, np.float32).reshape(-1, 2)
        out = np.zeros((max(det.shape[0], 1), 2), np.float32)
        keep = np.zeros(max(det.shape[0], 1), np.int32)
        n = C.c_int32()
        self._check(self.lib.svo_dedupe(self.h, seq, _ptr(det), det.shape[0], _ptr(tracked),
                                        tracked.shape[0], float(min_distance), _ptr(out),
                                        _ptr(keep), C.byref(n)), "svo_dedupe")
        return out[:n.value].copy(), keep[:n.value].copy()

    # -- taps -------------------------------------------------------------------------------
    def debug_min_eigen_map(self, seq: int = 0) -> np.ndarray:
        out = np.zeros((self.height, self.width), np.float32)
        self._check(self.lib.svo_debug_min_eigen_map(self.h, seq, _ptr(out)),
                    "svo_debug_min_eigen_map")
        return out

    def debug_pyramid_level(self, level: int, seq: int = 0) -> np.ndarray:
        w, h = C.c_int32(), C.c_int32()
        self._check(self.lib.svo_debug_pyramid_level(self.h, seq, level, None, C.byref(w),
                                                     C.byref(h)), "svo_debug_pyramid_level")
        out = np.zeros((h.value, w.value), np.uint8)
        self._check(self.lib.svo_debug_pyramid_level(self.h, seq, level, _ptr(out), C.byref(w),
                                                     C.byref(h)), "svo_debug_pyramid_level")
        return out

    def debug_candidates(self, seq: int = 0, cap: int = 1 << 20) -> np.ndarray:
        n = C.c_int32()
        self._check(self.lib.svo_debug_candidates(self.h, seq, None, 0, C.byref(n)),
                    "svo_debug_candidates")
        m = min(n.value, cap)
        xy = np.zeros((max(m, 1), 2), np.int32)
        self._check(self.lib.svo_debug_candidates(self.h, seq, _ptr(xy), m, C.byref(n)),
                    "svo_debug_candidates")
        return xy[:m]

    # -- fused batched path ------------------------------------------------------------------
    def _ptr_array(self, imgs: Sequence[np.ndarray]):
        arr = (C.c_void_p * len(imgs))()
        for i, im in enumerate(imgs):
            arr[i] = im.ctypes.data if isinstance(im, np.ndarray) else int(im)
        return arr

    def step(self, seqs, left, right, pitch: Optional[int] = None, device: bool = False,
             want_tracked: bool = True):
        """svo_step / svo_step_device. `left`/`right`: lists of uint8 images (or raw device
        pointers when device=True). Returns dict of per-listed-sequence results."""
        n = len(seqs)
        seqs_a = np.ascontiguousarray(seqs, np.int32)
        if not device:
            left = [_img(a) for a in left]
            right = [_img(a) for a in right]
            pitch = left[0].strides[0]
            for a in list(left) + list(right):
                if a.strides[0] != pitch:
                    raise ValueError("all images of a step must share one pitch")
        lp, rp = self._ptr_array(left), self._ptr_array(right)
        corners = np.zeros((n, self.max_corners, 2), np.float32)
        ncorn = np.zeros(n, np.int32)
        stats = np.zeros(n, STATS_DTYPE)
        tracked = np.zeros((n, self.max_features, 2), np.float32) if want_tracked else None
        tids = np.zeros((n, self.max_features), np.uint64) if want_tracked else None
        fn = self.lib.svo_step_device if device else self.lib.svo_step
        self._check(fn(self.h, n, _ptr(seqs_a), lp, rp, int(pitch), _ptr(corners), _ptr(ncorn),
                       _ptr(stats), _ptr(tracked), _ptr(tids)), "svo_step")
        return dict(corners=corners, n_corners=ncorn, stats=stats, tracked=tracked,
                    tracked_ids=tids)

    def prefetch(self, seqs, left, right):
        """svo_prefetch: announce the host frames of an upcoming step() (same arrays)."""
        seqs_a = np.ascontiguousarray(seqs, np.int32)
        left = [_img(a) for a in left]
        right = [_img(a) for a in right]
        pitch = left[0].strides[0]
        lp, rp = self._ptr_array(left), self._ptr_array(right)
        self._check(self.lib.svo_prefetch(self.h, len(seqs), _ptr(seqs_a), lp, rp, int(pitch)),
                    "svo_prefetch")

    def keyframe(self, seqs, inliers, n_inliers, poses, focal, cx, cy, baseline):
        n = len(seqs)
        seqs_a = np.ascontiguousarray(seqs, np.int32)
        inl = np.zeros((n, self.max_features, 2), np.float32)
        for i, a in enumerate(inliers):
            a = np.asarray(a, np.float32).reshape(-1, 2)
            inl[i, :a.shape[0]] = a
        ninl = np.ascontiguousarray(n_inliers, np.int32)
        poses = np.ascontiguousarray(poses, np.float32).reshape(n, 16)
        mc = self.max_corners
        new = np.zeros((n, mc, 2), np.float32)
        nnew = np.zeros(n, np.int32)
        v2 = np.zeros((n, mc, 2), np.float32)
        p3 = np.zeros((n, mc, 3), np.float32)
        vi = np.zeros((n, mc), np.int32)
        d16 = np.zeros((n, mc), np.int16)
        nv = np.zeros(n, np.int32)
        self._check(self.lib.svo_keyframe(self.h, n, _ptr(seqs_a), _ptr(inl), _ptr(ninl),
                                          _ptr(poses), float(focal), float(cx), float(cy),
                                          float(baseline), _ptr(new), _ptr(nnew), _ptr(v2),
                                          _ptr(p3), _ptr(vi), _ptr(d16), _ptr(nv)), "svo_keyframe")
        return dict(new_feats=new, n_new=nnew, valid_2d=v2, pts_3d=p3, valid_idx=vi, disp16=d16,
                    n_valid=nv)

    def keyframe_tracked(self, seqs, poses, first_new_id, focal, cx, cy, baseline):
        """svo_keyframe_tracked: every tracked feature is an inlier, tracker re-initialised on
        the device (see include/svo_c.h)."""
        n = len(seqs)
        seqs_a = np.ascontiguousarray(seqs, np.int32)
        poses = np.ascontiguousarray(poses, np.float32).reshape(n, 16)
        first = np.ascontiguousarray(first_new_id, np.uint64)
        mc = self.max_corners
        new = np.zeros((n, mc, 2), np.float32)
        nnew = np.zeros(n, np.int32)
        v2 = np.zeros((n, mc, 2), np.float32)
        p3 = np.zeros((n, mc, 3), np.float32)
        vi = np.zeros((n, mc), np.int32)
        d16 = np.zeros((n, mc), np.int16)
        nv = np.zeros(n, np.int32)
        self._check(self.lib.svo_keyframe_tracked(self.h, n, _ptr(seqs_a), _ptr(poses), _ptr(first),
                                                  float(focal), float(cx), float(cy),
                                                  float(baseline), _ptr(new), _ptr(nnew), _ptr(v2),
                                                  _ptr(p3), _ptr(vi), _ptr(d16), _ptr(nv)),
                    "svo_keyframe_tracked")
        return dict(new_feats=new, n_new=nnew, valid_2d=v2, pts_3d=p3, valid_idx=vi, disp16=d16,
                    n_valid=nv)

    def step_keyframe(self, seqs, left, right, poses, first_new_id, parallax_thresh, focal, cx, cy,
                      baseline):
        """svo_step_keyframe: svo_step + device-side decision + keyframe stage in one call.
        `right` may hold None for sequences whose right image is not sent."""
        n = len(seqs)
        seqs_a = np.ascontiguousarray(seqs, np.int32)
        left = [_img(a) for a in left]
        right = [None if a is None else _img(a) for a in right]
        pitch = left[0].strides[0]
        lp = self._ptr_array(left)
        rp = (C.c_void_p * n)(*[None if a is None else a.ctypes.data for a in right])
        poses = np.ascontiguousarray(poses, np.float32).reshape(n, 16)
        first = np.ascontiguousarray(first_new_id, np.uint64)
        mc, mf = self.max_corners, self.max_features
        corners = np.zeros((n, mc, 2), np.float32)
        ncorn = np.zeros(n, np.int32)
        stats = np.zeros(n, STATS_DTYPE)
        tracked = np.zeros((n, mf, 2), np.float32)
        tids = np.zeros((n, mf), np.uint64)
        dec = np.zeros(n, np.int32)
        new = np.zeros((n, mc, 2), np.float32)
        nnew = np.zeros(n, np.int32)
        v2 = np.zeros((n, mc, 2), np.float32)
        p3 = np.zeros((n, mc, 3), np.float32)
        vi = np.zeros((n, mc), np.int32)
        d16 = np.zeros((n, mc), np.int16)
        nv = np.zeros(n, np.int32)
        self._check(self.lib.svo_step_keyframe(
            self.h, n, _ptr(seqs_a), lp, rp, int(pitch), 0, _ptr(poses), _ptr(first),
            float(parallax_thresh), float(focal), float(cx), float(cy), float(baseline),
            _ptr(corners), _ptr(ncorn), _ptr(stats), _ptr(tracked), _ptr(tids), _ptr(dec), _ptr(new),
            _ptr(nnew), _ptr(v2), _ptr(p3), _ptr(vi), _ptr(d16), _ptr(nv)), "svo_step_keyframe")
        return dict(corners=corners, n_corners=ncorn, stats=stats, tracked=tracked,
                    tracked_ids=tids, decision=dec, new_feats=new, n_new=nnew, valid_2d=v2,
                    pts_3d=p3, valid_idx=vi, disp16=d16, n_valid=nv)

    def tracker_set(self, seqs, feats, ids):
        n = len(seqs)
        seqs_a = np.ascontiguousarray(seqs, np.int32)
        f = np.zeros((n, self.max_features, 2), np.float32)
        d = np.zeros((n, self.max_features), np.uint64)
        cnt = np.zeros(n, np.int32)
        for i in range(n):
            a = np.asarray(feats[i], np.float32).reshape(-1, 2)
            f[i, :a.shape[0]] = a
            d[i, :a.shape[0]] = np.asarray(ids[i], np.uint64).reshape(-1)
            cnt[i] = a.shape[0]
        self._check(self.lib.svo_tracker_set(self.h, n, _ptr(seqs_a), _ptr(f), _ptr(d), _ptr(cnt)),
                    "svo_tracker_set")

    # -- device-resident frames (bench kernel-only leg) --------------------------------------
    def pnp_ransac(self, world, image, focal, cx, cy, guess=None, iterations=100,
                   reproj_error=8.0, confidence=0.99):
        """svo_pnp_ransac on a list of point sets: world[i] (n_i, 3), image[i] (n_i, 2);
        guess (n, 6) rvec|tvec or None for zeros. Returns (results PNP_DTYPE[n], [inlier idx])."""
        n, mf = len(world), self.max_features
        w = np.zeros((n, mf, 3), np.float32)
        im = np.zeros((n, mf, 2), np.float32)
        counts = np.zeros(n, np.int32)
        for i in range(n):
            k = len(world[i])
            counts[i] = k
            w[i, :k] = np.asarray(world[i], np.float32).reshape(-1, 3)
            im[i, :k] = np.asarray(image[i], np.float32).reshape(-1, 2)
        g = np.zeros((n, 6), np.float64) if guess is None else \
            np.ascontiguousarray(guess, np.float64).reshape(n, 6)
        res = np.zeros(n, PNP_DTYPE)
        inl = np.zeros((n, mf), np.int32)
        self._check(self.lib.svo_pnp_ransac(self.h, n, _ptr(w), _ptr(im), _ptr(counts), float(focal),
                                            float(cx), float(cy), int(iterations),
                                            float(reproj_error), float(confidence), _ptr(g),
                                            _ptr(res), _ptr(inl)), "svo_pnp_ransac")
        return res, [inl[i, :res[i]["n_inliers"]].copy() for i in range(n)]

    def reprojection_evaluate(self, poses, points, pose_idx, point_idx, obs, focal, cx, cy):
        """svo_reprojection_evaluate: (residuals (n, 2), jac_pose (n, 2, 7), jac_point (n, 2, 3))"""
        poses = np.ascontiguousarray(poses, np.float64).reshape(-1, 7)
        points = np.ascontiguousarray(points, np.float64).reshape(-1, 3)
        pi = np.ascontiguousarray(pose_idx, np.int32)
        xi = np.ascontiguousarray(point_idx, np.int32)
        obs = np.ascontiguousarray(obs, np.float64).reshape(-1, 2)
        n = len(pi)
        r = np.zeros((n, 2))
        jp = np.zeros((n, 2, 7))
        jx = np.zeros((n, 2, 3))
        self._check(self.lib.svo_reprojection_evaluate(
            self.h, n, _ptr(pi), _ptr(xi), _ptr(obs), len(poses), _ptr(poses), len(points),
            _ptr(points), float(focal), float(cx), float(cy), _ptr(r), _ptr(jp), _ptr(jx)),
            "svo_reprojection_evaluate")
        return r, jp, jx

    def draw_track(self, image, from_pts, to_pts) -> np.ndarray:
        """svo_draw_track: (H, W, 3) RGB overlay of FeatureTracker::draw_track."""
        image = _img(image)
        a = np.ascontiguousarray(from_pts, np.float32).reshape(-1, 2)
        b = np.ascontiguousarray(to_pts, np.float32).reshape(-1, 2)
        out = np.zeros((self.height, self.width, 3), np.uint8)
        self._check(self.lib.svo_draw_track(self.h, _ptr(image), image.strides[0], _ptr(a), _ptr(b),
                                            len(a), _ptr(out), out.strides[0]), "svo_draw_track")
        return out

    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._check(self.lib.svo_device_alloc(self.h, nbytes, C.byref(p)), "svo_device_alloc")
        return p.value

    def device_free(self, ptr: int):
        self._check(self.lib.svo_device_free(self.h, C.c_void_p(ptr)), "svo_device_free")

    def device_upload(self, dst: int, src: np.ndarray):
        src = np.ascontiguousarray(src)
        self._check(self.lib.svo_device_upload(self.h, C.c_void_p(dst), _ptr(src), src.nbytes),
                    "svo_device_upload")
