"""ctypes binding of libsvo_host.so (include/svo_host_c.h): the C++ host layer - the
reference-shaped ImageProcessor / FeatureTracker facade and the multi-sequence BatchFrontEnd -
for the Python parity tests and bench.py. No computation happens in this file."""
from __future__ import annotations

import ctypes as C
import os
import re
from typing import Callable, Optional

import numpy as np

from . import capi
from .capi import STATS_DTYPE, SvoError, SvoParams

HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(HERE, "libsvo_host.so")
HEADER = os.path.join(HERE, "..", "include", "svo_host_c.h")

DECISIONS = {0: None, 1: "too_few", 2: "first", 3: "skip", 4: "keyframe"}

PNP_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_float),
                     C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_double),
                     C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_int32))


class VoRecord(C.Structure):
    _fields_ = [
        ("decision", C.c_int32),
        ("n_detected", C.c_int32), ("detected", C.c_void_p),
        ("av_parallax", C.c_float), ("percent_lost", C.c_float),
        ("n_tracked", C.c_int32), ("tracked", C.c_void_p), ("tracked_ids", C.c_void_p),
        ("n_inliers", C.c_int32), ("inliers", C.c_void_p),
        ("hmat", C.c_float * 16),
        ("n_new_features", C.c_int32), ("new_features", C.c_void_p), ("disp16", C.c_void_p),
        ("n_new", C.c_int32), ("valid_idx", C.c_void_p), ("new_2d", C.c_void_p),
        ("new_3d", C.c_void_p), ("new_ids", C.c_void_p),
        ("n_tracker", C.c_int32), ("tracker_pts", C.c_void_p), ("tracker_ids", C.c_void_p),
    ]


class BatchSummary(C.Structure):
    _fields_ = [("frames", C.c_int64), ("keyframes", C.c_int64), ("corners", C.c_int64),
                ("tracked", C.c_int64), ("new_points", C.c_int64), ("right_images", C.c_int64)]


class ImageMsg(C.Structure):
    """svo_image_msg: the fields of sensor_msgs/Image that cv_bridge reads"""
    _fields_ = [("stamp", C.c_double), ("height", C.c_uint32), ("width", C.c_uint32),
                ("encoding", C.c_char_p), ("is_bigendian", C.c_uint8), ("step", C.c_uint32),
                ("data", C.c_void_p), ("size", C.c_size_t)]


PAIR_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                      C.c_int32, C.c_double)


def declared_symbols():
    """Every SVO_API function svo_host_c.h declares."""
    with open(HEADER) as fh:
        text = fh.read()
    return re.findall(r"SVO_API\s+[\w\s\*]+?\b(svo_\w+)\s*\(", text)


_lib = None


def load_library():
    """dlopen libsvo_host.so (which pulls in libsvo.so); raises when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(HOST_LIB_PATH):
        raise SvoError(f"{HOST_LIB_PATH} is missing: build it with `python -m stereo_vo_b200.build`")
    capi.load_library()
    lib = C.CDLL(HOST_LIB_PATH)
    vp, i32, f32 = C.c_void_p, C.c_int32, C.c_float
    pf = C.POINTER(f32)
    lib.svo_camera_from_yaml.argtypes = [C.c_char_p, pf, pf, pf, pf, pf]
    lib.svo_vo_create.argtypes = [f32, f32, f32, f32, f32, f32, i32, C.POINTER(vp)]
    lib.svo_vo_destroy.argtypes = [vp]
    lib.svo_vo_destroy.restype = None
    lib.svo_vo_set_pnp.argtypes = [vp, PNP_FN, vp]
    lib.svo_vo_process.argtypes = [vp, vp, vp, i32, i32, i32, C.c_double]
    lib.svo_vo_last_record.argtypes = [vp, C.POINTER(VoRecord)]
    lib.svo_vo_get_drawing.argtypes = [vp, vp, i32, C.POINTER(i32), C.POINTER(i32)]
    lib.svo_vo_last_error.argtypes = [vp]
    lib.svo_vo_last_error.restype = C.c_char_p
    lib.svo_batch_create.argtypes = [C.POINTER(SvoParams), i32, i32, i32, f32, f32, f32, f32, f32,
                                     C.POINTER(vp)]
    lib.svo_batch_destroy.argtypes = [vp]
    lib.svo_batch_destroy.restype = None
    lib.svo_batch_last_error.argtypes = [vp]
    lib.svo_batch_last_error.restype = C.c_char_p
    lib.svo_batch_step.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp, C.POINTER(BatchSummary)]
    lib.svo_batch_submit.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp]
    lib.svo_batch_wait.argtypes = [vp, C.POINTER(BatchSummary)]
    lib.svo_batch_get.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32), vp, vp, C.POINTER(i32),
                                  vp, vp, C.POINTER(i32), vp, vp]
    lib.svo_batch_get_keyframe.argtypes = [vp, i32, C.POINTER(i32), vp, vp, C.POINTER(i32), vp]
    lib.svo_batch_set_closed_loop.argtypes = [vp, i32, i32, f32, C.c_double]
    lib.svo_batch_get_pose.argtypes = [vp, i32, C.POINTER(f32), C.POINTER(i32)]
    lib.svo_batch_set_blocking_sync.argtypes = [vp, i32]
    lib.svo_batch_kernel_launches.argtypes = [vp]
    i64 = C.c_int64
    lib.svo_kitti_open.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(vp)]
    lib.svo_kitti_close.argtypes = [vp]
    lib.svo_kitti_close.restype = None
    lib.svo_kitti_last_error.argtypes = [vp]
    lib.svo_kitti_last_error.restype = C.c_char_p
    lib.svo_kitti_info.argtypes = [vp, C.POINTER(i64), C.POINTER(i32), C.POINTER(i32)]
    lib.svo_kitti_read.argtypes = [vp, i64, vp, vp, i32, vp]
    lib.svo_kitti_prefetch_start.argtypes = [vp, vp, i32, i32, i64]
    lib.svo_kitti_prefetch_start_threads.argtypes = [vp, vp, i32, i32, i64, i32]
    lib.svo_kitti_next.argtypes = [vp, C.POINTER(i64), C.POINTER(vp), C.POINTER(vp), vp]
    lib.svo_batch_kernel_launches.restype = C.c_int64
    lib.svo_batch_timer_begin.argtypes = [vp]
    lib.svo_batch_timer_end.argtypes = [vp, pf]
    lib.svo_batch_set_profiling.argtypes = [vp, i32]
    lib.svo_batch_get_profile.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    lib.svo_batch_engine.argtypes = [vp, i32]
    lib.svo_batch_engine.restype = vp
    lib.svo_ba_create.argtypes = [i32, C.POINTER(vp)]
    lib.svo_ba_destroy.argtypes = [vp]
    lib.svo_ba_destroy.restype = None
    lib.svo_ba_last_error.argtypes = [vp]
    lib.svo_ba_last_error.restype = C.c_char_p
    lib.svo_ba_add_keyframe.argtypes = [vp, vp, i32, vp, i32, vp, C.POINTER(i32)]
    lib.svo_ba_get_world_points.argtypes = [vp, vp, i32, vp]
    lib.svo_ba_num_features.argtypes = [vp]
    lib.svo_ba_num_features.restype = C.c_int64
    lib.svo_image_to_mono8.argtypes = [C.POINTER(ImageMsg), vp, i32]
    lib.svo_image_last_error.argtypes = []
    lib.svo_image_last_error.restype = C.c_char_p
    lib.svo_stereo_queue_create.argtypes = [C.c_double, C.POINTER(vp)]
    lib.svo_stereo_queue_destroy.argtypes = [vp]
    lib.svo_stereo_queue_destroy.restype = None
    lib.svo_stereo_queue_last_error.argtypes = [vp]
    lib.svo_stereo_queue_last_error.restype = C.c_char_p
    lib.svo_stereo_queue_push.argtypes = [vp, C.POINTER(ImageMsg), C.POINTER(ImageMsg), C.POINTER(i32)]
    lib.svo_stereo_queue_size.argtypes = [vp]
    lib.svo_stereo_queue_drain.argtypes = [vp, PAIR_FN, vp, C.POINTER(i32)]
    lib.svo_vo_process_queue.argtypes = [vp, vp, C.POINTER(i32)]
    _lib = lib
    return lib


def camera_from_yaml(text: str):
    """(focal, cx, cy, baseline, frame_rate) as the float32 values vo_node would hold."""
    lib = load_library()
    v = [C.c_float() for _ in range(5)]
    rc = lib.svo_camera_from_yaml(text.encode(), *[C.byref(x) for x in v])
    if rc != 0:
        raise SvoError(f"svo_camera_from_yaml failed ({rc}): "
                       f"{lib.svo_vo_last_error(None).decode(errors='replace')}")
    return tuple(np.float32(x.value) for x in v)


def _arr(ptr, n, dtype, cols=None):
    if not ptr or n <= 0:
        shape = (0,) if cols is None else (0, cols)
        return np.zeros(shape, dtype)
    count = n * (cols or 1)
    buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr)
    a = np.frombuffer(buf, dtype=dtype, count=count).copy()
    return a if cols is None else a.reshape(n, cols)


class VoSequence:
    """One stereo sequence through the C++ ImageProcessor facade (svo_vo_*)."""

    def __init__(self, focal, cx, cy, baseline, min_feature_distance=30.0, parallax_thresh=20.0,
                 device=0, pnp: Optional[Callable] = None):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.svo_vo_create(float(focal), float(cx), float(cy), float(baseline),
                                    float(min_feature_distance), float(parallax_thresh), device,
                                    C.byref(h))
        if rc != 0:
            raise SvoError(f"svo_vo_create failed ({rc})")
        self.h = h
        self._cb = None
        self.log = []
        if pnp is not None:
            self.set_pnp(pnp)

    def set_pnp(self, fn: Callable):
        """fn(world (n,3) f32, image (n,2) f32, K (3,3) f32, rvec (3,) f64, tvec (3,) f64) ->
        (rvec, tvec, rmat (3,3), inlier indices)."""
        def trampoline(_user, n, world, image, k, rvec, tvec, rmat, inliers, n_inl):
            try:
                w = np.ctypeslib.as_array(world, shape=(n, 3)).copy()
                im = np.ctypeslib.as_array(image, shape=(n, 2)).copy()
                kk = np.ctypeslib.as_array(k, shape=(3, 3)).copy()
                rv = np.ctypeslib.as_array(rvec, shape=(3,)).copy()
                tv = np.ctypeslib.as_array(tvec, shape=(3,)).copy()
                rv2, tv2, rm, inl = fn(w, im, kk, rv, tv)
                rv2 = np.asarray(rv2, np.float64).reshape(3)
                tv2 = np.asarray(tv2, np.float64).reshape(3)
                rm = np.asarray(rm, np.float64).reshape(9)
                inl = np.asarray(inl, np.int32).reshape(-1)
                for i in range(3):
                    rvec[i] = rv2[i]
                    tvec[i] = tv2[i]
                for i in range(9):
                    rmat[i] = rm[i]
                for i, v in enumerate(inl):
                    inliers[i] = int(v)
                n_inl[0] = len(inl)
                return 0
            except Exception:                              # never unwind through the C++ frames
                import traceback
                traceback.print_exc()
                return -1
        self._cb = PNP_FN(trampoline)
        rc = self.lib.svo_vo_set_pnp(self.h, self._cb, None)
        if rc != 0:
            raise SvoError("svo_vo_set_pnp failed")

    def process(self, left: np.ndarray, right: np.ndarray, t: float = 0.0) -> dict:
        left = np.ascontiguousarray(left, np.uint8)
        right = np.ascontiguousarray(right, np.uint8)
        hgt, wid = left.shape
        rc = self.lib.svo_vo_process(self.h, left.ctypes.data, right.ctypes.data, wid, hgt,
                                     left.strides[0], float(t))
        if rc != 0:
            raise SvoError(f"svo_vo_process failed ({rc}): "
                           f"{self.lib.svo_vo_last_error(self.h).decode(errors='replace')}")
        r = VoRecord()
        self.lib.svo_vo_last_record(self.h, C.byref(r))
        rec = dict(
            decision=DECISIONS[r.decision],
            detected=_arr(r.detected, r.n_detected, np.float32, 2),
            av_parallax=np.float32(r.av_parallax), percent_lost=np.float32(r.percent_lost),
            tracked_pts=_arr(r.tracked, r.n_tracked, np.float32, 2),
            tracked_ids=[int(i) for i in _arr(r.tracked_ids, r.n_tracked, np.uint64)],
            inliers=_arr(r.inliers, r.n_inliers, np.int32),
            hmat=np.array(list(r.hmat), np.float32).reshape(4, 4),
            new_features=_arr(r.new_features, r.n_new_features, np.float32, 2),
            disp16=_arr(r.disp16, r.n_new_features, np.int16),
            valid_idx=_arr(r.valid_idx, r.n_new, np.int32),
            new2d=_arr(r.new_2d, r.n_new, np.float32, 2),
            new3d=_arr(r.new_3d, r.n_new, np.float32, 3),
            new_ids=[int(i) for i in _arr(r.new_ids, r.n_new, np.uint64)],
            tracker_pts=_arr(r.tracker_pts, r.n_tracker, np.float32, 2),
            tracker_ids=[int(i) for i in _arr(r.tracker_ids, r.n_tracker, np.uint64)],
        )
        self.log.append(rec)
        return rec

    def get_drawing(self) -> np.ndarray:
        """FeatureTracker::get_drawing: (H, W, 3) RGB overlay of the last keyframe's tracks."""
        w, h = C.c_int32(), C.c_int32()
        rc = self.lib.svo_vo_get_drawing(self.h, None, 0, C.byref(w), C.byref(h))
        if rc != 0:
            raise SvoError(f"svo_vo_get_drawing failed ({rc}): "
                           f"{self.lib.svo_vo_last_error(self.h).decode(errors='replace')}")
        out = np.zeros((h.value, w.value, 3), np.uint8)
        rc = self.lib.svo_vo_get_drawing(self.h, out.ctypes.data, out.strides[0], None, None)
        if rc != 0:
            raise SvoError(f"svo_vo_get_drawing failed ({rc})")
        return out

    def close(self):
        if self.h:
            self.lib.svo_vo_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BundleAdjusterBookkeeping:
    """The facade's Ceres-free BundleAdjuster on its own (svo_ba_*): ids, refcounts, sliding
    window, world points - src/bundle_adjuster.cpp:41-58,60-135,159-163. Host code, no GPU."""

    def __init__(self, window: int = 5):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.svo_ba_create(window, C.byref(h))
        if rc != 0:
            raise SvoError(f"svo_ba_create failed ({rc})")
        self.h = h

    def _err(self, what, rc):
        return SvoError(f"{what} failed ({rc}): "
                        f"{self.lib.svo_ba_last_error(self.h).decode(errors='replace')}")

    def add_keyframe(self, tracked_ids, new_3d) -> np.ndarray:
        tid = np.ascontiguousarray(tracked_ids, np.uint64).reshape(-1)
        n3 = np.ascontiguousarray(new_3d, np.float32).reshape(-1, 3)
        out = np.zeros(max(1, n3.shape[0]), np.uint64)
        kept = C.c_int32(0)
        rc = self.lib.svo_ba_add_keyframe(self.h, tid.ctypes.data, tid.shape[0], n3.ctypes.data,
                                          n3.shape[0], out.ctypes.data, C.byref(kept))
        if rc != 0:
            raise self._err("svo_ba_add_keyframe", rc)
        return out[:kept.value].copy()

    def world_points(self, ids) -> np.ndarray:
        ids = np.ascontiguousarray(ids, np.uint64).reshape(-1)
        out = np.zeros((max(1, ids.shape[0]), 3), np.float32)
        rc = self.lib.svo_ba_get_world_points(self.h, ids.ctypes.data, ids.shape[0], out.ctypes.data)
        if rc != 0:
            raise self._err("svo_ba_get_world_points", rc)
        return out[:ids.shape[0]]

    @property
    def num_features(self) -> int:
        return int(self.lib.svo_ba_num_features(self.h))

    def close(self):
        if self.h:
            self.lib.svo_ba_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def image_msg(data: np.ndarray, encoding: str, stamp: float = 0.0, step: Optional[int] = None,
              big_endian: bool = False, width: Optional[int] = None) -> ImageMsg:
    """An svo_image_msg over `data` (rows of `step` bytes; the array is kept alive by the
    message). width defaults to what the row length and the encoding imply."""
    buf = np.ascontiguousarray(data)
    h = buf.shape[0]
    row_bytes = buf.strides[0]
    m = ImageMsg(float(stamp), h, int(width if width is not None else buf.shape[1]),
                 encoding.encode(), 1 if big_endian else 0,
                 int(step if step is not None else row_bytes), buf.ctypes.data, buf.nbytes)
    m._keep = buf
    return m


def to_mono8(msg: ImageMsg, pitch: Optional[int] = None) -> np.ndarray:
    """cv_bridge::toCvCopy(msg, "mono8") (svo_image_to_mono8)."""
    lib = load_library()
    pitch = int(pitch if pitch is not None else msg.width)
    out = np.zeros((msg.height, max(pitch, 1)), np.uint8)
    rc = lib.svo_image_to_mono8(C.byref(msg), out.ctypes.data, pitch)
    if rc != 0:
        raise SvoError(f"svo_image_to_mono8 failed ({rc}): "
                       f"{lib.svo_image_last_error().decode(errors='replace')}")
    return out[:, :msg.width]


class StereoQueue:
    """handle_images and its queue<StereoPair> (src/vo_node.cpp:38-76) plus the node loop's
    drain (:140-143), svo_stereo_queue_*."""

    def __init__(self, drop_time: float = 0.05):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.svo_stereo_queue_create(drop_time, C.byref(h))
        if rc != 0:
            raise SvoError(f"svo_stereo_queue_create failed ({rc})")
        self.h = h

    def _err(self, what, rc):
        return SvoError(f"{what} failed ({rc}): "
                        f"{self.lib.svo_stereo_queue_last_error(self.h).decode(errors='replace')}")

    def push(self, left: ImageMsg, right: ImageMsg) -> bool:
        took = C.c_int32(0)
        rc = self.lib.svo_stereo_queue_push(self.h, C.byref(left), C.byref(right), C.byref(took))
        if rc != 0:
            raise self._err("svo_stereo_queue_push", rc)
        return bool(took.value)

    def __len__(self):
        return int(self.lib.svo_stereo_queue_size(self.h))

    def drain(self, consumer: Callable[[np.ndarray, np.ndarray, float], Optional[int]]) -> int:
        """consumer(left, right, t) for every queued pair, oldest first; a non-zero return
        stops the drain with that pair still queued. Returns the number of pairs popped."""
        def cb(_user, lp, rp, w, h, pl, pr, t):
            l = np.ctypeslib.as_array(C.cast(lp, C.POINTER(C.c_uint8)), (h, pl))[:, :w]
            r = np.ctypeslib.as_array(C.cast(rp, C.POINTER(C.c_uint8)), (h, pr))[:, :w]
            return int(consumer(l.copy(), r.copy(), t) or 0)
        n = C.c_int32(0)
        fn = PAIR_FN(cb)
        rc = self.lib.svo_stereo_queue_drain(self.h, fn, None, C.byref(n))
        self.last_drained = n.value
        if rc != 0:
            raise self._err("svo_stereo_queue_drain", rc)
        return n.value

    def process_all(self, vo: "VoSequence") -> int:
        """vo_node's loop body: every queued pair through ImageProcessor::process of `vo`."""
        n = C.c_int32(0)
        rc = self.lib.svo_vo_process_queue(vo.h, self.h, C.byref(n))
        if rc != 0:
            raise SvoError(f"svo_vo_process_queue failed ({rc}): "
                           f"{self.lib.svo_vo_last_error(vo.h).decode(errors='replace')}")
        return n.value

    def close(self):
        if self.h:
            self.lib.svo_stereo_queue_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class KittiSource:
    """KITTI odometry stereo frames + ground-truth poses (svo_kitti_*), the headless stand-in
    for the reference's kitti_node (src/kitti_node.cpp:16-68)."""

    def __init__(self, data_path: str, sequence: str):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.svo_kitti_open(data_path.encode(), sequence.encode(), C.byref(h))
        if rc != 0:
            raise SvoError(f"svo_kitti_open failed ({rc}): "
                           f"{self.lib.svo_kitti_last_error(None).decode(errors='replace')}")
        self.h = h
        n, w, hh = C.c_int64(), C.c_int32(), C.c_int32()
        self.lib.svo_kitti_info(self.h, C.byref(n), C.byref(w), C.byref(hh))
        self.n_frames, self.width, self.height = n.value, w.value, hh.value
        self._slab = None

    def _err(self, what, rc):
        return SvoError(f"{what} failed ({rc}): "
                        f"{self.lib.svo_kitti_last_error(self.h).decode(errors='replace')}")

    def read(self, i: int):
        """(left, right, pose 3x4) of frame i."""
        l = np.empty((self.height, self.width), np.uint8)
        r = np.empty((self.height, self.width), np.uint8)
        pose = np.zeros(12)
        rc = self.lib.svo_kitti_read(self.h, i, l.ctypes.data, r.ctypes.data, self.width,
                                     pose.ctypes.data)
        if rc != 0:
            raise self._err("svo_kitti_read", rc)
        return l, r, pose.reshape(3, 4)

    def prefetch(self, depth: int = 4, first: int = 0, threads: int = 1, pinned: bool = True):
        """Iterate (index, left, right, pose) with decoding running ahead on `threads` worker
        threads into a (pinned) ring; the arrays alias ring slots (valid for two iterations)."""
        shape = (depth, 2, self.height, self.width)
        self._slab = capi.pinned_empty(shape, np.uint8) if pinned else np.empty(shape, np.uint8)
        rc = self.lib.svo_kitti_prefetch_start_threads(self.h, self._slab.ctypes.data, depth,
                                                       self.width, first, threads)
        if rc != 0:
            raise self._err("svo_kitti_prefetch_start_threads", rc)
        idx, lp, rp = C.c_int64(), C.c_void_p(), C.c_void_p()
        pose = np.zeros(12)
        while True:
            rc = self.lib.svo_kitti_next(self.h, C.byref(idx), C.byref(lp), C.byref(rp), pose.ctypes.data)
            if rc < 0:
                raise self._err("svo_kitti_next", rc)
            if rc == 0:
                return
            slot = idx.value % depth
            assert lp.value == self._slab[slot, 0].ctypes.data
            yield idx.value, self._slab[slot, 0], self._slab[slot, 1], pose.reshape(3, 4).copy()

    def close(self):
        if self.h:
            self.lib.svo_kitti_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Batch:
    """Many independent sequences on one GPU through svo::BatchFrontEnd (svo_batch_*)."""

    def __init__(self, width, height, n_seq, n_groups, cam, device=0, parallax_thresh=20.0,
                 params: Optional[SvoParams] = None, closed_loop: bool = False):
        self.lib = load_library()
        p = params or capi.default_params(width, height)
        self.params = p
        f, cx, cy, b = cam.f32()
        h = C.c_void_p()
        rc = self.lib.svo_batch_create(C.byref(p), device, n_seq, n_groups, float(f), float(cx),
                                       float(cy), float(b), float(parallax_thresh), C.byref(h))
        if rc != 0:
            raise SvoError(f"svo_batch_create failed ({rc}): "
                           f"{self.lib.svo_batch_last_error(None).decode(errors='replace')}")
        self.h = h
        self.n_seq = n_seq
        self.n_groups = max(1, min(n_groups, n_seq))
        self.max_corners, self.max_features = p.max_corners, p.max_features
        self._keep = None
        self.closed_loop = closed_loop
        if closed_loop:
            # the pose step on the GPU with the reference's solvePnPRansac arguments
            rc = self.lib.svo_batch_set_closed_loop(self.h, 1, 100, 8.0, 0.99)
            if rc != 0:
                raise SvoError(f"svo_batch_set_closed_loop failed ({rc})")

    @staticmethod
    def _ptrs(items, device):
        if items is None:
            return None, None
        if device:
            arr = (C.c_void_p * len(items))(*[int(x) for x in items])
            return arr, None
        keep = [np.ascontiguousarray(a) if a.strides[1] != 1 else a for a in items]
        arr = (C.c_void_p * len(keep))(*[a.ctypes.data for a in keep])
        return arr, keep

    def step(self, left, right, pitch, device=False, poses=None, next_left=None, next_right=None):
        lp, k1 = self._ptrs(left, device)
        rp, k2 = self._ptrs(right, device)
        nlp, k3 = self._ptrs(next_left, False) if not device else (None, None)
        nrp, k4 = self._ptrs(next_right, False) if not device else (None, None)
        self._keep = (k1, k2, k3, k4)      # prefetched host frames must outlive this call
        pz = None
        if poses is not None:
            poses = np.ascontiguousarray(poses, np.float32).reshape(self.n_seq, 16)
            pz = poses.ctypes.data_as(C.c_void_p)
        s = BatchSummary()
        rc = self.lib.svo_batch_step(self.h, lp, rp, int(pitch), 1 if device else 0, pz, nlp, nrp,
                                     C.byref(s))
        if rc != 0:
            raise SvoError(f"svo_batch_step failed ({rc}): "
                           f"{self.lib.svo_batch_last_error(self.h).decode(errors='replace')}")
        return dict(frames=s.frames, keyframes=s.keyframes, corners=s.corners, tracked=s.tracked,
                    new_points=s.new_points)

    def step_raw(self, lp, rp, pitch, device, poses_ptr=None, nlp=None, nrp=None) -> BatchSummary:
        """svo_batch_step on prebuilt ctypes pointer arrays (bench inner loop)."""
        s = BatchSummary()
        rc = self.lib.svo_batch_step(self.h, lp, rp, int(pitch), 1 if device else 0, poses_ptr,
                                     nlp, nrp, C.byref(s))
        if rc != 0:
            raise SvoError(f"svo_batch_step failed ({rc}): "
                           f"{self.lib.svo_batch_last_error(self.h).decode(errors='replace')}")
        return s

    def submit_raw(self, lp, rp, pitch, device, poses_ptr=None, nlp=None, nrp=None):
        """svo_batch_submit on prebuilt ctypes pointer arrays (at most 2 steps in flight)."""
        rc = self.lib.svo_batch_submit(self.h, lp, rp, int(pitch), 1 if device else 0, poses_ptr,
                                       nlp, nrp)
        if rc != 0:
            raise SvoError(f"svo_batch_submit failed ({rc}): "
                           f"{self.lib.svo_batch_last_error(self.h).decode(errors='replace')}")

    def wait(self) -> BatchSummary:
        s = BatchSummary()
        rc = self.lib.svo_batch_wait(self.h, C.byref(s))
        if rc != 0:
            raise SvoError(f"svo_batch_wait failed ({rc}): "
                           f"{self.lib.svo_batch_last_error(self.h).decode(errors='replace')}")
        return s

    def submit(self, left, right, pitch, device=False, poses=None, next_left=None, next_right=None):
        """Pipelined step: returns at once; call wait() for the results (oldest first). The
        image arrays are kept alive until two more submits have gone by."""
        lp, k1 = self._ptrs(left, device)
        rp, k2 = self._ptrs(right, device)
        nlp, k3 = self._ptrs(next_left, False) if not device else (None, None)
        nrp, k4 = self._ptrs(next_right, False) if not device else (None, None)
        pz = None
        if poses is not None:
            poses = np.ascontiguousarray(poses, np.float32).reshape(self.n_seq, 16)
            pz = poses.ctypes.data_as(C.c_void_p)
        self._pipe_keep = (getattr(self, "_pipe_keep", []) + [(k1, k2, k3, k4, poses)])[-4:]
        self.submit_raw(lp, rp, pitch, device, pz, nlp, nrp)

    # device-resident frames (bench kernel-only leg): allocations through group 0's engine
    def device_alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        rc = capi.load_library().svo_device_alloc(self.engine_handle(0), nbytes, C.byref(p))
        if rc != 0:
            raise SvoError(f"svo_device_alloc failed ({rc})")
        return p.value

    def device_upload(self, dst: int, src: np.ndarray):
        src = np.ascontiguousarray(src)
        rc = capi.load_library().svo_device_upload(self.engine_handle(0), C.c_void_p(dst),
                                                   src.ctypes.data_as(C.c_void_p), src.nbytes)
        if rc != 0:
            raise SvoError(f"svo_device_upload failed ({rc})")

    def device_free(self, ptr: int):
        capi.load_library().svo_device_free(self.engine_handle(0), C.c_void_p(ptr))

    def get(self, seq: int) -> dict:
        dec, nc, nt, nn = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        corners = np.zeros((self.max_corners, 2), np.float32)
        stats = np.zeros(1, STATS_DTYPE)
        tp = np.zeros((self.max_features, 2), np.float32)
        ti = np.zeros(self.max_features, np.uint64)
        n2 = np.zeros((self.max_corners, 2), np.float32)
        n3 = np.zeros((self.max_corners, 3), np.float32)
        rc = self.lib.svo_batch_get(self.h, seq, C.byref(dec), C.byref(nc), corners.ctypes.data,
                                    stats.ctypes.data, C.byref(nt), tp.ctypes.data, ti.ctypes.data,
                                    C.byref(nn), n2.ctypes.data, n3.ctypes.data)
        if rc != 0:
            raise SvoError(f"svo_batch_get failed ({rc})")
        out = dict(decision=DECISIONS[dec.value], corners=corners[:nc.value], stats=stats[0],
                   tracker_pts=tp[:nt.value], tracker_ids=ti[:nt.value], new2d=n2[:nn.value],
                   new3d=n3[:nn.value])
        hm = (C.c_float * 16)()
        ni = C.c_int32()
        self.lib.svo_batch_get_pose(self.h, seq, hm, C.byref(ni))
        out.update(hmat=np.array(list(hm), np.float32).reshape(4, 4), n_inliers=ni.value)
        if dec.value in (2, 4):
            nf, nv = C.c_int32(), C.c_int32()
            feats = np.zeros((self.max_corners, 2), np.float32)
            d16 = np.zeros(self.max_corners, np.int16)
            vi = np.zeros(self.max_corners, np.int32)
            rc = self.lib.svo_batch_get_keyframe(self.h, seq, C.byref(nf), feats.ctypes.data,
                                                 d16.ctypes.data, C.byref(nv), vi.ctypes.data)
            if rc != 0:
                raise SvoError(f"svo_batch_get_keyframe failed ({rc})")
            out.update(new_features=feats[:nf.value], disp16=d16[:nf.value],
                       valid_idx=vi[:nv.value])
        return out

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.svo_batch_kernel_launches(self.h))

    def timer_begin(self):
        self.lib.svo_batch_timer_begin(self.h)

    def timer_end(self) -> float:
        ms = C.c_float()
        self.lib.svo_batch_timer_end(self.h, C.byref(ms))
        return float(ms.value)

    def set_profiling(self, on: bool):
        self.lib.svo_batch_set_profiling(self.h, 1 if on else 0)

    def set_blocking_sync(self, on: bool):
        self.lib.svo_batch_set_blocking_sync(self.h, 1 if on else 0)

    def get_profile(self) -> dict:
        lib = capi.load_library()
        n = lib.svo_profile_slots()
        ms = (C.c_double * n)()
        cnt = (C.c_int64 * n)()
        self.lib.svo_batch_get_profile(self.h, ms, cnt)
        return {lib.svo_profile_name(i).decode(): (ms[i], cnt[i]) for i in range(n)}

    def engine_handle(self, group: int = 0):
        return self.lib.svo_batch_engine(self.h, group)

    def close(self):
        if self.h:
            self.lib.svo_batch_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
