"""Host-side mirror of the reference's front-end interface on top of the C ABI (libsvo.so).

Same names, argument meaning and control flow as the reference's C++ classes, so the parity
tests read like a walk through src/image_processor.cpp / src/feature_tracker.cpp:

    FeatureTracker.init / track_features / get_tracked_features   (src/feature_tracker.hpp:20-44)
    ImageProcessor.process(stereo pair)                           (src/image_processor.hpp:46)
    BundleAdjuster bookkeeping: ids + world points only           (src/bundle_adjuster.cpp:60-135)

All pixel work goes through the CUDA library; the host keeps only what the reference keeps on
the host: the keyframe decision, PnP-RANSAC (a consumer of the GPU output) and id
assignment. The C++ twin of this file is include/stereo_vo/*.hpp + csrc/host/facade.cpp.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence

import numpy as np

from .capi import Engine

MAX_FEATURES = 400          # src/bundle_adjuster.hpp:75


@dataclass
class Keyframe:             # src/bundle_adjuster.hpp:22-46
    position: np.ndarray
    rmat: np.ndarray
    tracked_features_2d: np.ndarray
    tracked_ids: List[int]
    new_features_2d: np.ndarray
    new_features_3d: np.ndarray
    new_ids: List[int] = field(default_factory=list)


class BundleAdjuster:
    """Bookkeeping half of src/bundle_adjuster.cpp (ids, world points); Ceres is out of scope."""

    def __init__(self, max_features: int = MAX_FEATURES):
        self.features: List[np.ndarray] = []
        self.last_keyframe: Optional[Keyframe] = None
        self.max_features = max_features

    def get_last_keyframe(self):
        return self.last_keyframe

    def add_keyframe(self, kf: Keyframe):
        max_new = self.max_features - len(kf.tracked_ids)
        if kf.new_features_2d.shape[0] > max_new:
            # the reference's truncation branch corrupts new_ids; refuse loudly
            raise RuntimeError("tracked + new features exceed max_features (reference UB branch)")
        for i in range(kf.new_features_2d.shape[0]):
            self.features.append(kf.new_features_3d[i].astype(np.float64))
            kf.new_ids.append(len(self.features) - 1)       # ids are a running counter (B7)
        self.last_keyframe = kf

    def bundle_adjust(self):
        return

    def get_world_points(self, ids: Sequence[int]) -> np.ndarray:
        return np.array([self.features[i] for i in ids], np.float32).reshape(-1, 3)


class FeatureTracker:
    """src/feature_tracker.{hpp,cpp} on the GPU: stand-alone stage calls of the C ABI."""

    def __init__(self, engine: Engine, seq: int = 0):
        self.engine = engine
        self.seq = seq

    def init(self, image: np.ndarray, features: np.ndarray, ids: Sequence[int]):
        self.engine.tracker_init(image, features, np.asarray(ids, np.uint64), self.seq)

    def track_features(self, image: np.ndarray, flow_back: bool = True):
        """Returns (av_parallax, percent_lost) like the reference's out-parameters."""
        return self.engine.tracker_track(image, flow_back, self.seq)

    def get_tracked_features(self):
        feats, ids = self.engine.tracker_get(self.seq)
        return feats, [int(i) for i in ids]


def camera_matrix(focal, cx, cy) -> np.ndarray:
    """src/vo_node.cpp:104-108."""
    k = np.eye(3, dtype=np.float32)
    k[0, 0] = np.float32(focal)
    k[0, 2] = np.float32(cx)
    k[1, 1] = np.float32(focal)
    k[1, 2] = np.float32(cy)
    return k


def pose_inverse(rmat: np.ndarray, tvec: np.ndarray) -> np.ndarray:
    """src/image_processor.cpp:130-134: hmat = [R^T | -R^T t] (camera -> world), CV_32F."""
    hmat = np.eye(4, dtype=np.float32)
    rt = rmat.T
    hmat[:3, :3] = rt.astype(np.float32)
    hmat[:3, 3] = (-(rt @ tvec.reshape(3, 1))).reshape(3).astype(np.float32)
    return hmat


def opencv_pnp(world, pts2d, cam_matrix, rvec, tvec):
    """src/image_processor.cpp:75-92: host-side consumer, exactly the reference's call."""
    import cv2
    # float64 copies of the guess: cv2 4.13 mis-handles float32 in/out rvec/tvec (garbage poses
    # in ~3 % of calls); the guess values are unchanged
    rvec = np.array(rvec, np.float64).reshape(3, 1).copy()
    tvec = np.array(tvec, np.float64).reshape(3, 1).copy()
    _ok, rvec, tvec, inliers = cv2.solvePnPRansac(
        world.reshape(-1, 1, 3).astype(np.float32), pts2d.reshape(-1, 1, 2).astype(np.float32),
        cam_matrix, np.zeros((4, 1), np.float32), rvec, tvec, True, 100, 8.0, 0.99)
    inl = np.zeros((0,), np.int32) if inliers is None else inliers.reshape(-1).astype(np.int32)
    rmat, _ = cv2.Rodrigues(rvec)
    return rvec, tvec, rmat, inl


class ImageProcessor:
    """src/image_processor.{hpp,cpp}: process() keeps the reference's control flow and calls the
    fused GPU path (svo_step -> [host decision, PnP] -> svo_keyframe -> svo_tracker_set).

    `teacher`, when given, is a list of per-frame oracle records: the host-side decisions that
    are outside the accelerated path (PnP inlier indices and pose) are then replayed from it, so
    every GPU stage sees oracle-identical inputs (teacher-forced parity).
    """

    def __init__(self, engine: Engine, cam_matrix, adjuster: BundleAdjuster, baseline,
                 min_feature_distance: float = 30.0, parallax_thresh: float = 20.0, seq: int = 0,
                 pnp: Callable = opencv_pnp, teacher: Optional[list] = None):
        self.engine = engine
        self.seq = seq
        self.camera_matrix = np.asarray(cam_matrix, np.float32)
        self.bundle_adjuster = adjuster
        self.baseline = np.float32(baseline)
        self.min_feature_distance = np.float32(min_feature_distance)
        self.parallax_thresh = np.float32(parallax_thresh)
        self.pnp = pnp
        self.teacher = teacher
        self.rvec = None
        self.tvec = None
        self.frame = 0
        self.log: list = []
        # tracker state mirrored on the host (the device holds the authoritative copy)
        self.tracked = np.zeros((0, 2), np.float32)
        self.tracked_ids: List[int] = []

    def _intrinsics(self):
        k = self.camera_matrix
        return float(k[0, 0]), float(k[0, 2]), float(k[1, 2]), float(self.baseline)

    def process(self, left: np.ndarray, right: np.ndarray):
        e, s = self.engine, self.seq
        rec = {"decision": None}
        self.log.append(rec)
        t = self.frame
        self.frame += 1
        res = e.step([s], [left], [right])
        nc = int(res["n_corners"][0])
        detected = res["corners"][0, :nc].copy()
        rec["detected"] = detected
        if nc < 4:                                                   # image_processor.cpp:23-25
            rec["decision"] = "too_few"
            return rec
        f, cx, cy, b = self._intrinsics()
        last = self.bundle_adjuster.get_last_keyframe()
        if last is None:                                             # :30-58
            kr = e.keyframe([s], [np.zeros((0, 2), np.float32)], [0],
                            np.eye(4, dtype=np.float32).reshape(1, 16), f, cx, cy, b)
            nv, nn = int(kr["n_valid"][0]), int(kr["n_new"][0])
            kf = Keyframe(np.zeros(3, np.float32), np.eye(3, dtype=np.float32),
                          np.zeros((0, 2), np.float32), [], kr["valid_2d"][0, :nv].copy(),
                          kr["pts_3d"][0, :nv].copy())
            self.bundle_adjuster.add_keyframe(kf)
            e.tracker_set([s], [kf.new_features_2d], [kf.new_ids])   # :51-52
            self.tvec = np.zeros((3, 1), np.float32)
            self.rvec = np.zeros((3, 1), np.float32)
            rec.update(decision="first", new_features=kr["new_feats"][0, :nn].copy(),
                       valid_idx=kr["valid_idx"][0, :nv].copy(), disp16=kr["disp16"][0, :nn].copy(),
                       new2d=kf.new_features_2d, new3d=kf.new_features_3d,
                       new_ids=list(kf.new_ids), tracker_pts=kf.new_features_2d.copy(),
                       tracker_ids=list(kf.new_ids))
            return rec
        st = res["stats"][0]
        av, lost = np.float32(st["av_parallax"]), np.float32(st["percent_lost"])
        nk = int(st["n_kept"])
        tracked = res["tracked"][0, :nk].copy()
        tracked_ids = [int(i) for i in res["tracked_ids"][0, :nk]]
        rec.update(av_parallax=av, percent_lost=lost, tracked_pts=tracked, tracked_ids=tracked_ids)
        if av <= self.parallax_thresh and float(lost) < 0.4:         # :63-65
            rec["decision"] = "skip"
            return rec
        if self.teacher is not None:
            tr = self.teacher[t]
            assert tr["decision"] == "keyframe", "teacher log out of step with the GPU run"
            inl, hmat = tr["inliers"], tr["hmat"]
            self.rvec, self.tvec = tr["rvec"], tr["tvec"]
            rmat = hmat[:3, :3].T
        else:
            world = self.bundle_adjuster.get_world_points(tracked_ids)          # :72
            self.rvec, self.tvec, rmat, inl = self.pnp(world, tracked, self.camera_matrix,
                                                       self.rvec, self.tvec)
            hmat = pose_inverse(rmat, self.tvec)                                 # :130-134
        kf_ids = [tracked_ids[i] for i in inl]
        kf_2d = tracked[inl].reshape(-1, 2)
        kr = e.keyframe([s], [kf_2d], [kf_2d.shape[0]], hmat.reshape(1, 16), f, cx, cy, b)
        nv, nn = int(kr["n_valid"][0]), int(kr["n_new"][0])
        kf = Keyframe(np.asarray(self.tvec, np.float32).reshape(3), np.asarray(rmat, np.float32),
                      kf_2d, kf_ids, kr["valid_2d"][0, :nv].copy(), kr["pts_3d"][0, :nv].copy())
        self.bundle_adjuster.add_keyframe(kf)                                    # :144
        pts = np.concatenate([kf.tracked_features_2d, kf.new_features_2d], 0)    # :148-162
        ids = list(kf.tracked_ids) + list(kf.new_ids)
        e.tracker_set([s], [pts], [ids])
        rec.update(decision="keyframe", inliers=np.asarray(inl), hmat=hmat,
                   new_features=kr["new_feats"][0, :nn].copy(),
                   valid_idx=kr["valid_idx"][0, :nv].copy(), disp16=kr["disp16"][0, :nn].copy(),
                   new2d=kf.new_features_2d, new3d=kf.new_features_3d, new_ids=list(kf.new_ids),
                   tracker_pts=pts.copy(), tracker_ids=list(ids))
        return rec


class BatchFrontEnd:
    """Throughput driver: `n_seq` independent sequences stepped together through the batched C
    ABI (one svo_step per frame index, one svo_keyframe + svo_tracker_set for the sequences that
    became keyframes). Host PnP is outside the accelerated path: a keyframe
    here keeps every tracked feature (the survey measured inliers == all tracks on this data) and
    uses the pose supplied by `pose_fn(seq, t)` (ground truth of the synthetic trajectory).
    """

    def __init__(self, engine: Engine, cam, pose_fn: Optional[Callable] = None,
                 parallax_thresh: float = 20.0, max_features: int = MAX_FEATURES):
        self.e = engine
        self.n = engine.n_seq
        f, cx, cy, b = cam.f32()
        self.intr = (float(f), float(cx), float(cy), float(b))
        self.pose_fn = pose_fn
        self.parallax_thresh = np.float32(parallax_thresh)
        self.max_features = max_features
        self.next_id = [0] * self.n
        self.have_kf = [False] * self.n
        self.t = [0] * self.n
        self.keyframes = 0
        self.seqs = list(range(self.n))

    def step(self, left, right, pitch=None, device=False):
        e = self.e
        res = e.step(self.seqs, left, right, pitch=pitch, device=device)
        kf_seqs, inl, ninl, poses = [], [], [], []
        for i in range(self.n):
            t = self.t[i]
            self.t[i] += 1
            if res["n_corners"][i] < 4:
                continue
            if self.have_kf[i]:
                st = res["stats"][i]
                if st["av_parallax"] <= self.parallax_thresh and float(st["percent_lost"]) < 0.4:
                    continue
                nk = int(st["n_kept"])
                inl.append(res["tracked"][i, :nk])
                ninl.append(nk)
            else:
                inl.append(np.zeros((0, 2), np.float32))
                ninl.append(0)
            kf_seqs.append(i)
            # the first keyframe is triangulated with the identity (src/image_processor.cpp:34-39)
            poses.append(self.pose_fn(i, t) if self.pose_fn and self.have_kf[i]
                         else np.eye(4, dtype=np.float32))
        if kf_seqs:
            f, cx, cy, b = self.intr
            kr = e.keyframe(kf_seqs, inl, ninl, np.stack(poses).reshape(-1, 16), f, cx, cy, b)
            feats, ids = [], []
            for k, i in enumerate(kf_seqs):
                nv = int(kr["n_valid"][k])
                room = self.max_features - ninl[k]
                nv = min(nv, room)
                new_ids = np.arange(self.next_id[i], self.next_id[i] + nv, dtype=np.uint64)
                self.next_id[i] += nv
                old_ids = res["tracked_ids"][i, :ninl[k]] if self.have_kf[i] else \
                    np.zeros(0, np.uint64)
                feats.append(np.concatenate([inl[k], kr["valid_2d"][k, :nv]], 0))
                ids.append(np.concatenate([old_ids, new_ids]))
                self.have_kf[i] = True
            e.tracker_set(kf_seqs, feats, ids)
            self.keyframes += len(kf_seqs)
        return res, kf_seqs
