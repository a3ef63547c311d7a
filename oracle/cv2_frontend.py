"""ORACLE (test infrastructure, not product code): the reference's front end restated call for
call on top of Python cv2.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module. The product path (stereo_vo_b200/ + libsvo.so) never does.

Why cv2: every arithmetic step of the hot path lives in OpenCV, which the reference links but
does not vendor (CMakeLists.txt:27 `find_package(OpenCV REQUIRED)`, version unpinned; the code
is of the OpenCV 3.x era). The reference cannot be built as it stands here (needs catkin/roscpp,
OpenCV C++ SDK, Eigen, Ceres - none present). Python cv2 4.13.0 exports the same functions, so
this file follows src/image_processor.cpp:18-208 and src/feature_tracker.cpp:3-72 line by line
and calls them with the reference's literal arguments. BundleAdjuster is reduced to its
bookkeeping (ids and the world-point store, src/bundle_adjuster.cpp:60-135,159-163);
bundle_adjust() is a no-op because Ceres is absent (world points fed to PnP are un-refined on
both sides of every comparison).

Parity status: the reference has no tests, fixtures or golden vectors.
What pins this oracle: (0) the reference's OWN sources, compiled where they lie against stand-in
headers with OpenCV's algorithms served by cv2 (oracle/build_ref.py, oracle/ref_frontend.py),
do frame by frame what this file does - decisions, tracker ids and positions, LK / PnP inputs,
new features, ids, 3-D points to 1e-5, the literal arguments of every OpenCV call
(tests/test_reference_frontend.py; bookkeeping and reprojection factors:
tests/test_reference_ba.py); (1) it calls the very OpenCV entry points the reference calls;
(2) oracle/restate.py re-derives each of them from published arithmetic and is checked against
cv2 bit for bit in tests/test_oracle_restate.py; (3) golden vectors generated from this file are
committed under tests/golden/ and reproduced by the compiled reference. Still UNPINNED: OpenCV
4.13 here against the un-versioned 3.x the reference was written for.

Canonical arithmetic mode is cv2.setUseOptimized(False): the optimized cornerMinEigenVal uses
position-dependent FMA contraction.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional

import cv2
import numpy as np

# hard-coded constants of the reference
MAX_CORNERS = 300           # src/image_processor.cpp:22
QUALITY_LEVEL = 0.1         # src/image_processor.cpp:22
MIN_FEATURE_DISTANCE = 30.0  # src/vo_node.cpp:34
PARALLAX_THRESH = 20.0      # src/vo_node.cpp:35
MAX_FEATURES = 400          # src/bundle_adjuster.hpp:75
NUM_DISPARITIES = 16 * 3    # src/image_processor.cpp:174
BLOCK_SIZE = 21             # src/image_processor.cpp:174
LK_WIN = (21, 21)           # src/feature_tracker.cpp:24
LK_MAX_LEVEL = 3
LK_CRITERIA = (cv2.TERM_CRITERIA_COUNT + cv2.TERM_CRITERIA_EPS, 30, 0.01)
LK_MIN_EIG = 1e-2


def canonical_mode():
    """Select the arithmetic the parity tests are defined against."""
    cv2.setUseOptimized(False)


def detect(left: np.ndarray, max_corners=MAX_CORNERS, quality=QUALITY_LEVEL,
           min_distance=MIN_FEATURE_DISTANCE) -> np.ndarray:
    """src/image_processor.cpp:22. Returns (n, 2) float32, descending score order."""
    pts = cv2.goodFeaturesToTrack(left, max_corners, quality, min_distance)
    if pts is None:
        return np.zeros((0, 2), np.float32)
    return pts.reshape(-1, 2).astype(np.float32)


def stereo_disparity16(left: np.ndarray, right: np.ndarray) -> np.ndarray:
    """src/image_processor.cpp:174-175: a fresh StereoBM(48, 21) per call, CV_16S output."""
    sbm = cv2.StereoBM_create(NUM_DISPARITIES, BLOCK_SIZE)
    return sbm.compute(left, right)


def triangulate_stereo(cam_matrix: np.ndarray, baseline: np.float32, feats: np.ndarray,
                       left: np.ndarray, right: np.ndarray, camera_pose: np.ndarray):
    """src/image_processor.cpp:165-208. Returns (features_3d, valid_features_2d, valid_idx,
    disp16) - the last two are extra taps for the parity logs."""
    disp16 = stereo_disparity16(left, right)
    disparity = disp16.astype(np.float32) * np.float32(1.0 / 16)   # :176 convertTo(CV_32F, 1/16)
    focal = np.float32(cam_matrix[0, 0])
    cx = np.float32(cam_matrix[0, 2])
    cy = np.float32(cam_matrix[1, 2])
    q = np.zeros((4, 4), np.float32)                                # :183-189 (double -> float)
    q[0, 0] = np.float32(1.0 / float(focal))
    q[1, 1] = np.float32(1.0 / float(focal))
    q[0, 3] = np.float32(-cx / focal)
    q[1, 3] = np.float32(-cy / focal)
    q[2, 3] = np.float32(1.0)
    q[3, 2] = np.float32(1.0 / float(np.float32(baseline) * focal))
    pose = np.ascontiguousarray(camera_pose, dtype=np.float32)
    feats3d, valid2d, valid_idx, d16 = [], [], [], []
    for i in range(feats.shape[0]):
        x, y = feats[i]
        d = disparity[int(y), int(x)]                               # :193 float -> int truncation
        if d > 0:
            valid2d.append((x, y))
            valid_idx.append(i)
            d16.append(int(disp16[int(y), int(x)]))
            v = np.array([[x], [y], [d], [1.0]], np.float32)
            # :202 `camera_pose*Q*disparity_point` = two cv::gemm calls, left to right
            m = cv2.gemm(pose, q, 1.0, None, 0.0)
            w = cv2.gemm(m, v, 1.0, None, 0.0).reshape(4)
            feats3d.append((w[0] / w[3], w[1] / w[3], w[2] / w[3]))
    return (np.array(feats3d, np.float32).reshape(-1, 3),
            np.array(valid2d, np.float32).reshape(-1, 2),
            np.array(valid_idx, np.int32), np.array(d16, np.int16))


def lk(prev_img, next_img, pts):
    """src/feature_tracker.cpp:23-26 / 32-35 (identical arguments)."""
    if pts.shape[0] == 0:
        return np.zeros((0, 2), np.float32), np.zeros((0,), np.uint8)
    nxt, status, _err = cv2.calcOpticalFlowPyrLK(
        prev_img, next_img, pts.reshape(-1, 1, 2).astype(np.float32), None,
        winSize=LK_WIN, maxLevel=LK_MAX_LEVEL, criteria=LK_CRITERIA, flags=0,
        minEigThreshold=LK_MIN_EIG)
    return nxt.reshape(-1, 2), status.reshape(-1)


class FeatureTracker:
    """src/feature_tracker.{hpp,cpp} (draw_track/get_drawing omitted: debug overlay)."""

    def __init__(self):
        self.last_image = None
        self.initial_features = {}
        self.feature_set = np.zeros((0, 2), np.float32)
        self.feature_ids: List[int] = []
        self.taps = {}

    def init(self, image, features, ids):                           # :3-16
        self.feature_ids = list(ids)
        self.feature_set = np.array(features, np.float32).reshape(-1, 2)
        self.initial_features = {}
        for i, fid in enumerate(ids):
            # unordered_map::insert keeps the FIRST value for a duplicate key
            self.initial_features.setdefault(fid, self.feature_set[i].copy())
        self.last_image = image.copy()

    def track_features(self, image, flow_back=True):                # :18-67
        tracked, status1 = lk(self.last_image, image, self.feature_set)
        if flow_back:
            reverse, status2 = lk(image, self.last_image, tracked)
        old_set = self.feature_set
        old_ids = self.feature_ids
        new_set, new_ids = [], []
        av_parallax = np.float32(0)
        n = tracked.shape[0]
        kept_idx = []
        for i in range(n):
            ok = bool(status1[i])
            if ok and flow_back:
                dx = float(old_set[i, 0] - reverse[i, 0])           # Point2f difference is f32
                dy = float(old_set[i, 1] - reverse[i, 1])
                ok = bool(status2[i]) and math.sqrt(dx * dx + dy * dy) < 2   # cv::norm -> double
            if ok:
                init = self.initial_features[old_ids[i]]
                dx = np.float32(tracked[i, 0] - init[0])
                dy = np.float32(tracked[i, 1] - init[1])
                # :51 sqrt(dx*dx+dy*dy) on floats: std::sqrt(float) overload, f32 arithmetic
                parallax = np.float32(np.sqrt(np.float32(np.float32(dx * dx) + np.float32(dy * dy))))
                if parallax > 200:
                    continue
                new_set.append(tracked[i])
                new_ids.append(old_ids[i])
                kept_idx.append(i)
                av_parallax = np.float32(av_parallax + parallax)
        with np.errstate(divide="ignore", invalid="ignore"):
            av_parallax = np.float32(av_parallax / np.float32(n))   # SURVEY B2 (n == 0 -> NaN)
            # :64 `1.0 - size/(float)size` : float division, double subtraction, stored to float
            ratio = np.float32(np.float32(len(new_set)) / np.float32(len(self.initial_features)))
            percent_lost = np.float32(1.0 - float(ratio))
        self.taps = dict(tracked=tracked, status1=status1,
                         reverse=reverse if flow_back else None,
                         status2=status2 if flow_back else None,
                         kept_idx=np.array(kept_idx, np.int32))
        self.feature_set = np.array(new_set, np.float32).reshape(-1, 2)
        self.feature_ids = new_ids
        self.last_image = image.copy()
        return av_parallax, percent_lost

    def get_tracked_features(self):                                 # :69-72
        return self.feature_set.copy(), list(self.feature_ids)


@dataclass
class Keyframe:                                                     # src/bundle_adjuster.hpp:22-46
    position: np.ndarray
    rmat: np.ndarray          # the reference stores a quaternion; the bookkeeping keeps R
    tracked_features_2d: np.ndarray
    tracked_ids: List[int]
    new_features_2d: np.ndarray
    new_features_3d: np.ndarray
    new_ids: List[int] = field(default_factory=list)


class BundleAdjusterBookkeeping:
    """src/bundle_adjuster.cpp:60-135,159-163 without Ceres: ids + world-point store."""

    def __init__(self, max_features=MAX_FEATURES):
        self.features: List[np.ndarray] = []
        self.last_keyframe: Optional[Keyframe] = None
        self.max_features = max_features

    def get_last_keyframe(self):
        return self.last_keyframe

    def add_keyframe(self, kf: Keyframe):
        num_tracked = len(kf.tracked_ids)
        max_new = self.max_features - num_tracked
        # the reference's truncation branch corrupts new_ids; never hit in our runs
        assert kf.new_features_2d.shape[0] <= max_new, "reference UB branch (B6) would trigger"
        for i in range(kf.new_features_2d.shape[0]):
            # avail_ids is never refilled, so ids are a running counter
            self.features.append(kf.new_features_3d[i].astype(np.float64))
            kf.new_ids.append(len(self.features) - 1)
        self.last_keyframe = kf

    def bundle_adjust(self):
        return

    def get_world_points(self, ids):
        return np.array([self.features[i] for i in ids], np.float32).reshape(-1, 3)


def solve_pnp(world, pts2d, cam_matrix, rvec, tvec):
    """src/image_processor.cpp:75-92.

    The extrinsic guess is handed over as fresh float64 arrays: with cv2 4.13 a float32 (3,1)
    in/out rvec/tvec (what the reference's CV_32F zeros at :54-55 become in Python) makes
    solvePnPRansac return garbage in ~3 % of calls (uninitialised read inside the binding;
    reproduced single-threaded). The values of the guess are unchanged by the widening."""
    rvec = np.array(rvec, np.float64).reshape(3, 1).copy()
    tvec = np.array(tvec, np.float64).reshape(3, 1).copy()
    ok, rvec, tvec, inliers = cv2.solvePnPRansac(
        world.reshape(-1, 1, 3).astype(np.float32), pts2d.reshape(-1, 1, 2).astype(np.float32),
        cam_matrix, np.zeros((4, 1), np.float32), rvec, tvec, True, 100, 8.0, 0.99)
    inl = np.zeros((0,), np.int32) if inliers is None else inliers.reshape(-1).astype(np.int32)
    rmat, _ = cv2.Rodrigues(rvec)
    return rvec, tvec, rmat, inl


def pose_inverse(rmat, tvec):
    """src/image_processor.cpp:130-134: hmat = [R^T | -R^T t] as CV_32F."""
    hmat = np.eye(4, dtype=np.float32)
    rt = rmat.T
    hmat[:3, :3] = rt.astype(np.float32)
    hmat[:3, 3] = (-(rt @ tvec.reshape(3, 1))).reshape(3).astype(np.float32)
    return hmat


def dedupe(detected, tracked2d, min_distance=MIN_FEATURE_DISTANCE):
    """src/image_processor.cpp:113-128, f32 arithmetic (sqrt on float)."""
    keep = []
    md = np.float32(min_distance)
    for i in range(detected.shape[0]):
        tracked = False
        if tracked2d.shape[0]:
            dx = (detected[i, 0] - tracked2d[:, 0]).astype(np.float32)
            dy = (detected[i, 1] - tracked2d[:, 1]).astype(np.float32)
            dist = np.sqrt((dx * dx).astype(np.float32) + (dy * dy).astype(np.float32),
                           dtype=np.float32)
            tracked = bool(np.any(dist < md))
        if not tracked:
            keep.append(i)
    return np.array(keep, np.int32)


class ImageProcessor:
    """src/image_processor.{hpp,cpp}. `log` collects per-frame taps for the parity tests."""

    def __init__(self, cam_matrix, tracker, adjuster, baseline, min_feature_distance=30.0,
                 parallax_thresh=20.0, teacher=None):
        self.camera_matrix = np.asarray(cam_matrix, np.float32)
        self.feature_tracker = tracker
        self.bundle_adjuster = adjuster
        self.baseline = np.float32(baseline)
        self.min_feature_distance = np.float32(min_feature_distance)
        self.parallax_thresh = np.float32(parallax_thresh)
        self.rvec = None
        self.tvec = None
        self.log = []

    def process(self, left, right):
        rec = {"decision": None}
        self.log.append(rec)
        detected = detect(left, MAX_CORNERS, QUALITY_LEVEL, float(self.min_feature_distance))
        rec["detected"] = detected
        if detected.shape[0] < 4:                                   # :23-25
            rec["decision"] = "too_few"
            return rec
        last = self.bundle_adjuster.get_last_keyframe()
        if last is None:                                            # :30-58
            f3d, v2d, vidx, d16 = triangulate_stereo(self.camera_matrix, self.baseline, detected,
                                                     left, right, np.eye(4, dtype=np.float32))
            kf = Keyframe(np.zeros(3, np.float32), np.eye(3, dtype=np.float32),
                          np.zeros((0, 2), np.float32), [], v2d, f3d)
            self.bundle_adjuster.add_keyframe(kf)
            self.feature_tracker.init(left, kf.new_features_2d, kf.new_ids)
            self.tvec = np.zeros((3, 1), np.float32)
            self.rvec = np.zeros((3, 1), np.float32)
            rec.update(decision="first", new_features=detected, hmat=np.eye(4, dtype=np.float32),
                       valid_idx=vidx, disp16=d16, new2d=v2d, new3d=f3d, new_ids=list(kf.new_ids),
                       tracker_pts=self.feature_tracker.feature_set.copy(),
                       tracker_ids=list(self.feature_tracker.feature_ids))
            return rec
        prev_pts = self.feature_tracker.feature_set.copy()
        av, lost = self.feature_tracker.track_features(left, True)  # :62
        rec.update(prev_pts=prev_pts, av_parallax=av, percent_lost=lost,
                   tracked_pts=self.feature_tracker.feature_set.copy(),
                   tracked_ids=list(self.feature_tracker.feature_ids),
                   taps=self.feature_tracker.taps)
        if av <= self.parallax_thresh and float(lost) < 0.4:               # :63-65 (NaN falls through)
            rec["decision"] = "skip"
            return rec
        tracked, tracked_ids = self.feature_tracker.get_tracked_features()
        world = self.bundle_adjuster.get_world_points(tracked_ids)
        self.rvec, self.tvec, rmat, inl = solve_pnp(world, tracked, self.camera_matrix,
                                                    self.rvec, self.tvec)
        kf_ids = [tracked_ids[i] for i in inl]
        kf_2d = tracked[inl].reshape(-1, 2)
        keep = dedupe(detected, kf_2d, self.min_feature_distance)   # :113-128
        new_features = detected[keep].reshape(-1, 2)
        hmat = pose_inverse(rmat, self.tvec)                        # :130-134
        f3d, v2d, vidx, d16 = triangulate_stereo(self.camera_matrix, self.baseline, new_features,
                                                 left, right, hmat)
        kf = Keyframe(self.tvec.reshape(3).astype(np.float32), rmat.astype(np.float32),
                      kf_2d, kf_ids, v2d, f3d)
        self.bundle_adjuster.add_keyframe(kf)                       # :144
        pts = np.concatenate([kf.tracked_features_2d, kf.new_features_2d], 0)   # :148-162
        ids = list(kf.tracked_ids) + list(kf.new_ids)
        self.feature_tracker.init(left, pts, ids)
        rec.update(decision="keyframe", inliers=inl, hmat=hmat, dedupe_keep=keep,
                   new_features=new_features, valid_idx=vidx, disp16=d16, new2d=v2d, new3d=f3d,
                   new_ids=list(kf.new_ids), tracker_pts=pts.copy(), tracker_ids=list(ids),
                   rvec=np.array(self.rvec).copy(), tvec=np.array(self.tvec).copy())
        return rec


def camera_matrix(focal, cx, cy):
    """src/vo_node.cpp:104-108."""
    k = np.eye(3, dtype=np.float32)
    k[0, 0] = np.float32(focal)
    k[0, 2] = np.float32(cx)
    k[1, 1] = np.float32(focal)
    k[1, 2] = np.float32(cy)
    return k


def make_pipeline(cam):
    """src/vo_node.cpp:112-119."""
    tracker = FeatureTracker()
    adjuster = BundleAdjusterBookkeeping()
    f, cx, cy, b = cam.f32()
    proc = ImageProcessor(camera_matrix(f, cx, cy), tracker, adjuster, b,
                          MIN_FEATURE_DISTANCE, PARALLAX_THRESH)
    return proc
