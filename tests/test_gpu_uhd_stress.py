"""GPU parity at the high-resolution stress configuration (BASELINE configs[4]): 3840x2160
synthetic stereo, 4-level pyramid, goodFeaturesToTrack(20000, 0.01, 20), feature limit lifted.
Exercises the multi-block candidate sort (> 16k candidates), large corner / feature capacities
and image sizes far from the KITTI shape. Same bars as everywhere: corners, disparities, status
and ids bit-exact, flow <= 0.01 px."""
import numpy as np
import pytest

from stereo_vo_b200 import synth

pytestmark = pytest.mark.gpu
FLOW_TOL = 0.0           # px: flow is bit-exact (window sums in cv2's own f32 order)
PAR_TOL = 1e-4           # av_parallax: an f32 mean of f32 norms
MAXC, QUALITY, MIN_DIST = 20000, 0.01, 20.0


@pytest.fixture(scope="module")
def uhd(svo_lib):
    from stereo_vo_b200.capi import Engine
    cam = synth.UHD4K
    f0 = synth.stereo_frame(0, 0, cam, synth.UHD4K_MOTION)
    f1 = synth.stereo_frame(0, 1, cam, synth.UHD4K_MOTION)
    e = Engine(cam.width, cam.height, n_seq=1, max_corners=MAXC, quality_level=QUALITY,
               min_distance=MIN_DIST, max_features=MAXC)
    yield cam, f0, f1, e
    e.close()


def test_uhd_pyramid_bit_exact(uhd, oracle):
    import cv2
    cam, (l0, _), _, e = uhd
    e.good_features_to_track(l0)              # uploads the frame and builds its pyramid
    ref = l0
    for lvl in range(1, 4):
        ref = cv2.pyrDown(ref)
        got = e.debug_pyramid_level(lvl)
        assert got.shape == ref.shape
        assert np.array_equal(got, ref), f"level {lvl}"


def test_uhd_20k_corners_bit_exact(uhd, oracle):
    cam, (l0, _), _, e = uhd
    ref = oracle.detect(l0, MAXC, QUALITY, MIN_DIST)
    got = e.good_features_to_track(l0)
    assert len(ref) > 8000                    # min-distance 20 bounds it near 12k on this texture
    assert got.shape == ref.shape
    assert np.array_equal(got, ref)


def test_uhd_stereo_at_all_corners(uhd, oracle):
    cam, (l0, r0), _, e = uhd
    pts = e.good_features_to_track(l0)
    f, cx, cy, b = cam.f32()
    pose = np.eye(4, dtype=np.float32)
    p3, v2, vi, d16 = e.triangulate_stereo(l0, r0, pts, pose, f, cx, cy, b)
    ref16 = oracle.stereo_disparity16(l0, r0)
    want = ref16[pts[:, 1].astype(int), pts[:, 0].astype(int)]
    assert np.array_equal(d16, want)
    K = oracle.camera_matrix(f, cx, cy)
    r3, r2, ridx, _ = oracle.triangulate_stereo(K, b, pts, l0, r0, pose)
    assert np.array_equal(vi, ridx) and np.array_equal(v2, r2)
    assert len(ridx) > 5000
    assert np.abs(p3 - r3).max() <= 1e-4 * max(1.0, float(np.abs(r3).max()))


def test_uhd_track_thousands_of_points(uhd, oracle):
    cam, (l0, _), (l1, _), e = uhd
    pts = e.good_features_to_track(l0)
    ids = list(range(len(pts)))
    trk = oracle.FeatureTracker()
    trk.init(l0, pts, ids)
    av_ref, lost_ref = trk.track_features(l1, True)
    e.tracker_init(l0, pts, np.asarray(ids, np.uint64))
    av, lost = e.tracker_track(l1, True)
    dbg = e.tracker_debug()
    taps = trk.taps
    assert np.array_equal(dbg["status1"], taps["status1"])
    ok1 = taps["status1"] > 0
    assert ok1.sum() > 5000
    assert np.abs(dbg["tracked"][ok1] - taps["tracked"][ok1]).max() <= FLOW_TOL
    feats, fids = e.tracker_get()
    ref_feats, ref_ids = trk.get_tracked_features()
    assert list(fids) == list(ref_ids)
    assert np.abs(feats - ref_feats).max() <= FLOW_TOL
    assert abs(float(av) - float(av_ref)) <= PAR_TOL and lost == lost_ref
