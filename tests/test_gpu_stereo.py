"""GPU parity: sparse StereoBM + Q reprojection + dedupe against the cv2 oracle.
disp16 and valid indices bit-exact; 3-D points within 1e-4 relative (BASELINE.json)."""
import numpy as np
import pytest

from stereo_vo_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng(svo_lib):
    from stereo_vo_b200.capi import Engine
    cache = {}

    def get(w, h, **kw):
        key = (w, h, tuple(sorted(kw.items())))
        if key not in cache:
            cache[key] = Engine(w, h, n_seq=1, **kw)
        return cache[key]
    yield get
    for e in cache.values():
        e.close()


def _check_disp(e, oracle, left, right, pts):
    pose = np.eye(4, dtype=np.float32)
    f, cx, cy, b = synth.KITTI00.f32()
    p3, v2, vi, d16 = e.triangulate_stereo(left, right, pts, pose, f, cx, cy, b)
    ref16 = oracle.stereo_disparity16(left, right)
    want = ref16[pts[:, 1].astype(int), pts[:, 0].astype(int)]
    assert np.array_equal(d16, want), f"{(d16 != want).sum()} of {len(want)} disparities differ"
    return p3, v2, vi, d16


def test_disparity_at_random_pixels_kitti(eng, oracle):
    l, r = util.frames(0, 1)[0]
    e = eng(1241, 376, max_features=4000, max_corners=4000)
    rng = np.random.default_rng(1)
    pts = np.stack([rng.integers(0, 1241, 4000), rng.integers(0, 376, 4000)], 1).astype(np.float32)
    _, _, _, d16 = _check_disp(e, oracle, l, r, pts)
    assert (d16 > 0).sum() > 2000


@pytest.mark.parametrize("w,h", [(640, 480), (641, 479), (333, 211)])
def test_disparity_stress_images(eng, oracle, w, h):
    """flat patches / periodic stripes / odd heights: texture + uniqueness rejections, the
    odd-height last row and the ROI border."""
    left = util.stress_image(h, w, 3)
    right = util.shifted(left, 17)
    e = eng(w, h, max_features=6000, max_corners=6000)
    rng = np.random.default_rng(2)
    pts = np.stack([rng.integers(0, w, 6000), rng.integers(0, h, 6000)], 1).astype(np.float32)
    # make sure ROI corners are probed
    pts[:8] = [[57, 10], [w - 11, 10], [57, h - 11], [w - 11, h - 11],
               [56, 10], [w - 10, 10], [57, 9], [57, h - 10]]
    _, _, _, d16 = _check_disp(e, oracle, left, right, pts)
    assert (d16 == -16).sum() > 500 and (d16 > 0).sum() > 500


def test_triangulate_matches_oracle(eng, oracle):
    l, r = util.frames(2, 1)[0]
    e = eng(1241, 376)
    feats = oracle.detect(l)
    f, cx, cy, b = synth.KITTI00.f32()
    k = oracle.camera_matrix(f, cx, cy)
    rng = np.random.default_rng(0)
    for trial in range(3):
        pose = np.eye(4, dtype=np.float32)
        if trial:
            a = rng.normal(size=3) * 0.2
            import cv2
            pose[:3, :3] = cv2.Rodrigues(a)[0].astype(np.float32)
            pose[:3, 3] = rng.normal(size=3).astype(np.float32) * 3
        ref3, ref2, refi, refd = oracle.triangulate_stereo(k, b, feats, l, r, pose)
        p3, v2, vi, d16 = e.triangulate_stereo(l, r, feats, pose, f, cx, cy, b)
        assert np.array_equal(vi, refi)
        assert np.array_equal(v2, ref2)
        assert np.array_equal(d16[vi], refd)
        rel = np.abs(p3 - ref3).max() / np.abs(ref3).max()
        assert np.allclose(p3, ref3, rtol=1e-4, atol=1e-4 * np.abs(ref3).max()), rel


def test_empty_and_out_of_roi(eng, oracle):
    l, r = util.frames(2, 1)[0]
    e = eng(1241, 376)
    f, cx, cy, b = synth.KITTI00.f32()
    pose = np.eye(4, dtype=np.float32)
    p3, v2, vi, d16 = e.triangulate_stereo(l, r, np.zeros((0, 2), np.float32), pose, f, cx, cy, b)
    assert p3.shape[0] == 0 and d16.shape[0] == 0
    pts = np.array([[0, 0], [1240, 375], [30, 100], [1235, 200]], np.float32)
    p3, v2, vi, d16 = e.triangulate_stereo(l, r, pts, pose, f, cx, cy, b)
    assert (d16 == -16).all() and p3.shape[0] == 0


def test_dedupe_matches_oracle(eng, oracle):
    e = eng(1241, 376)
    rng = np.random.default_rng(5)
    det = np.stack([rng.integers(0, 1241, 300), rng.integers(0, 376, 300)], 1).astype(np.float32)
    trk = (det[rng.permutation(300)[:150]] + rng.normal(size=(150, 2)) * 25).astype(np.float32)
    trk[:5] = det[:5] + np.float32([30.0, 0.0])      # exactly at the threshold: kept (not <)
    out, keep = e.dedupe(det, trk, 30.0)
    ref = oracle.dedupe(det, trk, 30.0)
    assert np.array_equal(keep, ref)
    assert np.array_equal(out, det[ref])
    out, keep = e.dedupe(det, np.zeros((0, 2), np.float32), 30.0)
    assert np.array_equal(keep, np.arange(300))


@pytest.mark.parametrize("ndisp,win", [(64, 15), (32, 9), (48, 11)])
def test_other_stereo_bm_parameters(eng, ndisp, win):
    """StereoBM::create(ndisp, win) other than the reference's (48, 21): the generic kernel (the
    reference's parameters run a specialised one). Same bar: every probed disparity equals
    cv2's dense map there."""
    import cv2
    w, h = 640, 480
    left = util.stress_image(h, w, 9)
    right = util.shifted(left, 13)
    e = eng(w, h, max_features=3000, max_corners=3000, num_disparities=ndisp, sad_window=win)
    rng = np.random.default_rng(ndisp + win)
    pts = np.stack([rng.integers(0, w, 3000), rng.integers(0, h, 3000)], 1).astype(np.float32)
    pose = np.eye(4, dtype=np.float32)
    f, cx, cy, b = synth.KITTI00.f32()
    _, _, _, d16 = e.triangulate_stereo(left, right, pts, pose, f, cx, cy, b)
    ref16 = cv2.StereoBM_create(ndisp, win).compute(left, right)
    want = ref16[pts[:, 1].astype(int), pts[:, 0].astype(int)]
    assert np.array_equal(d16, want), f"{(d16 != want).sum()} of {len(want)} disparities differ"
    assert (d16 > 0).sum() > 300
