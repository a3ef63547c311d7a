"""Pins oracle/restate.py (numpy restatement of the OpenCV arithmetic)
against cv2 - the functions the reference actually calls - bit for bit. CPU only."""
import cv2
import numpy as np
import pytest

from oracle import restate as R
from stereo_vo_b200 import synth
from tests import util


@pytest.fixture(scope="module")
def pair(oracle):
    cam = util.small_camera(420, 200)
    return synth.sequence(7, 3, cam), cam


@pytest.mark.parametrize("w,h,seed", [(420, 200, 0), (333, 211, 1), (1241, 376, 2)])
def test_min_eigen_map_bit_exact(oracle, w, h, seed):
    img = synth.stereo_frame(seed, 2, util.small_camera(w, h))[0]
    got = R.min_eigen_map(img)
    ref = cv2.cornerMinEigenVal(img, 3, ksize=3)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("w,h,seed", [(420, 200, 0), (640, 480, 3), (1241, 376, 4)])
def test_good_features_bit_exact(oracle, w, h, seed):
    img = synth.stereo_frame(seed, 1, util.small_camera(w, h))[0]
    assert np.array_equal(R.good_features_to_track(img), oracle.detect(img))
    assert np.array_equal(R.good_features_to_track(img, 80, 0.03, 7.5), oracle.detect(img, 80, 0.03, 7.5))


def test_stereo_bm_sparse_bit_exact(oracle):
    w, h = 420, 201                               # odd height: last prefilter row is the cap
    left = util.stress_image(h, w, 5)
    right = util.shifted(left, 13)
    d16 = oracle.stereo_disparity16(left, right)
    rng = np.random.default_rng(0)
    pts = np.stack([rng.integers(0, w, 1500), rng.integers(0, h, 1500)], 1)
    got = R.stereo_bm_sparse(left, right, pts)
    want = d16[pts[:, 1], pts[:, 0]]
    assert np.array_equal(got, want)
    assert (want == -16).sum() > 100 and (want > 0).sum() > 100


def test_pyramid_and_scharr_bit_exact(oracle, pair):
    (l, _), _ = pair[0][0], None
    n, pyr = cv2.buildOpticalFlowPyramid(l, (21, 21), 3, withDerivatives=True)
    lv = R.build_pyramid(l)
    assert len(lv) == n + 1
    for i, im in enumerate(lv):
        assert np.array_equal(pyr[2 * i], im)
        ix, iy = R.scharr_deriv(im)
        assert np.array_equal(pyr[2 * i + 1][..., 0], ix)
        assert np.array_equal(pyr[2 * i + 1][..., 1], iy)


def test_lk_matches_cv2(oracle, pair):
    frames, cam = pair
    prev, nxt = frames[0][0], frames[1][0]
    pts = oracle.detect(prev, 120, 0.05, 10.0)
    edge = np.array([[0, 0], [cam.width - 1, cam.height - 1], [2.5, 100.25]], np.float32)
    pts = np.concatenate([pts, edge])
    ref, rst = oracle.lk(prev, nxt, pts)
    # sub-pixel start points (the tracker's normal input): continue from the tracked positions
    nxt2 = frames[2][0]
    start = ref[rst > 0]
    ref2, rst2 = oracle.lk(nxt, nxt2, start)
    for a, b, p, want, wst in ((prev, nxt, pts, ref, rst), (nxt, nxt2, start, ref2, rst2)):
        got, st = R.lk_track(a, b, p)
        assert np.array_equal(st, wst)
        ok = wst > 0
        # window sums accumulated in cv2's own f32 order: bit-identical flow
        assert np.array_equal(got[ok].view(np.uint32), want[ok].view(np.uint32))
    # exact integer sums (the round-1 kernel): close, not identical
    got, st = R.lk_track(prev, nxt, pts, exact_sums=True)
    assert np.array_equal(st, rst) and np.abs(got - ref)[rst > 0].max() < 1e-3


def test_reproject_matches_gemm(oracle):
    f, cx, cy, b = synth.KITTI00.f32()
    rng = np.random.default_rng(1)
    pts = np.stack([rng.integers(60, 1200, 50), rng.integers(12, 360, 50)], 1).astype(np.float32)
    d16 = rng.integers(-16, 700, 50).astype(np.int16)
    pose = np.eye(4, dtype=np.float32)
    pose[:3, :3] = cv2.Rodrigues(np.array([0.1, -0.2, 0.05]))[0].astype(np.float32)
    pose[:3, 3] = [1.5, -0.2, 7.0]
    idx, p3 = R.reproject(pts, d16, pose, f, cx, cy, b)
    k = oracle.camera_matrix(f, cx, cy)
    # the oracle's triangulate on a fake disparity map
    disp = np.full((376, 1241), -16, np.int16)
    disp[pts[:, 1].astype(int), pts[:, 0].astype(int)] = d16
    import unittest.mock as mock
    with mock.patch.object(oracle, "stereo_disparity16", return_value=disp):
        ref3, _, refi, _ = oracle.triangulate_stereo(k, b, pts, None, None, pose)
    assert np.array_equal(idx, refi)
    assert np.allclose(p3, ref3, rtol=2e-6, atol=1e-5)
