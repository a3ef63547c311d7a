"""GPU parity: svo_stereo_bm_dense against cv2.StereoBM_create(48, 21).compute (the call the
reference makes at src/image_processor.cpp:173-176) - every pixel of the CV_16S map bit-exact,
borders and rejected pixels (-16) included."""
import numpy as np
import pytest

from tests import util

pytestmark = pytest.mark.gpu


def _engine(w, h, n_seq=1):
    from stereo_vo_b200.capi import Engine
    return Engine(w, h, n_seq=n_seq)


def _compare(got, want):
    assert got.shape == want.shape and got.dtype == np.int16
    bad = np.argwhere(got != want)
    assert len(bad) == 0, (f"{len(bad)} of {want.size} pixels differ; first (y, x) {bad[:5].tolist()} "
                           f"got {[int(got[y, x]) for y, x in bad[:5]]} "
                           f"want {[int(want[y, x]) for y, x in bad[:5]]}")


def test_dense_map_kitti_frames(svo_lib, oracle):
    e = _engine(1241, 376)
    try:
        for seed in (0, 3):
            l, r = util.frames(seed, 1)[0]
            want = oracle.stereo_disparity16(l, r)
            got = e.stereo_bm_dense(l, r)
            _compare(got, want)
            assert (got > 0).mean() > 0.5          # the synthetic scene is matchable
            assert (got[:, :57] == -16).all() and (got[:10] == -16).all()
    finally:
        e.close()


@pytest.mark.parametrize("w,h", [(640, 480), (641, 479), (333, 211), (100, 45), (68, 32), (69, 33)])
def test_dense_map_stress_images(svo_lib, oracle, w, h):
    """flat patches / periodic stripes (texture + uniqueness rejections), odd and even heights
    (the odd-height last prefilter row), widths that leave a partial tile, down to one valid
    column (w = 68). (Narrower images have no valid pixel; cv2 leaves parts of its output
    uninitialised there, so there is nothing to compare with.)"""
    left = util.stress_image(h, w, 3)
    right = util.shifted(left, 17)
    want = oracle.stereo_disparity16(left, right)
    e = _engine(w, h)
    try:
        got = e.stereo_bm_dense(left, right)
    finally:
        e.close()
    _compare(got, want)
    if w >= 333:
        roi = got[10:h - 10, 57:w - 10]
        assert (roi == -16).sum() > 500 and (roi > 0).sum() > 500


def test_dense_map_random_noise_pair(svo_lib, oracle):
    """unrelated noise images: near-ties between disparities everywhere, so the first-minimum
    order, the uniqueness margin and the sub-pixel ends (winner at d = 0 / 47) all get hit."""
    rng = np.random.default_rng(7)
    left = rng.integers(0, 256, (200, 400)).astype(np.uint8)
    right = rng.integers(0, 256, (200, 400)).astype(np.uint8)
    # low-contrast variant: prefiltered values stay inside the cap
    left2 = (128 + (left.astype(np.int32) - 128) // 16).astype(np.uint8)
    right2 = (128 + (right.astype(np.int32) - 128) // 16).astype(np.uint8)
    e = _engine(400, 200)
    try:
        for l, r in ((left, right), (left2, right2), (left2, util.shifted(left2, 47)),
                     (left2, left2)):
            _compare(e.stereo_bm_dense(l, r), oracle.stereo_disparity16(l, r))
    finally:
        e.close()


def test_dense_resident_batch_and_sparse_agree(svo_lib, oracle):
    """three sequences in one launch on pairs that are already on the device; the sparse stage
    (svo_triangulate_stereo) reads the same values at its key-points."""
    from stereo_vo_b200 import synth
    pairs = [util.frames(s, 1)[0] for s in (1, 2, 4)]
    e = _engine(1241, 376, n_seq=3)
    try:
        for s, (l, r) in enumerate(pairs):
            _compare(e.stereo_bm_dense(l, r, seq=s), oracle.stereo_disparity16(l, r))
        order = [2, 0, 1]
        maps = e.stereo_bm_dense_resident(order, shape=(376, 1241))
        for s, m in zip(order, maps):
            _compare(m, oracle.stereo_disparity16(*pairs[s]))
        assert e.stereo_bm_dense_resident(order, fetch=False) is None
        f, cx, cy, b = synth.KITTI00.f32()
        rng = np.random.default_rng(5)
        pts = np.stack([rng.integers(0, 1241, 300), rng.integers(0, 376, 300)], 1).astype(np.float32)
        _, _, _, d16 = e.triangulate_stereo(pairs[0][0], pairs[0][1], pts, np.eye(4, dtype=np.float32),
                                            f, cx, cy, b, seq=0)
        m0 = e.stereo_bm_dense_resident([0], shape=(376, 1241))[0]
        assert np.array_equal(d16, m0[pts[:, 1].astype(int), pts[:, 0].astype(int)])
    finally:
        e.close()


def test_dense_many_sequences_tall_bands(svo_lib, oracle):
    """12 KITTI pairs in one launch: enough blocks for the launcher to pick the tall row bands."""
    pairs = [util.frames(s, 1)[0] for s in (0, 5, 6)]
    want = [oracle.stereo_disparity16(l, r) for l, r in pairs]
    e = _engine(1241, 376, n_seq=12)
    try:
        for s in range(12):
            _compare(e.stereo_bm_dense(*pairs[s % 3], seq=s), want[s % 3])
        order = list(range(11, -1, -1))
        for s, m in zip(order, e.stereo_bm_dense_resident(order, shape=(376, 1241))):
            _compare(m, want[s % 3])
    finally:
        e.close()


def test_dense_tall_bands_odd_height_partial_last_band(svo_lib, oracle):
    """12 pairs of 641x479 in one launch: tall bands of 156 rows, the last one 147 (not a multiple
    of the 4-row groups), odd height (capped last prefilter row), a partial last column tile."""
    w, h = 641, 479
    lefts = [util.stress_image(h, w, seed) for seed in (11, 12)]
    pairs = [(l, util.shifted(l, d)) for l, d in zip(lefts, (9, 31))]
    want = [oracle.stereo_disparity16(l, r) for l, r in pairs]
    e = _engine(w, h, n_seq=12)
    try:
        for s in range(12):
            _compare(e.stereo_bm_dense(*pairs[s % 2], seq=s), want[s % 2])
        order = list(range(12))
        for s, m in zip(order, e.stereo_bm_dense_resident(order, shape=(h, w))):
            _compare(m, want[s % 2])
    finally:
        e.close()


def test_dense_map_uhd(svo_lib, oracle):
    """3840x2160 (BASELINE configs[4] size): 8.3 M pixels, 119 column tiles x 12 tall bands."""
    w, h = 3840, 2160
    left = util.stress_image(h, w, 21)
    right = util.shifted(left, 23)
    want = oracle.stereo_disparity16(left, right)
    e = _engine(w, h)
    try:
        got = e.stereo_bm_dense(left, right)
    finally:
        e.close()
    _compare(got, want)
    assert (got > 0).mean() > 0.5


def test_dense_needs_a_stereo_pair(svo_lib):
    from stereo_vo_b200.capi import SvoError
    e = _engine(640, 480)
    try:
        with pytest.raises(SvoError):
            e.stereo_bm_dense_resident([0], shape=(480, 640))
    finally:
        e.close()
