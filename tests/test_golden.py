"""The cv2 oracle against the committed golden vectors (CPU). Pins the oracle across machines:
integer outputs (corners, decisions, ids, indices, disp16) bit-exact, floats to 1e-4."""
import cv2
import numpy as np
import pytest

from tests import golden_util as G


@pytest.mark.parametrize("name", list(G.CASES))
def test_oracle_reproduces_golden(oracle, name):
    recs, extra = G.load(name)
    frames, cam = G.frames_for(name, len(recs))
    G.check_inputs(recs, frames)
    proc = oracle.make_pipeline(cam)
    for t, (l, r) in enumerate(frames):
        rec = proc.process(l, r)
        g = recs[t]
        assert rec["decision"] == g["decision"], t
        assert np.array_equal(rec["detected"], g["detected"]), t
        if "av_parallax" in g:
            assert rec["tracked_ids"] == g["tracked_ids"]
            assert np.array_equal(rec["taps"]["status1"], g["status1"])
            assert np.array_equal(rec["taps"]["status2"], g["status2"])
            assert np.array_equal(rec["taps"]["kept_idx"], g["kept_idx"])
            assert np.allclose(rec["tracked_pts"], g["tracked_pts"], atol=1e-4)
            assert abs(float(rec["av_parallax"]) - float(g["av_parallax"])) < 1e-4
            assert rec["percent_lost"] == g["percent_lost"]
        if g["decision"] in ("first", "keyframe"):
            assert np.array_equal(rec["new_features"], g["new_features"])
            assert np.array_equal(rec["valid_idx"], g["valid_idx"])
            assert np.array_equal(rec["disp16"], g["disp16"])
            assert rec["new_ids"] == g["new_ids"] and rec["tracker_ids"] == g["tracker_ids"]
            assert np.allclose(rec["new3d"], g["new3d"], rtol=1e-4, atol=1e-3)
    d16 = oracle.stereo_disparity16(*frames[0])
    p = extra["probe_pts"]
    assert np.array_equal(d16[p[:, 1], p[:, 0]], extra["probe_disp16"])
    eig = cv2.cornerMinEigenVal(frames[0][0], 3, ksize=3)
    assert np.float32(eig.max()).view(np.uint32) == extra["eig_max_bits"]
    assert np.array_equal(eig[100].view(np.uint32), extra["eig_row100"].view(np.uint32))


def test_restatement_reproduces_golden(oracle):
    """The numpy restatement (no cv2 on the math path) against the same vectors."""
    from oracle import restate as R
    recs, extra = G.load("kitti_seq0")
    frames, cam = G.frames_for("kitti_seq0", 2)
    assert np.array_equal(R.good_features_to_track(frames[0][0]), recs[0]["detected"])
    eig = R.min_eigen_map(frames[0][0])
    assert np.array_equal(eig[100].view(np.uint32), extra["eig_row100"].view(np.uint32))
    p = extra["probe_pts"][:400]
    assert np.array_equal(R.stereo_bm_sparse(frames[0][0], frames[0][1], p), extra["probe_disp16"][:400])
    g = recs[1]
    pts = recs[0]["tracker_pts"]
    got, st = R.lk_track(frames[0][0], frames[1][0], pts)
    assert np.array_equal(st, g["status1"])
