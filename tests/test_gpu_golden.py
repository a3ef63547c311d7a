"""CUDA path (through the C ABI) against the committed golden vectors - no oracle run here.
Keypoints, match indices, disp16 and track ids bit-exact; flow <= 0.01 px; 3-D <= 1e-4 rel."""
import numpy as np
import pytest

from tests import golden_util as G

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(G.CASES))
def test_gpu_reproduces_golden(svo_lib, name):
    from stereo_vo_b200 import frontend
    from stereo_vo_b200.capi import Engine
    recs, extra = G.load(name)
    frames, cam = G.frames_for(name, len(recs))
    G.check_inputs(recs, frames)
    e = Engine(cam.width, cam.height, n_seq=1, max_features=2000, max_corners=2000)
    f, cx, cy, b = cam.f32()
    p = extra["probe_pts"].astype(np.float32)
    _, _, _, d16 = e.triangulate_stereo(frames[0][0], frames[0][1], p, np.eye(4, dtype=np.float32),
                                        f, cx, cy, b)
    assert np.array_equal(d16, extra["probe_disp16"])
    e.close()

    e = Engine(cam.width, cam.height, n_seq=1, keep_eig_map=1)
    gpu = frontend.ImageProcessor(e, frontend.camera_matrix(f, cx, cy), frontend.BundleAdjuster(), b,
                                  teacher=recs)
    for t, (l, r) in enumerate(frames):
        rec = gpu.process(l, r)
        g = recs[t]
        assert rec["decision"] == g["decision"], t
        assert np.array_equal(rec["detected"], g["detected"]), f"frame {t}: corners"
        if t == 0:
            eig = e.debug_min_eigen_map()
            assert np.float32(eig.max()).view(np.uint32) == extra["eig_max_bits"]
            assert np.array_equal(eig[100].view(np.uint32), extra["eig_row100"].view(np.uint32))
        if "av_parallax" in g:
            assert rec["tracked_ids"] == g["tracked_ids"], f"frame {t}: track ids"
            assert np.array_equal(rec["tracked_pts"], g["tracked_pts"])      # flow is bit-exact
            assert abs(float(rec["av_parallax"]) - float(g["av_parallax"])) <= 1e-4
            assert rec["percent_lost"] == g["percent_lost"]
        if g["decision"] in ("first", "keyframe"):
            assert np.array_equal(rec["new_features"], g["new_features"])
            assert np.array_equal(rec["valid_idx"], g["valid_idx"])
            assert np.array_equal(rec["disp16"][rec["valid_idx"]], g["disp16"])
            assert rec["new_ids"] == g["new_ids"] and rec["tracker_ids"] == g["tracker_ids"]
            scale = max(1.0, float(np.abs(g["new3d"]).max())) if len(g["new3d"]) else 1.0
            if len(g["new3d"]):
                assert np.abs(rec["new3d"] - g["new3d"]).max() <= 1e-4 * scale
    e.close()
