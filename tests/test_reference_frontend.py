"""CPU: the CONTROL FLOW of the oracle pinned by the reference's own compiled code.

oracle/_ref/libref_frontend.so is /root/reference/src/image_processor.cpp + feature_tracker.cpp
(+ bundle_adjuster.cpp, reprojection_factor.cpp) compiled where they lie; the OpenCV algorithms
they call are served by the real OpenCV through callbacks (oracle/build_ref.py,
oracle/ref_frontend.py). tests/golden/ref_frontend.npz is what that build did on two synthetic
sequences (tests/golden/make_golden_ref_frontend.py). Here:

  * the Python restatement (oracle/cv2_frontend.py) is run on the same frames with the pose step
    replayed from the reference's log (RANSAC amplifies last-bit differences of its inputs, so
    it is taken out of the loop; its inputs are compared instead) and must do, frame by frame,
    exactly what the compiled reference did: same frames skipped / made keyframes, same tracker
    ids and sub-pixel positions, same features handed to LK, same image points (and world points
    to 1e-5) handed to PnP, same new features and ids, 3-D points to 1e-5 relative (the reference computes hmat in
    f32 OpenCV calls, the oracle in numpy);
  * the literal arguments the reference's code hands to OpenCV are the constants of the oracle
    and the defaults of the product;
  * with the reference tree on this machine: the fixture is reproduced from a live run, and a
    closed-loop run (nothing replayed) agrees on every discrete output for 30 frames.
"""
import os

import numpy as np
import pytest

from oracle import build_ref, cv2_frontend as O
from stereo_vo_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURE = os.path.join(HERE, "golden", "ref_frontend.npz")
CASES = {"kitti_seed1": (synth.KITTI00, synth.KITTI_MOTION, 1),
         "d435i_seed3": (synth.D435I, synth.D435I_MOTION, 3)}


@pytest.fixture(scope="module")
def G():
    return np.load(FIXTURE)


def _frame(G, case, t):
    prefix = f"{case}/{t}/"
    return {k[len(prefix):]: G[k] for k in G.files if k.startswith(prefix)}


@pytest.mark.parametrize("case", list(CASES))
def test_oracle_does_what_the_compiled_reference_did(G, case, monkeypatch):
    import cv2
    cam, motion, seed = CASES[case]
    n = int(G[f"{case}/n_frames"])
    ref = [_frame(G, case, t) for t in range(n)]
    O.canonical_mode()
    state = {"t": 0}

    def replay_pnp(world, pts2d, cam_matrix, rvec, tvec):
        r = ref[state["t"]]
        # the oracle hands PnP exactly what the reference handed it
        # (world points triangulated after the first keyframe carry the 1e-7 of the hmat arithmetic)
        assert np.allclose(np.asarray(world, np.float32).reshape(-1, 3), r["pnp_world"], rtol=1e-5, atol=1e-5), state["t"]
        assert np.array_equal(np.asarray(pts2d, np.float32).reshape(-1, 2), r["pnp_image"]), state["t"]
        # the reference keeps rvec / tvec as CV_32F and takes Rodrigues of that (OpenCV 3.x types)
        rv32 = r["pnp_rvec"].astype(np.float32).reshape(3, 1)
        rmat, _ = cv2.Rodrigues(rv32)
        return (r["pnp_rvec"].reshape(3, 1).copy(), r["pnp_tvec"].astype(np.float32).astype(np.float64).reshape(3, 1),
                rmat, r["pnp_inliers"].astype(np.int32))

    monkeypatch.setattr(O, "solve_pnp", replay_pnp)
    orc = O.make_pipeline(cam)
    keyframes = 0
    worst = 0.0
    for t, (l, r) in enumerate(synth.sequence(seed, n, cam, motion)):
        state["t"] = t
        o, a = orc.process(l, r), ref[t]
        assert (o["decision"] in ("first", "keyframe")) == bool(a["made_keyframe"]), f"frame {t}: {o['decision']}"
        assert np.array_equal(o["detected"], a["detected"]), f"frame {t}: corners"
        # which OpenCV calls the reference made on this frame: gftt, LK (fwd + bwd), PnP, StereoBM
        want = {"first": [1, 0, 0, 1], "skip": [1, 2, 0, 0], "keyframe": [1, 2, 1, 1]}[o["decision"]]
        assert list(a["n_calls"]) == want, f"frame {t}: {o['decision']} {a['n_calls']}"
        if o["decision"] != "first":
            assert np.array_equal(o["prev_pts"], a["lk_in"]), f"frame {t}: features handed to LK"
            assert np.array_equal(o["taps"]["tracked"], a["lk_out"]) and np.array_equal(o["taps"]["status1"], a["lk_status"])
        ids = o["tracker_ids"] if "tracker_ids" in o else o["tracked_ids"]
        pts = o["tracker_pts"] if "tracker_pts" in o else o["tracked_pts"]
        assert [int(i) for i in ids] == [int(i) for i in a["tracker_ids"]], f"frame {t}: tracker ids"
        assert np.array_equal(pts, a["tracker_pts"]), f"frame {t}: tracker positions"
        if a["made_keyframe"]:
            keyframes += 1
            assert np.array_equal(o["new2d"], a["new2d"]), f"frame {t}: new features"
            assert [int(i) for i in o["new_ids"]] == [int(i) for i in a["new_ids"]], f"frame {t}: new ids"
            assert a["new3d"].shape == o["new3d"].shape
            if a["new3d"].size:
                worst = max(worst, float(np.abs(a["new3d"] - o["new3d"]).max() / np.abs(a["new3d"]).max()))
            if o["decision"] == "keyframe":
                assert np.array_equal(o["inliers"], a["pnp_inliers"])
                kept = [int(i) for i in a["kf_tracked_ids"]]
                assert kept == [int(ids[k]) for k in range(len(kept))], f"frame {t}: inlier ids lead the tracker set"
    assert keyframes >= 4 and worst <= 1e-5, (keyframes, worst)


def test_the_reference_hands_opencv_the_constants_we_use(G):
    from stereo_vo_b200 import capi
    p = capi.default_params(1241, 376)
    for case in CASES:
        a = G[f"{case}/args"]
        gftt, lk, sbm, pnp = a[0:3], a[3:11], a[11:13], a[13:18]
        assert tuple(gftt) == (O.MAX_CORNERS, O.QUALITY_LEVEL, O.MIN_FEATURE_DISTANCE) == (300, 0.1, 30.0)
        assert tuple(gftt) == (p.max_corners, p.quality_level, p.min_distance)
        # winSize (21, 21), maxLevel 3, criteria (COUNT + EPS = 3, 30, 0.01), flags 0, minEigThreshold 1e-2
        assert tuple(lk) == (21, 21, 3, 3, 30, 0.01, 0, 1e-2)
        assert (p.lk_max_level, p.lk_max_iters, p.lk_epsilon, p.lk_min_eig_threshold) == (3, 30, 0.01, 1e-2)
        assert tuple(sbm) == (O.NUM_DISPARITIES, O.BLOCK_SIZE) == (48, 21) == (p.num_disparities, p.sad_window)
        # useExtrinsicGuess, 100 iterations, 8 px, 0.99; rvec / tvec arrive as CV_32F (depth 5)
        assert tuple(pnp) == (1, 100, 8.0, 0.99, 5)


@pytest.mark.skipif(not build_ref.available(), reason="the reference tree is not on this machine")
def test_fixture_is_what_the_compiled_reference_does(G):
    import importlib.util
    spec = importlib.util.spec_from_file_location("mk", os.path.join(HERE, "golden", "make_golden_ref_frontend.py"))
    mk = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mk)
    cam, motion, seed = CASES["d435i_seed3"]
    n = int(G["d435i_seed3/n_frames"])
    log, _ = mk.run_reference(cam, motion, seed, n)
    for t, rec in enumerate(log):
        for k, v in rec.items():
            assert np.array_equal(np.asarray(v), G[f"d435i_seed3/{t}/{k}"]), (t, k)


@pytest.mark.skipif(not build_ref.available(), reason="the reference tree is not on this machine")
def test_closed_loop_agrees_with_the_compiled_reference():
    """nothing replayed: both sides call cv2.solvePnPRansac themselves"""
    from oracle import ref_frontend
    O.canonical_mode()
    cam = synth.KITTI00
    f, cx, cy, b = cam.f32()
    ref = ref_frontend.RefPipeline(f, cx, cy, b)
    orc = O.make_pipeline(cam)
    keyframes = 0
    for t, (l, r) in enumerate(synth.sequence(2, 30)):
        a, o = ref.process(l, r, t / 11.0), orc.process(l, r)
        made = o["decision"] in ("first", "keyframe")
        assert made == a["made_keyframe"], t
        ids = o["tracker_ids"] if "tracker_ids" in o else o["tracked_ids"]
        pts = o["tracker_pts"] if "tracker_pts" in o else o["tracked_pts"]
        assert [int(i) for i in ids] == [int(i) for i in a["tracker_ids"]], t
        assert np.array_equal(pts, a["tracker_pts"]), t
        if made:
            keyframes += 1
            assert np.array_equal(o["new2d"], a["new2d"]) and [int(i) for i in o["new_ids"]] == [int(i) for i in a["new_ids"]]
            assert np.allclose(o["new3d"], a["new3d"], rtol=1e-5, atol=1e-4)
    ref.close()
    assert keyframes >= 6


@pytest.mark.skipif(not build_ref.available(), reason="the reference tree is not on this machine")
@pytest.mark.parametrize("name", ["kitti_seq0", "d435i_seq1"])
def test_compiled_reference_reproduces_the_gpu_golden_logs(name):
    """tests/golden/<name>.npz is what tests/test_gpu_golden.py replays on the CUDA path. It was
    written by the Python oracle; here the reference's own compiled front end (closed loop,
    nothing replayed) walks the same frames and must arrive at the same log: decisions, corners,
    tracked set, forward-LK status, new features, ids, disparities-implied valid set, inliers,
    3-D points to 1e-5. GPU == golden and golden == compiled reference close the chain."""
    from oracle import ref_frontend
    from tests import golden_util as GU
    recs, _ = GU.load(name)
    frames, cam = GU.frames_for(name, len(recs))
    GU.check_inputs(recs, frames)
    O.canonical_mode()
    f, cx, cy, b = cam.f32()
    ref = ref_frontend.RefPipeline(f, cx, cy, b)
    for t, ((l, r), g) in enumerate(zip(frames, recs)):
        a = ref.process(l, r, t / 11.0)
        calls = {}
        for cname, c in a["calls"]:
            calls.setdefault(cname, []).append(c)
        assert a["made_keyframe"] == (g["decision"] in ("first", "keyframe")), (t, g["decision"])
        assert np.array_equal(calls["goodFeaturesToTrack"][0]["corners"], g["detected"]), t
        if "tracked_pts" in g:
            assert np.array_equal(calls["calcOpticalFlowPyrLK"][0]["status"], g["status1"]), t
        if g["decision"] == "skip":
            assert [int(i) for i in a["tracker_ids"]] == g["tracked_ids"], t
            assert np.array_equal(a["tracker_pts"], g["tracked_pts"]), t
        if a["made_keyframe"]:
            assert [int(i) for i in a["tracker_ids"]] == g["tracker_ids"], t
            assert np.array_equal(a["tracker_pts"], g["tracker_pts"]), t
            assert np.array_equal(a["new2d"], g["new2d"]) and [int(i) for i in a["new_ids"]] == g["new_ids"], t
            assert np.allclose(a["new3d"], g["new3d"], rtol=1e-5, atol=1e-4), t
            # the disparity map the reference's StereoBM call produced, read where the golden log reads it
            d16 = calls["StereoBM"][0]["disp16"]
            nf = g["new_features"].astype(np.int32)
            at_features = d16[nf[:, 1], nf[:, 0]]
            assert np.array_equal(np.flatnonzero(at_features > 0), g["valid_idx"]), t
            assert np.array_equal(at_features[g["valid_idx"]], g["disp16"]), t
        if g["decision"] == "keyframe":
            assert np.array_equal(calls["solvePnPRansac"][0]["inliers"], g["inliers"]), t
    ref.close()


@pytest.mark.skipif(not build_ref.available(), reason="the reference tree is not on this machine")
def test_edge_frames_go_the_same_way_in_the_compiled_reference():
    """frames the synthetic sequences never produce: a flat image (fewer than 4 corners: the
    reference returns before touching the tracker, image_processor.cpp:23-25), a flat image as the
    very first frame, a frame whose tracks are all lost (large jump: percent_lost = 1 forces a
    keyframe with next to no inliers), and the same stereo pair twice (zero parallax: skip)."""
    from oracle import ref_frontend
    O.canonical_mode()
    cam = synth.KITTI00
    f, cx, cy, b = cam.f32()
    seq = synth.sequence(4, 8)
    flat = np.full_like(seq[0][0], 127)
    jump = (np.roll(seq[5][0], 400, axis=1).copy(), np.roll(seq[5][1], 400, axis=1).copy())
    frames = [(flat, flat), seq[0], seq[1], (flat, flat), seq[2], seq[2], seq[3], jump, seq[6], seq[7]]
    ref = ref_frontend.RefPipeline(f, cx, cy, b)
    orc = O.make_pipeline(cam)
    decisions = []
    for t, (l, r) in enumerate(frames):
        a, o = ref.process(l, r, t / 11.0), orc.process(l, r)
        decisions.append(o["decision"])
        names = [c[0] for c in a["calls"]]
        assert a["made_keyframe"] == (o["decision"] in ("first", "keyframe")), (t, o["decision"])
        if o["decision"] == "too_few":
            assert names == ["goodFeaturesToTrack"], (t, names)          # nothing else was called
        if o["decision"] == "skip":
            assert names == ["goodFeaturesToTrack", "calcOpticalFlowPyrLK", "calcOpticalFlowPyrLK"], (t, names)
        ids = o.get("tracker_ids", o.get("tracked_ids"))
        pts = o.get("tracker_pts", o.get("tracked_pts"))
        if ids is not None:
            assert [int(i) for i in ids] == [int(i) for i in a["tracker_ids"]], t
            assert np.array_equal(np.asarray(pts).reshape(-1, 2), a["tracker_pts"]), t
        if a["made_keyframe"]:
            assert np.array_equal(o["new2d"], a["new2d"]) and [int(i) for i in o["new_ids"]] == [int(i) for i in a["new_ids"]], t
            if o["decision"] == "keyframe":
                pn = [c for c in a["calls"] if c[0] == "solvePnPRansac"][0][1]
                assert np.array_equal(pn["inliers"], o["inliers"]), t
    ref.close()
    assert decisions[0] == "too_few" and decisions[1] == "first" and decisions[3] == "too_few"
    assert decisions[5] == "skip" and "keyframe" in decisions
