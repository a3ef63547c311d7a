"""GPU parity of the whole front end, frame by frame, against the cv2 oracle's process()
(reference src/image_processor.cpp:18-162): teacher-forced (host PnP decisions replayed from the
oracle log) and closed-loop (cv2.solvePnPRansac on the GPU outputs)."""
import numpy as np
import pytest

from stereo_vo_b200 import synth
from tests import util

pytestmark = pytest.mark.gpu
FLOW_TOL = 0.0           # px: flow is bit-exact (window sums in cv2's own f32 order), at every frame
PAR_TOL = 1e-4           # av_parallax: an f32 mean of f32 norms


def _compare(recs_ref, recs_gpu, closed_loop=False):
    kf = 0
    for t, (a, b) in enumerate(zip(recs_ref, recs_gpu)):
        assert a["decision"] == b["decision"], f"frame {t}: {a['decision']} vs {b['decision']}"
        assert np.array_equal(a["detected"], b["detected"]), f"frame {t}: corners"
        if "av_parallax" in a:
            assert a["tracked_ids"] == b["tracked_ids"], f"frame {t}: track ids"
            d = float(np.abs(a["tracked_pts"] - b["tracked_pts"]).max()) if len(a["tracked_pts"]) else 0.0
            assert d <= FLOW_TOL, f"frame {t}: flow differs by {d} px"
            assert abs(float(a["av_parallax"]) - float(b["av_parallax"])) <= PAR_TOL
            assert a["percent_lost"] == b["percent_lost"]
        if a["decision"] in ("first", "keyframe"):
            kf += 1
            assert np.array_equal(a["new_features"], b["new_features"]), f"frame {t}: dedupe"
            assert np.array_equal(a["valid_idx"], b["valid_idx"]), f"frame {t}: match indices"
            assert np.array_equal(a["disp16"], b["disp16"][b["valid_idx"]]), f"frame {t}: disp16"
            assert np.array_equal(a["new2d"], b["new2d"])
            assert a["new_ids"] == b["new_ids"] and a["tracker_ids"] == b["tracker_ids"]
            tol = 1e-4 if not closed_loop else 1e-3
            scale = max(1.0, float(np.abs(a["new3d"]).max())) if len(a["new3d"]) else 1.0
            assert np.abs(a["new3d"] - b["new3d"]).max() <= tol * scale if len(a["new3d"]) else True
    return kf


@pytest.mark.parametrize("cam,motion,seed,n", [
    # run length: on the synthetic sequences the reference's own keyframe bookkeeping reaches
    # its undefined 400-feature truncation (src/bundle_adjuster.cpp:75,98-116) from
    # frame 89 on (KITTI seed 0; other seeds 92..200) - 80 frames stay
    # inside the range where the reference is defined
    (synth.KITTI00, synth.KITTI_MOTION, 0, 80),
    (synth.D435I, synth.D435I_MOTION, 1, 80),
])
def test_teacher_forced_sequence(svo_lib, oracle, cam, motion, seed, n):
    from stereo_vo_b200 import frontend
    from stereo_vo_b200.capi import Engine
    fr = synth.sequence(seed, n, cam, motion)
    ref = oracle.make_pipeline(cam)
    for l, r in fr:
        ref.process(l, r)
    e = Engine(cam.width, cam.height, n_seq=1)
    f, cx, cy, b = cam.f32()
    gpu = frontend.ImageProcessor(e, frontend.camera_matrix(f, cx, cy), frontend.BundleAdjuster(), b,
                                  teacher=ref.log)
    for l, r in fr:
        gpu.process(l, r)
    kf = _compare(ref.log, gpu.log)
    assert kf >= 2
    e.close()


def test_closed_loop_sequence(svo_lib, oracle):
    """No teacher: PnP runs on the GPU-produced tracks. Flow differs from cv2 by <= 1e-4 px, so
    RANSAC sees (numerically) different inputs; decisions, ids and integer outputs must still
    agree and the pose must stay on the synthetic ground truth."""
    from stereo_vo_b200 import frontend
    from stereo_vo_b200.capi import Engine
    cam = synth.KITTI00
    fr = util.frames(0, 48)
    e = Engine(cam.width, cam.height, n_seq=1)
    f, cx, cy, b = cam.f32()
    gpu = frontend.ImageProcessor(e, frontend.camera_matrix(f, cx, cy), frontend.BundleAdjuster(), b)
    for l, r in fr:
        gpu.process(l, r)
    kfs = [(t, r) for t, r in enumerate(gpu.log) if r["decision"] == "keyframe"]
    assert len(kfs) >= 2
    for t, r in kfs:
        gt = np.array(synth.ground_truth(t))
        assert np.abs(r["hmat"][:3, 3] - gt).max() < 0.05, (t, r["hmat"][:3, 3], gt)
    e.close()


def test_batched_sequences_match_single(svo_lib, oracle):
    """4 sequences stepped together through the batched ABI == each one processed alone."""
    from stereo_vo_b200 import frontend
    from stereo_vo_b200.capi import Engine
    cam = synth.KITTI00
    n_seq, n_fr = 4, 6
    seqs = [util.frames(s, n_fr) for s in range(n_seq)]
    e = Engine(cam.width, cam.height, n_seq=n_seq)
    bf = frontend.BatchFrontEnd(e, cam)
    logs = []
    for t in range(n_fr):
        res, kfs = bf.step([seqs[s][t][0] for s in range(n_seq)], [seqs[s][t][1] for s in range(n_seq)])
        logs.append((res, list(kfs)))
    e.close()
    for s in range(n_seq):
        e1 = Engine(cam.width, cam.height, n_seq=1)
        b1 = frontend.BatchFrontEnd(e1, cam)
        for t in range(n_fr):
            res, kfs = b1.step([seqs[s][t][0]], [seqs[s][t][1]])
            rb, kb = logs[t]
            assert (s in kb) == (0 in kfs)
            n = res["n_corners"][0]
            assert n == rb["n_corners"][s]
            assert np.array_equal(res["corners"][0, :n], rb["corners"][s, :n])
            assert res["stats"][0].tobytes() == rb["stats"][s].tobytes()
            k = res["stats"][0]["n_kept"]
            assert np.array_equal(res["tracked"][0, :k], rb["tracked"][s, :k])
            assert np.array_equal(res["tracked_ids"][0, :k], rb["tracked_ids"][s, :k])
        e1.close()


def test_keyframe_tracked_equals_keyframe_plus_tracker_set(svo_lib):
    """svo_keyframe_tracked (inliers = the tracker's device-resident set, tracker re-initialised
    on the device) == svo_keyframe with the tracked list sent from the host + svo_tracker_set,
    over a run with a keyframe every third frame, two sequences per engine."""
    from stereo_vo_b200.capi import Engine
    cam = synth.KITTI00
    n_seq, n_fr = 2, 8
    seqs = [util.frames(10 + s, n_fr) for s in range(n_seq)]
    ea = Engine(cam.width, cam.height, n_seq=n_seq)
    eb = Engine(cam.width, cam.height, n_seq=n_seq)
    ids_all = list(range(n_seq))
    next_id = [1000 * (s + 1) for s in range(n_seq)]
    mf = ea.max_features
    for t in range(n_fr):
        left = [seqs[s][t][0] for s in range(n_seq)]
        right = [seqs[s][t][1] for s in range(n_seq)]
        ra, rb = ea.step(ids_all, left, right), eb.step(ids_all, left, right)
        for key in ("corners", "n_corners", "tracked", "tracked_ids"):
            assert np.array_equal(ra[key], rb[key]), (t, key)
        assert ra["stats"].tobytes() == rb["stats"].tobytes(), t
        if t % 3 != 0:
            continue
        kept = [int(ra["stats"][s]["n_kept"]) if t > 0 else 0 for s in range(n_seq)]
        poses = np.tile(np.eye(4, dtype=np.float32).reshape(1, 16), (n_seq, 1))
        poses[:, 3] = 0.1 * t
        ka = ea.keyframe(ids_all, [ra["tracked"][s, :kept[s]] for s in range(n_seq)], kept, poses,
                         cam.focal, cam.cx, cam.cy, cam.baseline)
        feats, ids = [], []
        for s in range(n_seq):
            nv = min(int(ka["n_valid"][s]), mf - kept[s])
            feats.append(np.concatenate([ra["tracked"][s, :kept[s]], ka["valid_2d"][s, :nv]]))
            ids.append(np.concatenate([ra["tracked_ids"][s, :kept[s]],
                                       np.arange(next_id[s], next_id[s] + nv, dtype=np.uint64)]))
        ea.tracker_set(ids_all, feats, ids)
        kb = eb.keyframe_tracked(ids_all, poses, next_id, cam.focal, cam.cx, cam.cy, cam.baseline)
        for key in ka:
            assert np.array_equal(ka[key], kb[key]), (t, key)
        for s in range(n_seq):
            fa, ia = ea.tracker_get(s)
            fb, ib = eb.tracker_get(s)
            assert np.array_equal(fa, feats[s]) and np.array_equal(ia, ids[s])
            assert np.array_equal(fa, fb) and np.array_equal(ia, ib), (t, s)
            next_id[s] += len(ids[s]) - kept[s]
    ea.close()
    eb.close()


def test_too_few_corners_frame_is_skipped_entirely(svo_lib, oracle):
    """src/image_processor.cpp:23-25: a frame with fewer than 4 corners returns before the
    tracker is touched, so the next good frame is tracked from the last good one. Featureless
    frames in the middle of a run (also as the very first frame), python facade and C++ batch
    driver, against the oracle."""
    from stereo_vo_b200 import frontend, hostapi
    from stereo_vo_b200.capi import Engine
    cam = synth.KITTI00
    fr = list(synth.sequence(3, 9, cam, synth.KITTI_MOTION))
    flat = np.full((cam.height, cam.width), 120, np.uint8)
    fr[0] = (flat, flat)
    fr[4] = (flat, flat.copy())
    fr[5] = (flat.copy(), flat)
    ref = oracle.make_pipeline(cam)
    for l, r in fr:
        ref.process(l, r)
    assert [ref.log[t]["decision"] for t in (0, 1, 4, 5)] == ["too_few", "first", "too_few", "too_few"]
    e = Engine(cam.width, cam.height, n_seq=1)
    f, cx, cy, b = cam.f32()
    gpu = frontend.ImageProcessor(e, frontend.camera_matrix(f, cx, cy), frontend.BundleAdjuster(), b,
                                  teacher=ref.log)
    for l, r in fr:
        gpu.process(l, r)
    _compare(ref.log, gpu.log)
    assert sum(r["decision"] == "too_few" for r in gpu.log) == 3
    e.close()

    # C++ batch driver: "skipped entirely" means the good frames give exactly what they give
    # when the featureless frames are not there at all
    good = [(l, r) for t, (l, r) in enumerate(fr) if t not in (0, 4, 5)]
    with_flat = hostapi.Batch(cam.width, cam.height, 2, 2, cam)
    without = hostapi.Batch(cam.width, cam.height, 2, 1, cam)
    pitch = cam.width
    k = 0
    for t, (l, r) in enumerate(fr):
        with_flat.step([l, l], [r, r], pitch)
        a = with_flat.get(1)
        if t in (0, 4, 5):
            assert a["decision"] == "too_few" and len(a["corners"]) == 0, t
            continue
        gl, gr = good[k]
        k += 1
        without.step([gl, gl], [gr, gr], pitch)
        b = without.get(1)
        assert a["decision"] == b["decision"], t
        assert np.array_equal(a["corners"], b["corners"]), t
        assert a["stats"].tobytes() == b["stats"].tobytes(), t
        assert np.array_equal(a["tracker_pts"], b["tracker_pts"]), t
        assert np.array_equal(a["tracker_ids"], b["tracker_ids"]), t
        assert np.array_equal(a["new3d"], b["new3d"]), t
    assert k == len(good)
    with_flat.close()
    without.close()


def test_step_keyframe_one_call_equals_step_then_keyframe(svo_lib):
    """svo_step_keyframe (decision taken on the device, keyframe stage in the same call) against
    svo_step + the same decision on the host + svo_keyframe_tracked, frame by frame. Sequence 1
    never gets its right image with the call: its keyframes come back PENDING and are finished
    with svo_set_right + svo_keyframe_tracked."""
    import ctypes as C
    from stereo_vo_b200.capi import Engine, _ptr
    cam = synth.KITTI00
    f, cx, cy, b = cam.f32()
    n_seq, n_fr, thresh = 2, 12, 20.0
    seqs = [util.frames(20 + s, n_fr) for s in range(n_seq)]
    ea = Engine(cam.width, cam.height, n_seq=n_seq)      # separate calls
    eb = Engine(cam.width, cam.height, n_seq=n_seq)      # one call
    ids_all = list(range(n_seq))
    next_a = [100, 5000]
    next_b = [100, 5000]
    decisions = []

    def set_right(e, seq, img):
        s = np.asarray([seq], np.int32)
        arr = (C.c_void_p * 1)(img.ctypes.data)
        e._check(e.lib.svo_set_right(e.h, 1, _ptr(s), arr, int(img.strides[0]), 0), "svo_set_right")

    for t in range(n_fr):
        left = [seqs[s][t][0] for s in range(n_seq)]
        right = [seqs[s][t][1] for s in range(n_seq)]
        poses = np.tile(np.eye(4, dtype=np.float32).reshape(1, 16), (n_seq, 1))
        poses[:, 3] = 0.05 * t
        ra = ea.step(ids_all, left, right)
        rb = eb.step_keyframe(ids_all, left, [right[0], None], poses, next_b, thresh, f, cx, cy, b)
        for key in ("corners", "n_corners", "tracked", "tracked_ids"):
            assert np.array_equal(ra[key], rb[key]), (t, key)
        assert ra["stats"].tobytes() == rb["stats"].tobytes(), t
        for s in range(n_seq):
            st = ra["stats"][s]
            if ra["n_corners"][s] < 4:
                want = 1
            elif t == 0:
                want = 2
            elif st["av_parallax"] <= np.float32(thresh) and float(st["percent_lost"]) < 0.4:
                want = 3
            else:
                want = 4
            got = int(rb["decision"][s])
            assert (got & 7) == want, (t, s, got, want)
            assert bool(got & 8) == (want in (2, 4) and s == 1), (t, s, got)
            decisions.append(want)
            if want not in (2, 4):
                continue
            ka = ea.keyframe_tracked([s], poses[s:s + 1], [next_a[s]], f, cx, cy, b)
            if got & 8:
                set_right(eb, s, right[s])
                kb = eb.keyframe_tracked([s], poses[s:s + 1], [next_b[s]], f, cx, cy, b)
                kb = {k: v[0] for k, v in kb.items()}
            else:
                kb = {k: rb[k][s] for k in ka}
            for k in ka:
                assert np.array_equal(ka[k][0], kb[k]), (t, s, k)
            kept = int(st["n_kept"]) if want == 4 else 0
            nv = min(int(ka["n_valid"][0]), ea.max_features - kept)
            next_a[s] += nv
            next_b[s] += nv
            fa, ia = ea.tracker_get(s)
            fb, ib = eb.tracker_get(s)
            assert np.array_equal(fa, fb) and np.array_equal(ia, ib), (t, s)
    assert decisions.count(4) >= 2 and decisions.count(3) >= 4
    ea.close()
    eb.close()
