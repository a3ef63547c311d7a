"""GPU pose step (svo_pnp_ransac) against cv2.solvePnPRansac called with
the reference's literal arguments (src/image_processor.cpp:75-92: useExtrinsicGuess = true, 100
iterations, 8 px, confidence 0.99) + cv2.Rodrigues.

Tolerance (north_star states none for the pose). OpenCV's 5-point samples
come from its own RNG and are fitted by EPnP, the kernel's from a counter hash and LM from the
extrinsic guess; both then apply the same consensus rule (inlier <=> error <= 8 px under the
winning 5-point hypothesis) and minimise the same reprojection cost over the inliers. Two
winning hypotheses differ by a few tenths of a pixel across the image, so:
  * judged by cv2's own FINAL pose, the kernel must accept every point with error < 8 - BORDER px
    that cv2 accepted, and reject every point with error > 8 + BORDER px. (cv2 itself sometimes
    rejects points that fit its final pose to 0.1 px: its winning EPnP hypothesis on 5 noisy,
    coplanar points can be off by > 8 px in a far corner of the image and the loop stops after
    1-2 samples at 99 % confidence. The kernel's LM-fitted hypotheses do not lose those points;
    that difference is reported, not hidden;)
  * identical inlier sets: translation within 1 mm, rotation within 1e-4 rad of cv2;
  * sets that differ in borderline points: 5 mm / 5e-4 rad (one 8-px residual more or less in a
    least-squares fit over ~250 points)."""
import numpy as np
import pytest

from stereo_vo_b200 import synth

pytestmark = pytest.mark.gpu
T_TOL, R_TOL = 1e-3, 1e-4
T_TOL_BORDER, R_TOL_BORDER, BORDER = 5e-3, 5e-4, 2.0


def _angle(ra, rb):
    d = ra @ rb.T
    return float(np.arccos(np.clip((np.trace(d) - 1.0) * 0.5, -1.0, 1.0)))


def _check_pose(res, inl, ref_rvec, ref_tvec, ref_inl, what, world=None, image=None, intr=None):
    import cv2
    assert res["ok"] == 1, what
    rref, _ = cv2.Rodrigues(np.asarray(ref_rvec, np.float64).reshape(3, 1))
    same = np.array_equal(inl, np.sort(ref_inl))
    if not same:
        # only points on the edge of the threshold may be classified differently
        assert world is not None, f"{what}: inlier sets differ"
        f, cx, cy = intr
        xc = world.astype(np.float64) @ rref.T + np.asarray(ref_tvec, np.float64).reshape(3)
        err = np.hypot(f * xc[:, 0] / xc[:, 2] + cx - image[:, 0], f * xc[:, 1] / xc[:, 2] + cy - image[:, 1])
        lost = np.setdiff1d(ref_inl, inl)            # cv2 inliers the kernel rejected
        extra = np.setdiff1d(inl, ref_inl)           # kernel inliers cv2 rejected
        assert np.all(err[lost] >= 8.0 - BORDER), f"{what}: lost inliers {lost} with errors {err[lost]} px"
        assert np.all(err[extra] <= 8.0 + BORDER), f"{what}: accepted {extra} with errors {err[extra]} px"
        assert len(lost) + len(extra) <= max(3, len(ref_inl) // 40), f"{what}: {len(lost)} + {len(extra)} points differ"
    rgpu = res["rmat"].reshape(3, 3)
    r2, _ = cv2.Rodrigues(res["rvec"].reshape(3, 1))
    assert np.abs(r2 - rgpu).max() < 1e-12, f"{what}: rmat is not Rodrigues(rvec)"
    dt = float(np.abs(res["tvec"] - np.asarray(ref_tvec, np.float64).reshape(3)).max())
    da = _angle(rgpu, rref)
    tt, rt = (T_TOL, R_TOL) if same else (T_TOL_BORDER, R_TOL_BORDER)
    assert dt <= tt and da <= rt, f"{what}: dt {dt} m, rotation {da} rad (inlier sets {'equal' if same else 'differ'})"
    return dt, da, same


@pytest.fixture(scope="module")
def keyframe_calls(oracle):
    """Every solve_pnp call of a closed-loop oracle run (cv2 PnP in the loop): inputs + outputs."""
    calls = []
    inner = oracle.solve_pnp
    cur = {"cam": None}

    def recording(world, pts2d, cam_matrix, rvec, tvec):
        g = np.concatenate([np.asarray(rvec, np.float64).reshape(3), np.asarray(tvec, np.float64).reshape(3)])
        out = inner(world, pts2d, cam_matrix, rvec, tvec)
        calls.append(dict(world=world.copy(), image=pts2d.copy(), guess=g, rvec=np.array(out[0]).reshape(3),
                          tvec=np.array(out[1]).reshape(3), inl=np.array(out[3]), cam=cur["cam"]))
        return out
    oracle.solve_pnp = recording
    try:
        for cam, motion, seed in ((synth.KITTI00, synth.KITTI_MOTION, 0), (synth.D435I, synth.D435I_MOTION, 1)):
            cur["cam"] = cam
            proc = oracle.make_pipeline(cam)
            for t in range(48):
                l, r = synth.stereo_frame(seed, t, cam, motion)
                proc.process(l, r)
    finally:
        oracle.solve_pnp = inner
    return calls


def test_pnp_matches_cv2_on_the_sequences(svo_lib, keyframe_calls):
    from stereo_vo_b200.capi import Engine
    assert len(keyframe_calls) >= 12
    e = Engine(1241, 376, n_seq=4)
    worst_t = worst_r = 0.0
    n_same = 0
    for k, c in enumerate(keyframe_calls):
        f, cx, cy, _ = c["cam"].f32()
        res, inl = e.pnp_ransac([c["world"]], [c["image"]], f, cx, cy, guess=c["guess"][None])
        dt, da, same = _check_pose(res[0], inl[0], c["rvec"], c["tvec"], c["inl"], f"keyframe call {k}",
                                   c["world"], c["image"], (float(f), float(cx), float(cy)))
        n_same += same
        if same:
            worst_t, worst_r = max(worst_t, dt), max(worst_r, da)
    print(f"\n[pnp] {len(keyframe_calls)} keyframe calls: identical inlier sets in {n_same}; on those "
          f"max |dt| {worst_t:.2e} m, max rotation difference {worst_r:.2e} rad vs cv2.solvePnPRansac")
    assert n_same >= 0.6 * len(keyframe_calls)
    e.close()


def _scene(rng, n, outlier_frac, noise_px, planar):
    f, cx, cy = 718.856, 607.19, 185.22
    X = np.stack([rng.uniform(-12, 12, n), rng.uniform(-4, 4, n),
                  np.full(n, 15.0) if planar else rng.uniform(8, 40, n)], 1)
    import cv2
    rvec = rng.normal(0, 0.03, 3)
    tvec = np.array([rng.normal(0, 0.5), rng.normal(0, 0.1), rng.normal(0, 0.8)])
    R, _ = cv2.Rodrigues(rvec.reshape(3, 1))
    xc = X @ R.T + tvec
    uv = np.stack([f * xc[:, 0] / xc[:, 2] + cx, f * xc[:, 1] / xc[:, 2] + cy], 1)
    uv += rng.normal(0, noise_px, uv.shape)
    bad = rng.choice(n, int(outlier_frac * n), replace=False)
    uv[bad] += rng.choice([-1, 1], (len(bad), 2)) * rng.uniform(30, 200, (len(bad), 2))
    return X.astype(np.float32), uv.astype(np.float32), (f, cx, cy)


@pytest.mark.parametrize("planar", [True, False], ids=["planar", "3d"])
@pytest.mark.parametrize("outlier_frac", [0.0, 0.2, 0.45])
def test_pnp_with_outliers_batched(svo_lib, oracle, planar, outlier_frac):
    """Gross outliers (30..200 px): same inliers and pose as cv2, several point sets per launch,
    and the batched launch equals the one-by-one launches bit for bit."""
    from stereo_vo_b200.capi import Engine
    rng = np.random.default_rng(11 + int(100 * outlier_frac) + planar)
    sets = [_scene(rng, n, outlier_frac, 0.25, planar) for n in (60, 150, 300, 400)]
    f, cx, cy = sets[0][2]
    e = Engine(1241, 376, n_seq=4)
    K = oracle.camera_matrix(f, cx, cy)
    res, inl = e.pnp_ransac([s[0] for s in sets], [s[1] for s in sets], f, cx, cy)
    for i, (X, uv, _) in enumerate(sets):
        rv, tv, _rm, ref_inl = oracle.solve_pnp(X, uv, K, np.zeros(3), np.zeros(3))
        _check_pose(res[i], inl[i], rv, tv, ref_inl, f"set {i}", X, uv, (f, cx, cy))
        assert len(inl[i]) >= int((1 - outlier_frac) * len(X)) - 3
    e.close()


def test_pnp_degenerate_inputs(svo_lib):
    """Fewer than 5 points, or no consensus: ok = 0 and the guess comes back, as
    cv::solvePnPRansac returns false and leaves the pose to the caller."""
    from stereo_vo_b200.capi import Engine
    rng = np.random.default_rng(3)
    e = Engine(1241, 376, n_seq=2)
    X, uv, (f, cx, cy) = _scene(rng, 4, 0.0, 0.1, False)
    g = np.array([[0.01, -0.02, 0.005, 0.1, 0.0, -0.2]])
    res, inl = e.pnp_ransac([X], [uv], f, cx, cy, guess=g)
    assert res[0]["ok"] == 0 and res[0]["n_inliers"] == 0 and len(inl[0]) == 0
    assert np.allclose(res[0]["rvec"], g[0, :3]) and np.allclose(res[0]["tvec"], g[0, 3:])
    X = rng.uniform(-5, 5, (50, 3)).astype(np.float32) + np.float32([0, 0, 20])
    uv = rng.uniform(0, 1000, (50, 2)).astype(np.float32)          # unrelated image points
    res, inl = e.pnp_ransac([X], [uv], f, cx, cy)
    assert res[0]["n_inliers"] < 12
    e.close()


# ---------------------------------------------------------------- closed loop, many sequences
CL_S, CL_F, CL_STEPS = 16, 40, 40


def _oracle_closed_loop(arg):
    """The cv2 oracle with ITS OWN pose step (cv2.solvePnPRansac) over one sequence."""
    seed, order = arg
    import cv2
    from oracle import cv2_frontend as O
    cv2.setNumThreads(1)
    O.canonical_mode()
    cam = synth.KITTI00
    frames = np.empty((CL_F, 2, cam.height, cam.width), np.uint8)
    for t in range(CL_F):
        frames[t, 0], frames[t, 1] = synth.stereo_frame(seed, t)
    proc = O.make_pipeline(cam)
    log = []
    for t in order:
        rec = proc.process(frames[t, 0], frames[t, 1])
        log.append({k: rec[k] for k in ("decision", "detected", "hmat", "inliers", "tracked_ids",
                                        "tracker_ids") if k in rec})
    return seed, frames, log


def test_closed_loop_batch_with_gpu_pose_step(svo_lib):
    """svo::BatchFrontEnd with the pose step on the GPU (closed loop, 16 sequences, 4 groups,
    pipelined) beside the cv2 oracle running its own solvePnPRansac. Corners never depend on the
    pose and must be identical; keyframe poses must agree to the borderline tolerance and stay on
    the synthetic ground truth; decisions and track ids agree until the first borderline inlier
    (the two solvers may classify a point with ~8 px error differently), and the fraction of
    frames with identical decisions is reported."""
    import multiprocessing as mp
    import os
    from stereo_vo_b200 import build, capi, hostapi, workload
    if build.host_needs_build():
        build.build_host_lib()
    cam = synth.KITTI00
    sched = workload.Schedule(CL_S, CL_F, phase_step=0)          # every sequence starts at t = 0
    slab = capi.pinned_empty((CL_S, CL_F, 2, cam.height, cam.width), np.uint8)
    jobs = [(s, [sched.frame(k, s) for k in range(CL_STEPS)]) for s in range(CL_S)]
    logs = [None] * CL_S
    with mp.get_context("spawn").Pool(min(CL_S, len(os.sched_getaffinity(0)))) as pool:
        for seed, frames, log in pool.imap_unordered(_oracle_closed_loop, jobs):
            slab[seed] = frames
            logs[seed] = log
    batch = hostapi.Batch(cam.width, cam.height, CL_S, 4, cam, closed_loop=True)
    tables = workload.PointerTables(sched, slab.ctypes.data, cam)
    agree = {"frames": 0, "decision": 0, "ids": 0, "kf_both": 0}
    worst = {"gt": 0.0, "dt": 0.0, "da": 0.0}

    def on_step(k, _sm):
        for s in range(CL_S):
            a, g = logs[s][k], batch.get(s)
            assert np.array_equal(a["detected"], g["corners"]), f"step {k} sequence {s}: corners"
            agree["frames"] += 1
            agree["decision"] += a["decision"] == g["decision"]
            if "tracker_ids" in a and a["decision"] == g["decision"]:
                agree["ids"] += a["tracker_ids"] == [int(i) for i in g["tracker_ids"]]
            if g["decision"] == "keyframe":
                t = sched.frame(k, s)
                gt = np.array(synth.ground_truth(t)) - np.array(synth.ground_truth(sched.frame(0, s)))
                err = float(np.abs(g["hmat"][:3, 3] - gt).max())
                worst["gt"] = max(worst["gt"], err)
                assert err < 0.05, f"step {k} sequence {s}: pose {g['hmat'][:3, 3]} vs ground truth {gt}"
                assert g["n_inliers"] >= 0.85 * g["stats"]["n_kept"], (k, s, g["n_inliers"])
                if a["decision"] == "keyframe":
                    agree["kf_both"] += 1
                    worst["dt"] = max(worst["dt"], float(np.abs(a["hmat"][:3, 3] - g["hmat"][:3, 3]).max()))
                    worst["da"] = max(worst["da"], _angle(a["hmat"][:3, :3].astype(np.float64),
                                                          g["hmat"][:3, :3].astype(np.float64)))

    workload.run_steps(batch, lambda k: tables.args(k, cam.width, False, True), 0, CL_STEPS,
                       pipeline=True, on_step=on_step)
    batch.close()
    print(f"\n[closed loop] {agree['frames']} frames: decisions identical on {agree['decision']}, "
          f"keyframes in both {agree['kf_both']}; pose vs cv2 closed loop: max |dt| {worst['dt']:.2e} m, "
          f"rotation {worst['da']:.2e} rad; vs ground truth {worst['gt']:.3f} m")
    assert agree["kf_both"] >= 6 * CL_S
    assert agree["decision"] >= 0.95 * agree["frames"]
    assert worst["dt"] <= 2e-2 and worst["da"] <= 2e-3      # two closed loops, each on its own tracks
