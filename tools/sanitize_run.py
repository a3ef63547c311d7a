#!/usr/bin/env python
"""Small end-to-end workload for compute-sanitizer (one tool per run):

    compute-sanitizer --tool memcheck  python tools/sanitize_run.py [kitti|uhd|dense|extras]
    compute-sanitizer --tool racecheck python tools/sanitize_run.py ...

kitti : 3 KITTI-shaped sequences x 6 frames through svo::BatchFrontEnd (steps + keyframes,
        teacher-forced and closed loop with the GPU pose step)
uhd   : one 3840x2160 step + keyframe stage with the 20 k-corner detector
dense : one dense StereoBM map
extras: reprojection factors, draw_track overlay
Exit code 0 and "SANITIZE_RUN_OK" on stdout when every call returned; the sanitizer's own
summary line says whether it saw errors / hazards."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def kitti():
    from stereo_vo_b200 import capi, hostapi, synth, workload
    cam = synth.KITTI00
    S, F = 3, 6
    slab = capi.pinned_empty((S, F, 2, cam.height, cam.width), np.uint8)
    workload.generate_frames(list(range(S)), F, slab, workers=1)
    for closed in (False, True):
        batch = hostapi.Batch(cam.width, cam.height, S, 2, cam, closed_loop=closed)
        tables = workload.PointerTables(workload.Schedule(S, F, phase_step=1), slab.ctypes.data, cam)
        tot = workload.run_steps(batch, lambda k: tables.args(k, cam.width, False, True), 0, F, True)
        assert tot["keyframes"] >= S, tot
        batch.close()


def uhd():
    from stereo_vo_b200 import synth
    from stereo_vo_b200.capi import Engine
    cam = synth.UHD4K
    f0 = synth.stereo_frame(0, 0, cam, synth.UHD4K_MOTION)
    f1 = synth.stereo_frame(0, 1, cam, synth.UHD4K_MOTION)
    e = Engine(cam.width, cam.height, n_seq=1, max_corners=20000, quality_level=0.01,
               min_distance=20.0, max_features=20000)
    pts = e.good_features_to_track(f0[0])
    e.tracker_init(f0[0], pts[:3000], np.arange(3000, dtype=np.uint64))
    e.tracker_track(f1[0], True)
    f, cx, cy, b = cam.f32()
    e.triangulate_stereo(f1[0], f1[1], pts[:3000], np.eye(4, dtype=np.float32), f, cx, cy, b)
    e.close()


def dense():
    from stereo_vo_b200 import synth
    from stereo_vo_b200.capi import Engine
    cam = synth.KITTI00
    l, r = synth.stereo_frame(1, 0)
    e = Engine(cam.width, cam.height, n_seq=1)
    m = e.stereo_bm_dense(l, r)
    assert (m > 0).mean() > 0.3
    e.close()


def extras():
    from stereo_vo_b200 import synth
    from stereo_vo_b200.capi import Engine
    rng = np.random.default_rng(1)
    e = Engine(640, 480, n_seq=1)
    poses = np.concatenate([rng.normal(size=(5, 4)) * 0.05 + [1, 0, 0, 0], rng.normal(size=(5, 3))], 1)
    pts = rng.normal(size=(300, 3)) + [0, 0, 20]
    e.reprojection_evaluate(poses, pts, rng.integers(0, 5, 900), rng.integers(0, 300, 900),
                            rng.uniform(0, 600, (900, 2)), 500.0, 320.0, 240.0)
    img = synth.stereo_frame(0, 0, synth.D435I, synth.D435I_MOTION)[0]
    a = rng.uniform(0, 480, (200, 2)).astype(np.float32)
    e.draw_track(img, a, a + rng.normal(scale=15, size=a.shape).astype(np.float32))
    e.close()


if __name__ == "__main__":
    which = sys.argv[1:] or ["kitti", "uhd", "dense", "extras"]
    for name in which:
        {"kitti": kitti, "uhd": uhd, "dense": dense, "extras": extras}[name]()
        print(f"[sanitize_run] {name} done", flush=True)
    print("SANITIZE_RUN_OK")
