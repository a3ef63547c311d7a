#!/usr/bin/env python
"""Per-frame latency of ONE stereo sequence through the C++ ImageProcessor facade
(svo_vo_process: host frames in, results out) - BASELINE configs[1] (KITTI 1241x376) and
configs[2] (d435i 640x480, the small-image latency path) - next to the cv2 oracle on one core.
PnP is the oracle's cv2.solvePnPRansac through the facade's solver hook (closed loop)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import cv2
    from oracle import cv2_frontend as O
    from stereo_vo_b200 import hostapi, synth
    out = {}
    for name, cam, motion in (("kitti00_1241x376", synth.KITTI00, synth.KITTI_MOTION),
                              ("d435i_640x480", synth.D435I, synth.D435I_MOTION)):
        frames = synth.sequence(3, 40, cam, motion)
        f, cx, cy, b = cam.f32()
        pnp_ms = [0.0]

        def pnp(world, image, K, rvec, tvec):
            t0 = time.perf_counter()
            r = O.solve_pnp(world, image, K, rvec, tvec)
            pnp_ms[0] += (time.perf_counter() - t0) * 1e3
            return r
        vo = hostapi.VoSequence(f, cx, cy, b, pnp=pnp)
        lat, kinds = [], []
        for t, (l, r) in enumerate(frames):
            pnp_ms[0] = 0.0
            t0 = time.perf_counter()
            rec = vo.process(l, r, t * 0.1)
            dt = (time.perf_counter() - t0) * 1e3
            if t >= 5:
                lat.append(dt - pnp_ms[0])           # front end only; host PnP timed apart
                kinds.append(rec["decision"])
        vo.close()
        lat = np.array(lat)
        kf = np.array([k == "keyframe" for k in kinds])
        cv2.setNumThreads(1)
        cv2.setUseOptimized(True)
        ref = O.make_pipeline(cam)
        rl = []
        for t, (l, r) in enumerate(frames):
            t0 = time.perf_counter()
            ref.process(l, r)
            if t >= 5:
                rl.append((time.perf_counter() - t0) * 1e3)
        out[name] = {
            "gpu_ms_per_frame_median": float(np.median(lat)), "gpu_ms_p95": float(np.percentile(lat, 95)),
            "gpu_ms_skip_frames": float(np.median(lat[~kf])) if (~kf).any() else None,
            "gpu_ms_keyframes": float(np.median(lat[kf])) if kf.any() else None,
            "keyframes": int(kf.sum()), "frames": int(len(lat)),
            "cpu_cv2_1thread_ms_per_frame_median": float(np.median(rl)),
            "note": "gpu_*: wall clock around svo_vo_process (H2D + kernels + D2H + C++ host "
                    "bookkeeping + ctypes record copy) minus the host PnP callback; cpu_*: the cv2 "
                    "oracle's process() on one core, its own PnP included (~0.4 ms on keyframes)",
        }
    print(json.dumps(out))


if __name__ == "__main__":
    main()
