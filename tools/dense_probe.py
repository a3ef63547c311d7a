"""Times stereo_bm_dense (svo_stereo_bm_dense_resident) on S KITTI-shaped stereo pairs that
are already on the device and beside it cv2.StereoBM on the host cores (one image, one thread).
Prints one JSON line; `python tools/dense_probe.py [S] [reps]`."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from stereo_vo_b200 import synth                      # noqa: E402
from stereo_vo_b200.capi import Engine                # noqa: E402


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    cam = synth.KITTI00
    W, H = cam.width, cam.height
    pairs = [synth.sequence(s % 8, 1)[0] for s in range(min(S, 8))]
    e = Engine(W, H, n_seq=S)
    for s in range(S):
        l, r = pairs[s % len(pairs)]
        m = e.stereo_bm_dense(l, r, seq=s)
    seqs = list(range(S))
    for _ in range(3):
        e.stereo_bm_dense_resident(seqs, fetch=False)
    ms = []
    for _ in range(reps):
        e.stereo_bm_dense_resident(seqs, fetch=False)
        ms.append(e.last_timing()[1])               # events around the two kernels
    e.close()
    ms = float(np.median(ms))
    n = W * H
    out = {"kernel": "stereo_bm_dense", "pairs_per_launch": S, "ms_per_launch": ms,
           "us_per_pair": 1e3 * ms / S, "pairs_per_s": S / (ms * 1e-3),
           "algorithmic_bytes_per_pair": 4 * n, "algo_gbs": 4 * n * S / (ms * 1e-3) / 1e9,
           "valid_fraction": float((m > 0).mean())}
    try:
        import cv2
        cv2.setNumThreads(1)
        l, r = pairs[0]
        sbm = cv2.StereoBM_create(48, 21)
        sbm.compute(l, r)
        t0 = time.perf_counter()
        for _ in range(5):
            sbm.compute(l, r)
        out["cv2_ms_per_pair_1_thread"] = (time.perf_counter() - t0) / 5 * 1e3
    except ImportError:
        pass
    print(json.dumps(out))


if __name__ == "__main__":
    main()
