"""Generates the committed golden vectors under tests/golden/ from the cv2 oracle
(oracle/cv2_frontend.py, canonical mode cv2.setUseOptimized(False)) on deterministic synthetic
sequences. Run from the repo root:  python tests/golden/make_golden.py

The reference has no tests or fixtures of its own; these files are what
pins the oracle across machines: tests/test_golden.py re-runs the oracle against them on the
CPU, tests/test_gpu_golden.py checks the CUDA path against them without running the oracle.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import cv2_frontend as O          # noqa: E402
from stereo_vo_b200 import synth              # noqa: E402

CASES = {
    "kitti_seq0": (synth.KITTI00, synth.KITTI_MOTION, 0, 10),
    "d435i_seq1": (synth.D435I, synth.D435I_MOTION, 1, 8),
}


def flatten(log, frames):
    out = {"n_frames": np.int32(len(log))}
    for t, (rec, (l, r)) in enumerate(zip(log, frames)):
        p = f"f{t}_"
        out[p + "sha"] = np.frombuffer(bytes.fromhex(synth.digest(l, r)), np.uint8)
        out[p + "decision"] = np.array(rec["decision"])
        out[p + "detected"] = rec["detected"]
        if "av_parallax" in rec:
            out[p + "av_parallax"] = np.float32(rec["av_parallax"])
            out[p + "percent_lost"] = np.float32(rec["percent_lost"])
            out[p + "tracked_pts"] = rec["tracked_pts"]
            out[p + "tracked_ids"] = np.array(rec["tracked_ids"], np.int64)
            out[p + "status1"] = rec["taps"]["status1"]
            out[p + "status2"] = rec["taps"]["status2"]
            out[p + "kept_idx"] = rec["taps"]["kept_idx"]
        if rec["decision"] in ("first", "keyframe"):
            for k in ("new_features", "valid_idx", "disp16", "new2d", "new3d", "hmat",
                      "tracker_pts"):
                out[p + k] = np.asarray(rec[k])
            out[p + "new_ids"] = np.array(rec["new_ids"], np.int64)
            out[p + "tracker_ids"] = np.array(rec["tracker_ids"], np.int64)
        if rec["decision"] == "keyframe":
            out[p + "inliers"] = np.asarray(rec["inliers"], np.int32)
            out[p + "rvec"] = np.asarray(rec["rvec"])
            out[p + "tvec"] = np.asarray(rec["tvec"])
    return out


def main():
    O.canonical_mode()
    here = os.path.dirname(os.path.abspath(__file__))
    for name, (cam, motion, seed, n) in CASES.items():
        frames = synth.sequence(seed, n, cam, motion)
        proc = O.make_pipeline(cam)
        for l, r in frames:
            proc.process(l, r)
        data = flatten(proc.log, frames)
        # one standalone StereoBM probe set per case: disp16 at 2000 seeded pixels of frame 0
        rng = np.random.default_rng(42)
        pts = np.stack([rng.integers(0, cam.width, 2000), rng.integers(0, cam.height, 2000)], 1)
        d16 = O.stereo_disparity16(*frames[0])
        data["probe_pts"] = pts.astype(np.int32)
        data["probe_disp16"] = d16[pts[:, 1], pts[:, 0]]
        import cv2
        eig = cv2.cornerMinEigenVal(frames[0][0], 3, ksize=3)
        data["eig_max_bits"] = np.float32(eig.max()).view(np.uint32)
        data["eig_row100"] = eig[100].copy()
        path = os.path.join(here, name + ".npz")
        np.savez_compressed(path, **data)
        print(name, os.path.getsize(path), "bytes", [str(data[f"f{t}_decision"]) for t in range(n)])


if __name__ == "__main__":
    main()
