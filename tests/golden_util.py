"""Loader for the committed golden vectors (tests/golden/*.npz, made by make_golden.py)."""
import os

import numpy as np

from stereo_vo_b200 import synth

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {
    "kitti_seq0": (synth.KITTI00, synth.KITTI_MOTION, 0),
    "d435i_seq1": (synth.D435I, synth.D435I_MOTION, 1),
}


def load(name):
    z = np.load(os.path.join(HERE, name + ".npz"))
    n = int(z["n_frames"])
    recs = []
    for t in range(n):
        p = f"f{t}_"
        rec = {k[len(p):]: z[k] for k in z.files if k.startswith(p)}
        rec["decision"] = str(rec["decision"])
        for k in ("tracked_ids", "new_ids", "tracker_ids"):
            if k in rec:
                rec[k] = [int(v) for v in rec[k]]
        recs.append(rec)
    extra = {k: z[k] for k in ("probe_pts", "probe_disp16", "eig_max_bits", "eig_row100")}
    return recs, extra


def frames_for(name, n):
    cam, motion, seed = CASES[name]
    fr = synth.sequence(seed, n, cam, motion)
    return fr, cam


def check_inputs(recs, frames):
    """The deterministic generator must reproduce the exact bytes the vectors were made from."""
    for t, (rec, (l, r)) in enumerate(zip(recs, frames)):
        assert bytes(rec["sha"]).hex() == synth.digest(l, r), \
            f"frame {t}: synthetic generator drifted from the golden inputs"
