"""Golden log of the REFERENCE'S OWN front end: oracle/_ref/libref_frontend.so is
/root/reference/src/{image_processor,feature_tracker,bundle_adjuster,reprojection_factor}.cpp
compiled where they lie (oracle/build_ref.py), with the OpenCV algorithms it calls served by the
real OpenCV through callbacks (oracle/ref_frontend.py). This script runs it on synthetic
sequences and stores, per frame, what the reference did - so the tests can hold the Python
restatement to it on machines without the reference tree. Run in the build container:

    python tests/golden/make_golden_ref_frontend.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

from oracle import cv2_frontend as O, ref_frontend  # noqa: E402
from stereo_vo_b200 import synth  # noqa: E402

CASES = {"kitti_seed1": (synth.KITTI00, synth.KITTI_MOTION, 1, 60),
         "d435i_seed3": (synth.D435I, synth.D435I_MOTION, 3, 40)}


def run_reference(cam, motion, seed, n_frames, guess="carried"):
    """-> list of per-frame dicts (plain arrays only)"""
    O.canonical_mode()
    f, cx, cy, b = cam.f32()
    ref = ref_frontend.RefPipeline(f, cx, cy, b, guess=guess)
    out = []
    for t, (l, r) in enumerate(synth.sequence(seed, n_frames, cam, motion)):
        a = ref.process(l, r, t / 11.0)
        calls = dict()
        for name, c in a["calls"]:
            calls.setdefault(name, []).append(c)
        rec = dict(made_keyframe=np.int32(a["made_keyframe"]),
                   n_calls=np.array([len(calls.get(k, [])) for k in
                                     ("goodFeaturesToTrack", "calcOpticalFlowPyrLK", "solvePnPRansac", "StereoBM")], np.int32),
                   detected=calls["goodFeaturesToTrack"][0]["corners"],
                   tracker_pts=a["tracker_pts"], tracker_ids=a["tracker_ids"])
        if "calcOpticalFlowPyrLK" in calls:
            fwd = calls["calcOpticalFlowPyrLK"][0]
            rec.update(lk_in=fwd["pts"], lk_out=fwd["out"], lk_status=fwd["status"])
        if "solvePnPRansac" in calls:
            p = calls["solvePnPRansac"][0]
            rec.update(pnp_inliers=p["inliers"], pnp_rvec=p["rvec"], pnp_tvec=p["tvec"],
                       pnp_world=p["world"], pnp_image=p["image"])
        if a["made_keyframe"]:
            rec.update(kf_tracked_ids=a["kf_tracked_ids"], kf_tracked_2d=a["kf_tracked_2d"], new2d=a["new2d"],
                       new3d=a["new3d"], new_ids=a["new_ids"], pose7=a["pose7"])
        out.append(rec)
    params = dict()
    for frame in ref.log:
        for name, c in frame["calls"]:
            params.setdefault(name, {k: v for k, v in c.items() if np.isscalar(v) or isinstance(v, tuple)})
    ref.close()
    return out, params


def main():
    store = {}
    for case, (cam, motion, seed, n) in CASES.items():
        log, params = run_reference(cam, motion, seed, n)
        store[f"{case}/n_frames"] = np.int32(n)
        # the literal arguments the reference's code handed to OpenCV
        g, k, p, sb = (params[x] for x in ("goodFeaturesToTrack", "calcOpticalFlowPyrLK", "solvePnPRansac", "StereoBM"))
        store[f"{case}/args"] = np.array([g["max_corners"], g["quality"], g["min_distance"], k["win"][0], k["win"][1],
                                          k["max_level"], k["criteria"][0], k["criteria"][1], k["criteria"][2],
                                          k["flags"], k["min_eig"], sb["ndisp"], sb["block"], p["use_guess"],
                                          p["iterations"], p["reproj"], p["confidence"], p["depth"]], np.float64)
        for t, rec in enumerate(log):
            for k, v in rec.items():
                store[f"{case}/{t}/{k}"] = np.asarray(v)
        print(case, "keyframes", sum(int(r["made_keyframe"]) for r in log), params)
    path = os.path.join(HERE, "ref_frontend.npz")
    np.savez_compressed(path, **store)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
