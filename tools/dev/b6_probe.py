"""Where the reference's own code reaches its feature-limit branch on the synthetic
KITTI-shaped sequence of seed 0: runs oracle/_ref/libref_frontend.so (the reference's compiled
front end, OpenCV served by cv2) for 110 frames and prints every keyframe's counts. Build
container only (needs /root/reference).   python tools/dev/b6_probe.py [seed] [frames]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from oracle import cv2_frontend as O, ref_frontend  # noqa: E402
from stereo_vo_b200 import synth  # noqa: E402

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
n = int(sys.argv[2]) if len(sys.argv) > 2 else 110
O.canonical_mode()
f, cx, cy, bl = synth.KITTI00.f32()
ref = ref_frontend.RefPipeline(f, cx, cy, bl)
for t, (l, r) in enumerate(synth.sequence(seed, n)):
    a = ref.process(l, r, t / 11.0)
    if a["made_keyframe"]:
        flag = "   <-- new_ids != new features: B6" if len(a["new_ids"]) != len(a["new2d"]) else ""
        print(f"frame {t:3d} keyframe: tracked {len(a['kf_tracked_ids']):3d}  new features {len(a['new2d']):3d}  "
              f"new_ids {len(a['new_ids']):3d}  tracker set {len(a['tracker_ids']):3d}{flag}", flush=True)
