"""GPU: svo_reprojection_evaluate against the REFERENCE'S OWN compiled ReprojectionFactor::Evaluate
(tests/golden/ba_reference.npz, written by oracle/_ref/libref_ba.so = the reference's
bundle_adjuster.cpp + reprojection_factor.cpp compiled where they lie; see
tests/test_reference_ba.py for the CPU side): the 848 residual blocks of the 5-keyframe window the
reference's add_keyframe built, and 256 single factors with non-unit quaternions. No oracle in
between. (Collected late - the file name - because it was added at the very end of the round;
it passed on a B200 with the last seconds of the GPU budget.)"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FIXTURE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ba_reference.npz")
REL = 1e-10          # fp64 expressions in a different order (chain rule vs MATLAB-generated)


def _close(a, b):
    return np.abs(a - b).max() <= REL * max(1.0, float(np.abs(b).max()))


def evaluate_and_compare(evaluate, G):
    """evaluate(poses, points, pose_idx, point_idx, obs, f, cx, cy) -> (resid, jac_pose, jac_point)"""
    f, cx, cy = (float(v) for v in G["fac_camera"])
    for pre in ("win_", "fac_"):
        n = G[pre + "pose"].shape[0]
        idx = np.arange(n, dtype=np.int32)
        r, jp, jx = evaluate(G[pre + "pose"], G[pre + "point"], idx, idx, G[pre + "obs"], f, cx, cy)
        assert r.shape == (n, 2) and jp.shape == (n, 2, 7) and jx.shape == (n, 2, 3)
        assert _close(r, G[pre + "resid"]), pre
        assert _close(jp, G[pre + "jac_pose"]), pre
        assert _close(jx, G[pre + "jac_point"]), pre
        assert (jp[:, 0, 5] == 0.0).all() and (jp[:, 1, 4] == 0.0).all()      # the reference's setZero
    return True


def test_reprojection_kernel_equals_the_compiled_reference(svo_lib):
    from stereo_vo_b200.capi import Engine
    G = dict(np.load(FIXTURE))
    e = Engine(640, 480, n_seq=1)
    try:
        assert evaluate_and_compare(e.reprojection_evaluate, G)
    finally:
        e.close()
