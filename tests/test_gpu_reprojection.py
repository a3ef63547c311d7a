"""GPU: svo_reprojection_evaluate (batched ReprojectionFactor::Evaluate)
against the numpy oracle and the reference's own Jacobian expressions (golden vectors)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reprojection_factor.npz")
REL = 1e-12


def _close(a, b):
    return np.abs(a - b).max() <= REL * max(1.0, float(np.abs(b).max()))


def test_golden_reference_expressions(svo_lib):
    from stereo_vo_b200.capi import Engine
    z = np.load(GOLD)
    f, cx, cy = z["intr"]
    n = len(z["poses"])
    e = Engine(640, 480, n_seq=1)
    idx = np.arange(n, dtype=np.int32)
    r, jp, jx = e.reprojection_evaluate(z["poses"], z["points"], idx, idx, z["obs"], f, cx, cy)
    for i in range(n):
        assert _close(jp[i].reshape(14), z["jac_pose"][i]), i
        assert _close(jx[i].reshape(6), z["jac_point"][i]), i
        assert jp[i, 0, 5] == 0.0 and jp[i, 1, 4] == 0.0
    e.close()


def test_sliding_window_matches_oracle(svo_lib):
    """A window as BundleAdjuster builds it (src/vo_node.cpp:36, src/bundle_adjuster.hpp:75):
    5 keyframes x <= 400 observations each over a shared set of points."""
    from oracle import reprojection_factor as RF
    from stereo_vo_b200.capi import Engine
    rng = np.random.default_rng(9)
    n_kf, n_pts = 5, 900
    q = rng.normal(size=(n_kf, 4)) * 0.05 + np.array([1.0, 0, 0, 0])
    poses = np.concatenate([q, rng.normal(scale=0.5, size=(n_kf, 3))], 1)
    points = np.stack([rng.uniform(-12, 12, n_pts), rng.uniform(-4, 4, n_pts), rng.uniform(8, 40, n_pts)], 1)
    pi, xi = [], []
    for k in range(n_kf):
        seen = rng.choice(n_pts, 400, replace=False)
        pi += [k] * 400
        xi += list(seen)
    obs = rng.uniform(0, 1241, (len(pi), 2))
    f, cx, cy = 718.856, 607.1928, 185.2157
    e = Engine(640, 480, n_seq=1)
    r, jp, jx = e.reprojection_evaluate(poses, points, pi, xi, obs, f, cx, cy)
    rr, rjp, rjx = RF.evaluate_batch(poses, points, pi, xi, obs, f, cx, cy)
    assert _close(r, rr) and _close(jp, rjp) and _close(jx, rjx)
    # outputs are optional and indices are validated
    from stereo_vo_b200.capi import SvoError
    with pytest.raises(SvoError):
        e.reprojection_evaluate(poses, points, [n_kf], [0], obs[:1], f, cx, cy)
    e.close()
