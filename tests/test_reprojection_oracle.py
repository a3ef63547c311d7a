"""CPU: the numpy restatement of ReprojectionFactor::Evaluate (oracle/reprojection_factor.py)
against the reference's own MATLAB-generated Jacobian expressions (golden vectors evaluated from
/root/reference/src/reprojection_factor.cpp:62-85 by tests/golden/make_golden_reprojection.py)
and against central differences of the residual."""
import os

import numpy as np

from oracle import reprojection_factor as RF

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reprojection_factor.npz")


def test_jacobians_equal_the_reference_expressions():
    z = np.load(GOLD)
    f, cx, cy = z["intr"]
    for i in range(len(z["poses"])):
        _r, jp, jx = RF.evaluate(z["poses"][i], z["points"][i], z["obs"][i], f, cx, cy)
        ref_p, ref_x = z["jac_pose"][i], z["jac_point"][i]
        assert ref_p[5] == 0.0 and ref_p[11] == 0.0            # :58 setZero, never assigned
        assert np.abs(jp.reshape(14) - ref_p).max() <= 1e-12 * max(1.0, np.abs(ref_p).max())
        assert np.abs(jx.reshape(6) - ref_x).max() <= 1e-12 * max(1.0, np.abs(ref_x).max())


def test_jacobians_are_the_derivatives_of_the_residual():
    z = np.load(GOLD)
    f, cx, cy = z["intr"]
    h = 1e-6
    for i in range(0, len(z["poses"]), 4):
        pose, pt, ob = z["poses"][i], z["points"][i], z["obs"][i]
        _r, jp, jx = RF.evaluate(pose, pt, ob, f, cx, cy)
        for k in range(7):
            d = np.zeros(7)
            d[k] = h
            num = (RF.evaluate(pose + d, pt, ob, f, cx, cy)[0] - RF.evaluate(pose - d, pt, ob, f, cx, cy)[0]) / (2 * h)
            assert np.abs(num - jp[:, k]).max() <= 1e-5 * max(1.0, np.abs(jp).max())
        for k in range(3):
            d = np.zeros(3)
            d[k] = h
            num = (RF.evaluate(pose, pt + d, ob, f, cx, cy)[0] - RF.evaluate(pose, pt - d, ob, f, cx, cy)[0]) / (2 * h)
            assert np.abs(num - jx[:, k]).max() <= 1e-5 * max(1.0, np.abs(jx).max())


def test_residual_is_the_pinhole_reprojection():
    """unit quaternion: r = K (R p + t) / z - obs"""
    rng = np.random.default_rng(5)
    q = rng.normal(size=4)
    q /= np.linalg.norm(q)
    w, x, y, zq = q
    R = np.array([[1 - 2 * (y * y + zq * zq), 2 * (x * y - w * zq), 2 * (x * zq + w * y)],
                  [2 * (x * y + w * zq), 1 - 2 * (x * x + zq * zq), 2 * (y * zq - w * x)],
                  [2 * (x * zq - w * y), 2 * (y * zq + w * x), 1 - 2 * (x * x + y * y)]])
    t = np.array([0.3, -0.1, 0.5])
    p = np.array([1.0, -2.0, 12.0])
    g = R @ p + t
    r, _, _ = RF.evaluate(np.concatenate([q, t]), p, [10.0, 20.0], 700.0, 600.0, 180.0)
    assert np.allclose(r, [700 * g[0] / g[2] + 600 - 10, 700 * g[1] / g[2] + 180 - 20], rtol=0, atol=1e-10)
