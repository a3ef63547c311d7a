"""ORACLE (test infrastructure, not product code): ReprojectionFactor::Evaluate of the reference
(src/reprojection_factor.cpp:10-88, a ceres::SizedCostFunction<2, 7, 3>) restated in numpy.

Parameters as the reference lays them out: pose = (qw, qx, qy, qz, tx, ty, tz), the orientation
and translation of the WORLD frame with respect to the camera (:16-17), point = feature position
in the world frame (:19). The residual follows :24-37 literally:

    ss = [v]x,  ll = ss + qw I,  LR = v v^T + ll ll      (v = (qx, qy, qz); LR = |q|^2 R(q))
    gamma = LR p / |q|^2 + t,   psi = 1 / gamma_z,   r = K psi gamma - obs,  K = [f 0 cx; 0 f cy]

The Jacobians of the reference are MATLAB-generated expressions (:62-85) of exactly this
function, INCLUDING the derivative of the 1/|q|^2 normalisation, 2 x 7 row-major with
d r0/d ty = d r1/d tx = 0 (:58 setZero) and 2 x 3 for the point. Here they are derived by hand
(chain rule through gamma); tests/golden/reprojection_factor.npz holds the values of the
reference's own expressions (evaluated from the reference source by
tests/golden/make_golden_reprojection.py) and tests/test_reprojection_oracle.py checks this file
against them and against central differences.

Only tests/ may import this module.
"""
from __future__ import annotations

import numpy as np


def _skew(v):
    return np.array([[0.0, -v[2], v[1]], [v[2], 0.0, -v[0]], [-v[1], v[0], 0.0]])


def evaluate(pose, point, obs, focal, cx, cy):
    """-> (residual (2,), jac_pose (2, 7), jac_point (2, 3)), float64."""
    pose = np.asarray(pose, np.float64)
    p = np.asarray(point, np.float64)
    w, v, t = pose[0], pose[1:4], pose[4:7]
    ss = _skew(v)                                            # :26-29
    ll = ss + w * np.eye(3)                                  # :31
    lr = np.outer(v, v) + ll @ ll                            # :33
    n = float(pose[:4] @ pose[:4])
    gamma = lr @ p / n + t                                   # :35
    psi = 1.0 / gamma[2]                                     # :37
    k = np.array([[focal, 0.0, cx], [0.0, focal, cy]])
    resid = k @ (psi * gamma) - np.asarray(obs, np.float64)  # :40
    # d r / d gamma
    drg = np.array([[focal * psi, 0.0, -focal * gamma[0] * psi * psi],
                    [0.0, focal * psi, -focal * gamma[1] * psi * psi]])
    # d (LR p) / d q : LR p = (w^2 - |v|^2) p + 2 v (v.p) + 2 w (v x p)
    mp = lr @ p
    dmp = np.zeros((3, 4))
    dmp[:, 0] = 2.0 * w * p + 2.0 * np.cross(v, p)
    for i in range(3):
        e = np.zeros(3)
        e[i] = 1.0
        dmp[:, 1 + i] = -2.0 * v[i] * p + 2.0 * e * float(v @ p) + 2.0 * v * p[i] + 2.0 * w * np.cross(e, p)
    dgq = dmp / n - np.outer(mp, 2.0 * pose[:4]) / (n * n)  # normalisation differentiated too
    jac_pose = np.zeros((2, 7))
    jac_pose[:, :4] = drg @ dgq
    jac_pose[:, 4:] = drg                                    # d gamma / d t = I
    jac_point = drg @ (lr / n)
    return resid, jac_pose, jac_point


def evaluate_batch(poses, points, pose_idx, point_idx, obs, focal, cx, cy):
    """All observations of a window: residuals (n, 2), jac_pose (n, 2, 7), jac_point (n, 2, 3)."""
    n = len(pose_idx)
    r = np.zeros((n, 2))
    jp = np.zeros((n, 2, 7))
    jx = np.zeros((n, 2, 3))
    for i in range(n):
        r[i], jp[i], jx[i] = evaluate(poses[pose_idx[i]], points[point_idx[i]], obs[i], focal, cx, cy)
    return r, jp, jx
