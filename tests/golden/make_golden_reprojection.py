"""Golden vectors for ReprojectionFactor::Evaluate: evaluates the REFERENCE'S OWN expressions
(read from /root/reference/src/reprojection_factor.cpp where they lie; nothing is copied into
this repository) on seeded random inputs and stores inputs + outputs in
tests/golden/reprojection_factor.npz. Run in the build container (the reference tree does not
exist on the GPU box):

    python tests/golden/make_golden_reprojection.py
"""
import os
import re

import numpy as np

REF = "/root/reference/src/reprojection_factor.cpp"
HERE = os.path.dirname(os.path.abspath(__file__))


def reference_expressions():
    """{('pose'|'point', flat index): python expression} from the `jacobians[b][i] = ...;` lines
    (:62-85) - MATLAB-generated C, plain arithmetic + pow()."""
    src = open(REF).read()
    out = {}
    for blk, idx, expr in re.findall(r"jacobians\[(\d)\]\[(\d+)\]\s*=\s*([^;]+);", src):
        py = re.sub(r"pow\(([^,()]*(?:\([^()]*\)[^,()]*)*(?:\((?:[^()]|\([^()]*\))*\))*[^,()]*),\s*2\.0\)", r"((\1)**2.0)", expr)
        while "pow(" in py:                       # nested parentheses: resolve innermost-first
            py2 = _replace_pow(py)
            if py2 == py:
                raise RuntimeError("could not translate pow()")
            py = py2
        py = py.replace("camera_info.focal", "focal").replace("camera_info.cx", "cx").replace("camera_info.cy", "cy")
        out[("pose" if blk == "0" else "point", int(idx))] = py
    return out


def _replace_pow(s):
    i = s.find("pow(")
    depth, j = 0, i + 3
    args_start = j + 1
    comma = None
    while j < len(s):
        c = s[j]
        if c == "(":
            depth += 1
        elif c == ")":
            depth -= 1
            if depth == 0:
                break
        elif c == "," and depth == 1:
            comma = j
        j += 1
    base, expo = s[args_start:comma], s[comma + 1:j]
    return s[:i] + f"(({base})**({expo}))" + s[j + 1:]


def main():
    exprs = reference_expressions()
    assert len(exprs) == 12 + 6, sorted(exprs)
    rng = np.random.default_rng(20261105)
    n = 64
    poses = np.zeros((n, 7))
    q = rng.normal(size=(n, 4))
    q[: n // 2] /= np.linalg.norm(q[: n // 2], axis=1, keepdims=True)     # half unit, half not
    poses[:, :4] = q
    poses[:, 4:] = rng.normal(scale=2.0, size=(n, 3))
    points = rng.normal(scale=5.0, size=(n, 3)) + np.array([0, 0, 20.0])
    obs = rng.uniform(0, 1200, size=(n, 2))
    focal, cx, cy = 718.856, 607.1928, 185.2157
    jp = np.zeros((n, 14))
    jx = np.zeros((n, 6))
    for i in range(n):
        env = dict(focal=focal, cx=cx, cy=cy)
        env.update({f"q{k + 1}": poses[i, k] for k in range(4)})
        env.update({f"t{k + 1}": poses[i, 4 + k] for k in range(3)})
        env.update({f"p{k + 1}": points[i, k] for k in range(3)})
        for (blk, idx), e in exprs.items():
            val = eval(e, {"__builtins__": {}}, env)
            (jp if blk == "pose" else jx)[i, idx] = val
    np.savez(os.path.join(HERE, "reprojection_factor.npz"), poses=poses, points=points, obs=obs,
             intr=np.array([focal, cx, cy]), jac_pose=jp, jac_point=jx)
    print("wrote", n, "cases;", len(exprs), "reference expressions")


if __name__ == "__main__":
    main()
