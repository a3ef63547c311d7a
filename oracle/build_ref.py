"""TEST INFRASTRUCTURE. Builds oracle/_ref/libref_ba.so: the reference's OWN bundle-adjustment
sources, compiled where they lie -

    /root/reference/src/bundle_adjuster.cpp      (add_keyframe, remove_oldest_pose, ids, refcounts,
                                                  sliding window, world-point store, bundle_adjust)
    /root/reference/src/reprojection_factor.cpp  (residual + both Jacobians)

- together with oracle/ref_harness.cpp (C entry points) against the stand-in headers under
oracle/ref_stubs/ for the three SDKs this image lacks: a small fixed-size Eigen, an OpenCV with
three value types, and a Ceres that RECORDS the problem instead of solving it. No reference
source is copied; outputs go to oracle/_ref/ only (git-ignored). The pixel path of the reference
(image_processor.cpp, feature_tracker.cpp) cannot be built this way: its arithmetic is inside
OpenCV itself, so compiling it against stand-ins would pin nothing.

    python -m oracle.build_ref          # needs /root/reference; used by tests/golden/make_golden_ba.py
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src"
OUT_DIR = os.path.join(HERE, "_ref")
LIB = os.path.join(OUT_DIR, "libref_ba.so")
SOURCES = [os.path.join(REF_SRC, "bundle_adjuster.cpp"), os.path.join(REF_SRC, "reprojection_factor.cpp"),
           os.path.join(HERE, "ref_harness.cpp")]


def available() -> bool:
    return all(os.path.exists(s) for s in SOURCES)


def build(force: bool = False) -> str:
    """Compile the library when the reference tree is present; returns its path."""
    if not available():
        raise RuntimeError("the reference tree is not on this machine (/root/reference/src)")
    deps = SOURCES + [os.path.join(dp, f) for dp, _, fs in os.walk(os.path.join(HERE, "ref_stubs")) for f in fs]
    if not force and os.path.exists(LIB) and all(os.path.getmtime(LIB) >= os.path.getmtime(d) for d in deps):
        return LIB
    os.makedirs(OUT_DIR, exist_ok=True)
    tmp = LIB + f".tmp{os.getpid()}"
    # -O1 -ffp-contract=off: one rounding per operation in the MATLAB-generated expressions
    cmd = ["g++", "-std=c++14", "-O1", "-ffp-contract=off", "-fPIC", "-shared", "-w",
           "-I", os.path.join(HERE, "ref_stubs"), "-I", REF_SRC, *SOURCES, "-o", tmp]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("building oracle/_ref failed:\n" + r.stdout.decode(errors="replace"))
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
