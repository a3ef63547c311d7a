"""TEST INFRASTRUCTURE. Builds oracle/_ref/libref_ba.so: the reference's OWN bundle-adjustment
sources, compiled where they lie (and oracle/_ref/libref_frontend.so, below) -

    /root/reference/src/bundle_adjuster.cpp      (add_keyframe, remove_oldest_pose, ids, refcounts,
                                                  sliding window, world-point store, bundle_adjust)
    /root/reference/src/reprojection_factor.cpp  (residual + both Jacobians)

- together with oracle/ref_harness.cpp (C entry points) against the stand-in headers under
oracle/ref_stubs/ for the three SDKs this image lacks: a small fixed-size Eigen, an OpenCV with
three value types, and a Ceres that RECORDS the problem instead of solving it. No reference
source is copied; outputs go to oracle/_ref/ only (git-ignored).

The pixel path of the reference (image_processor.cpp, feature_tracker.cpp) keeps its arithmetic
inside OpenCV, which is not here as an SDK - but it IS here as Python cv2. libref_frontend.so is
those two files (plus the two above) compiled against a stand-in <opencv2/opencv.hpp> whose
value types (Mat, Point2f, ...) are implemented and whose ALGORITHMS (goodFeaturesToTrack,
calcOpticalFlowPyrLK, StereoBM, solvePnPRansac, Rodrigues, the matrix product) are forwarded
through C callbacks to the real OpenCV in the test process (oracle/ref_frontend.py). That pins
the control flow of the Python restatement (decisions, id plumbing, dedupe, filter, statistics,
reprojection loop, the literal arguments of every OpenCV call) against the reference's compiled
code; what stays unpinned is OpenCV 4.13 versus the 3.x the reference was written for.

    python -m oracle.build_ref          # needs /root/reference; used by tests/golden/make_golden_ba.py
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
# (both overridable so that the tests can play a machine without the reference and without the
# built libraries)
REF_SRC = os.environ.get("ORACLE_REF_SRC", "/root/reference/src")
OUT_DIR = os.environ.get("ORACLE_REF_OUT", os.path.join(HERE, "_ref"))
LIB = os.path.join(OUT_DIR, "libref_ba.so")
SOURCES = [os.path.join(REF_SRC, "bundle_adjuster.cpp"), os.path.join(REF_SRC, "reprojection_factor.cpp"),
           os.path.join(HERE, "ref_harness.cpp")]
# the whole front end: process() / track_features() / triangulate_stereo() as compiled reference
# code, the OpenCV algorithms they call served by callbacks (oracle/ref_frontend.py)
FRONTEND_LIB = os.path.join(OUT_DIR, "libref_frontend.so")
FRONTEND_SOURCES = [os.path.join(REF_SRC, f) for f in ("image_processor.cpp", "feature_tracker.cpp",
                                                       "bundle_adjuster.cpp", "reprojection_factor.cpp")] + \
                   [os.path.join(HERE, "ref_frontend_harness.cpp")]


def available() -> bool:
    return all(os.path.exists(s) for s in SOURCES + FRONTEND_SOURCES)


def _compile(sources, lib, force):
    deps = sources + [os.path.join(dp, f) for dp, _, fs in os.walk(os.path.join(HERE, "ref_stubs")) for f in fs]
    if not force and os.path.exists(lib) and all(os.path.getmtime(lib) >= os.path.getmtime(d) for d in deps):
        return lib
    os.makedirs(OUT_DIR, exist_ok=True)
    tmp = lib + f".tmp{os.getpid()}"
    # -O1 -ffp-contract=off: one rounding per operation (the reference's f32 loops, the
    # MATLAB-generated expressions)
    cmd = ["g++", "-std=c++14", "-O1", "-ffp-contract=off", "-fPIC", "-shared", "-w",
           "-I", os.path.join(HERE, "ref_stubs"), "-I", REF_SRC, *sources, "-o", tmp]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("building oracle/_ref failed:\n" + r.stdout.decode(errors="replace"))
    os.replace(tmp, lib)
    return lib


def build(force: bool = False) -> str:
    """Compile both libraries when the reference tree is present; returns the path of libref_ba.so."""
    if not available():
        raise RuntimeError("the reference tree is not on this machine (/root/reference/src)")
    _compile(FRONTEND_SOURCES, FRONTEND_LIB, force)
    return _compile(SOURCES, LIB, force)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    print(FRONTEND_LIB)
