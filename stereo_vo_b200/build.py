"""Build libsvo.so (hand-written sm_100a CUDA + the C ABI of include/svo_c.h) in-tree with nvcc.

The library is compiled for sm_100a only (no PTX fallback for other architectures, no CPU
path). `--fmad=false` keeps every float operation un-fused: parity with the reference's OpenCV
arithmetic is defined on separately rounded IEEE operations.
"""
from __future__ import annotations

import contextlib
import fcntl
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsvo.so")
SOURCES = ["engine.cu", "pyr_down.cu", "corner_score.cu", "sort_candidates.cu",
           "greedy_select.cu", "stereo.cu", "stereo_dense.cu", "lk_track.cu", "track_filter.cu",
           "pnp_ransac.cu", "reprojection_factor.cu", "draw_track.cu", "dedupe_static.cu"]
NVCC_FLAGS = ["-O3", "-std=c++17", "--fmad=false", "-gencode", "arch=compute_100a,code=sm_100a",
              "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
              "--threads", "0"]


HOST_LIB = os.path.join(HERE, "libsvo_host.so")
HOST_DIR = os.path.join(HERE, "host")
INCLUDE = os.path.join(HERE, "..", "include")
HOST_SOURCES = ["bundle_adjuster.cpp", "camera_config.cpp", "feature_tracker.cpp",
                "image_processor.cpp", "batch_front_end.cpp", "kitti_source.cpp", "image_msg.cpp",
                "host_c.cpp"]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall", "-Wextra", "-pthread"]


@contextlib.contextmanager
def _build_lock():
    """Serialise builds across processes (torchrun ranks, pytest-xdist workers): every rank may
    call build_lib(); the first one in compiles, the others wait and find the library fresh.
    Libraries are linked under a temporary name and renamed into place, so a process that is
    dlopen()ing the old file never sees a half-written one."""
    fd = os.open(os.path.join(HERE, ".build.lock"), os.O_CREAT | os.O_RDWR, 0o644)
    try:
        fcntl.flock(fd, fcntl.LOCK_EX)
        yield
    finally:
        fcntl.flock(fd, fcntl.LOCK_UN)
        os.close(fd)


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, "..", "include", "svo_c.h"))
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def host_needs_build() -> bool:
    if not os.path.exists(HOST_LIB):
        return True
    t = os.path.getmtime(HOST_LIB)
    deps = [os.path.join(HOST_DIR, f) for f in os.listdir(HOST_DIR)]
    for root, _d, files in os.walk(INCLUDE):
        deps += [os.path.join(root, f) for f in files]
    deps.append(LIB)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build_host_lib(force: bool = False) -> str:
    """g++ -> stereo_vo_b200/libsvo_host.so: the C++ host layer (ImageProcessor / FeatureTracker
    facade, BundleAdjuster bookkeeping, BatchFrontEnd, YAML reader) linked against libsvo.so."""
    if not force and not host_needs_build():
        return HOST_LIB
    with _build_lock():
        if not force and not host_needs_build():       # another process built it meanwhile
            return HOST_LIB
        # the system compiler, whose libstdc++ is the shared one every other library of the
        # process uses (a toolchain under /opt that links libstdc++ statically puts a second copy
        # of the iostream / locale state into this library: crashes in stream extraction once
        # another libstdc++ user is loaded). SVO_CXX overrides.
        cxx = os.environ.get("SVO_CXX") or ("/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++")
        tmp = f"{HOST_LIB}.tmp.{os.getpid()}"
        cmd = [cxx, *CXX_FLAGS, "-shared", "-I", INCLUDE,
               *[os.path.join(HOST_DIR, f) for f in HOST_SOURCES], "-o", tmp,
               "-L", HERE, "-l:libsvo.so", "-lz", "-Wl,-rpath,$ORIGIN"]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
        text = r.stdout.decode(errors="replace")
        if r.returncode != 0:
            sys.stderr.write(f"[build] libsvo_host.so FAILED\n{text}\n")
            raise RuntimeError("g++ failed")
        if text.strip():
            sys.stderr.write(f"[build] libsvo_host.so\n{text}\n")
        os.replace(tmp, HOST_LIB)
    return HOST_LIB


def build_all(force: bool = False) -> None:
    build_lib(force=force)
    build_host_lib(force=force)


def build_lib(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    with _build_lock():
        if not force and not needs_build():            # another process built it meanwhile
            return LIB
        return _build_lib_locked(verbose)


def _build_lib_locked(verbose: bool) -> str:
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [_nvcc(), *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        text = out.decode(errors="replace")
        if p.returncode != 0:
            failed = True
            sys.stderr.write(f"[build] {src} FAILED\n{text}\n")
        elif verbose or text.strip():
            sys.stderr.write(f"[build] {src}\n{text}\n")
    if failed:
        raise RuntimeError("nvcc failed")
    tmp = f"{LIB}.tmp.{os.getpid()}"
    link = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs]
    subprocess.run(link, check=True)
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
    print(build_host_lib(force="--force" in sys.argv))
