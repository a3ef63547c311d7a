"""Build libsvo.so (hand-written sm_100a CUDA + the C ABI of include/svo_c.h) in-tree with nvcc.

The library is compiled for sm_100a only (no PTX fallback for other architectures, no CPU
path). `--fmad=false` keeps every float operation un-fused: parity with the reference's OpenCV
arithmetic is defined on separately rounded IEEE operations.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsvo.so")
SOURCES = ["engine.cu", "pyr_down.cu", "corner_score.cu", "sort_candidates.cu",
           "greedy_select.cu", "stereo.cu", "lk_track.cu", "track_filter.cu"]
NVCC_FLAGS = ["-O3", "-std=c++17", "--fmad=false", "-gencode", "arch=compute_100a,code=sm_100a",
              "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
              "--threads", "0"]


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


def build_lib(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
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
    link = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs]
    subprocess.run(link, check=True)
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
