"""CPU: race / memory checks of the host-side I/O code (race detection). The KITTI
source, its multi-threaded decode-ahead ring and the image-message queue are compiled from their
sources together with tests/host_sanitizer_driver.cpp under -fsanitize=thread and
-fsanitize=address,undefined and run on a small KITTI-shaped tree: any report fails the test.
(The CUDA side cannot be checked this way: compute-sanitizer is closed on this pool.)"""
import os
import shutil
import subprocess

import cv2
import numpy as np
import pytest

from stereo_vo_b200 import synth
from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SOURCES = [os.path.join(ROOT, "tests", "host_sanitizer_driver.cpp"),
           os.path.join(ROOT, "stereo_vo_b200", "host", "kitti_source.cpp"),
           os.path.join(ROOT, "stereo_vo_b200", "host", "image_msg.cpp")]
REPORTS = ("WARNING: ThreadSanitizer", "ERROR: AddressSanitizer", "ERROR: LeakSanitizer",
           "runtime error:")


@pytest.fixture(scope="module")
def tree(tmp_path_factory):
    root = tmp_path_factory.mktemp("kitti_san")
    cam = util.small_camera(333, 97)
    os.makedirs(root / "00" / "image_0")
    os.makedirs(root / "00" / "image_1")
    os.makedirs(root / "data_odometry_poses" / "dataset" / "poses")
    n = 14
    for i in range(n):
        l, r = synth.stereo_frame(3, i, cam)
        cv2.imwrite(str(root / "00" / "image_0" / f"{i:06d}.png"), l)
        cv2.imwrite(str(root / "00" / "image_1" / f"{i:06d}.png"), np.dstack([r, r, r]) if i % 5 == 0 else r)
    with open(root / "data_odometry_poses" / "dataset" / "poses" / "00.txt", "w") as fh:
        for i in range(n):
            fh.write("1 0 0 %d 0 1 0 0 0 0 1 0\n" % i)
    return str(root) + "/"


@pytest.mark.parametrize("sanitizer", ["thread", "address,undefined"])
def test_host_io_code_is_clean_under(sanitizer, tree, tmp_path):
    gxx = shutil.which("g++")
    if gxx is None:
        pytest.skip("no g++")
    exe = str(tmp_path / "driver")
    build = subprocess.run([gxx, "-std=c++17", "-O1", "-g", f"-fsanitize={sanitizer}",
                            "-fno-omit-frame-pointer", "-I", os.path.join(ROOT, "include"), *SOURCES,
                            "-lz", "-pthread", "-o", exe], capture_output=True, text=True)
    if build.returncode != 0 and "sanitizer" in build.stderr.lower() and "cannot find" in build.stderr.lower():
        pytest.skip("sanitizer runtime not installed")
    assert build.returncode == 0, build.stderr[-3000:]
    env = dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=66",
               ASAN_OPTIONS="detect_leaks=1 exitcode=66", UBSAN_OPTIONS="print_stacktrace=1")
    run = subprocess.run([exe, tree], capture_output=True, text=True, env=env, timeout=600)
    out = run.stdout + run.stderr
    if "FATAL: ThreadSanitizer" in out or "Shadow memory range interleaves" in out:
        pytest.skip("the sanitizer cannot map its shadow memory in this container")
    assert run.returncode == 0, out[-4000:]
    assert not any(r in out for r in REPORTS), out[-4000:]
    assert "threads 8 depth 11" in out and "queued+drained" in out
