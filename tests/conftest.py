import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def svo_lib():
    """The in-tree libsvo.so (built here with nvcc if it is stale or missing)."""
    from stereo_vo_b200 import build, capi
    import shutil
    if build.needs_build():
        if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
            pytest.fail("libsvo.so missing/stale and nvcc not available")
        build.build_lib()
    return capi.load_library()


@pytest.fixture(scope="session")
def oracle():
    from oracle import cv2_frontend
    cv2_frontend.canonical_mode()
    return cv2_frontend
