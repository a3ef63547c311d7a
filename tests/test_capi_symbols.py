"""The C-ABI library builds for sm_100a, loads, and exports every symbol include/svo_c.h
declares. No compute calls here (no GPU needed)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "svo_c.h")).read()
    return sorted(set(re.findall(r"SVO_API[^;(]*?\b(svo_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(svo_lib):
    from stereo_vo_b200 import capi
    declared = _declared_symbols()
    assert len(declared) >= 30
    assert sorted(capi.EXPORTS) == declared
    for name in declared:
        assert hasattr(svo_lib, name), f"libsvo.so does not export {name}"


def test_library_is_sm100a_only(svo_lib):
    from stereo_vo_b200 import capi
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    out = subprocess.run([cuobjdump, "-lelf", capi.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_default_params_are_the_reference_constants(svo_lib):
    from stereo_vo_b200 import capi
    p = capi.default_params(1241, 376)
    assert (p.max_corners, p.quality_level, p.min_distance) == (300, 0.1, 30.0)   # image_processor.cpp:22, vo_node.cpp:34
    assert (p.num_disparities, p.sad_window) == (48, 21)                          # image_processor.cpp:174
    assert (p.prefilter_cap, p.texture_threshold, p.uniqueness_ratio) == (31, 10, 15)
    assert (p.lk_max_level, p.lk_max_iters, p.lk_epsilon, p.lk_min_eig_threshold) == (3, 30, 0.01, 1e-2)
    assert (p.fb_max_error, p.max_parallax, p.max_features) == (2.0, 200.0, 400)


def test_no_cpu_fallback_without_gpu(svo_lib):
    """Without a CUDA device engine creation must fail loudly (never compute on the CPU)."""
    import torch
    from stereo_vo_b200 import capi
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.SvoError) as ei:
        capi.Engine(1241, 376)
    assert "SVO_ERR_CUDA" in str(ei.value)


def test_bad_arguments_are_rejected(svo_lib):
    from stereo_vo_b200 import capi
    h = ctypes.c_void_p()
    p = capi.default_params(10, 10)
    assert svo_lib.svo_create(ctypes.byref(p), 0, 1, ctypes.byref(h)) == -1   # SVO_ERR_INVALID
    assert b"image size" in svo_lib.svo_last_error(None)
    p = capi.default_params(640, 480)
    assert svo_lib.svo_create(ctypes.byref(p), 0, 0, ctypes.byref(h)) == -1
    assert svo_lib.svo_num_levels(None) == -1
