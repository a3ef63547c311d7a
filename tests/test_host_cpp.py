"""CPU-side checks of the C++ host layer (libsvo_host.so): it builds with g++, loads, exports
every symbol include/svo_host_c.h declares, reads the reference's camera YAML the way vo_node
does (src/vo_node.cpp:82-98) and - with no GPU - fails loudly instead of falling back."""
import ctypes as C
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# contents of the reference's config/kitti00.yaml and config/d435i.yaml (7 flat keys each)
KITTI00_YAML = """focal_length: 7.18856e+02
cx: 6.071928e+02
cy: 1.852157e+02
baseline: 0.537165718864418
left_topic: "/leftImage"
right_topic: "/rightImage"
frame_rate: 11
"""


@pytest.fixture(scope="module")
def host(svo_lib):
    from stereo_vo_b200 import build, hostapi
    if build.host_needs_build():
        build.build_host_lib()
    return hostapi.load_library()


def test_host_library_exports_its_header(host):
    from stereo_vo_b200 import hostapi
    declared = sorted(set(hostapi.declared_symbols()))
    assert len(declared) >= 18, declared
    for name in declared:
        assert hasattr(host, name), f"libsvo_host.so does not export {name}"


def test_facade_headers_keep_the_reference_signatures():
    """The public member functions of the reference classes (src/image_processor.hpp:31-46,
    src/feature_tracker.hpp:20-54) are declared with the same names and argument lists."""
    inc = os.path.join(ROOT, "include", "stereo_vo")
    ft = " ".join(open(os.path.join(inc, "feature_tracker.hpp")).read().split())
    ip = " ".join(open(os.path.join(inc, "image_processor.hpp")).read().split())
    for sig in ["void init(const cv::Mat& image, const std::vector<cv::Point2f>& features, const std::vector<size_t>& ids);",
                "void track_features(float& av_parallax, float& percent_lost, const cv::Mat& image, bool flow_back);",
                "void get_tracked_features(std::vector<cv::Point2f>& features, std::vector<size_t>& ids);",
                "void draw_track();", "cv::Mat get_drawing();"]:
        assert sig in ft, sig
    for sig in ["ImageProcessor(cv::Mat cam_mat, std::shared_ptr<FeatureTracker> tracker, std::shared_ptr<BundleAdjuster> adjuster, float bline, float min_feature_distance, float parallax_thresh);",
                "void process(const StereoPair& stereo_pair);",
                "StereoPair(const cv::Mat& l, const cv::Mat& r, double stamp)"]:
        assert sig in ip, sig


def test_camera_yaml_semantics(host):
    from stereo_vo_b200 import hostapi, synth
    f, cx, cy, b, rate = hostapi.camera_from_yaml(KITTI00_YAML)
    # narrowed to float exactly like vo_node's float variables
    assert f == np.float32(718.856) and cx == np.float32(607.1928)
    assert cy == np.float32(185.2157) and b == np.float32(0.537165718864418)
    assert rate == np.float32(11)
    assert (f, cx, cy, b) == tuple(synth.KITTI00.f32())
    # comments, quoting and unknown keys are tolerated; a missing intrinsic is an error
    f2, *_ = hostapi.camera_from_yaml("# cam\nfocal_length: '385.75'\ncx: 1\ncy: 2\nbaseline: 0.05\nfoo: bar\n")
    assert f2 == np.float32(385.75)
    from stereo_vo_b200.capi import SvoError
    with pytest.raises(SvoError):
        hostapi.camera_from_yaml("focal_length: 1\ncx: 2\ncy: 3\n")


def test_host_layer_has_no_cpu_fallback(host):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from stereo_vo_b200 import capi, hostapi, synth
    p = capi.default_params(64, 48)
    h = C.c_void_p()
    rc = host.svo_batch_create(C.byref(p), 0, 2, 2, 100.0, 32.0, 24.0, 0.1, 20.0, C.byref(h))
    assert rc != 0 and not h.value
    assert b"svo_create" in host.svo_batch_last_error(None)
    vo = hostapi.VoSequence(*synth.KITTI00.f32())
    img = np.zeros((48, 64), np.uint8)
    with pytest.raises(capi.SvoError):
        vo.process(img, img)
    vo.close()


def test_bad_arguments_are_rejected(host):
    assert host.svo_vo_process(None, None, None, 0, 0, 0, 0.0) != 0
    assert host.svo_batch_step(None, None, None, 0, 0, None, None, None, None) != 0
    assert host.svo_camera_from_yaml(None, None, None, None, None, None) != 0
