// camera_info.hpp - interface type of the reference kept as is (src/camera_info.hpp:4-18):
// pinhole intrinsics, the four (unused, always 0: src/vo_node.cpp:110) distortion terms and the
// stereo baseline, all double. Plus the YAML reader that reproduces how vo_node fills it.
#pragma once
#include <string>

struct CameraInfo {
    double focal = 0, cx = 0, cy = 0;        // intrinsics
    double k1 = 0, k2 = 0, p1 = 0, p2 = 0;   // distortion (never read by the front end)
    double baseline = 0;
};

namespace svo {

// What `config/*.yaml` carries (7 keys each). vo_node reads the four numbers into FLOAT
// variables (src/vo_node.cpp:82-98) and only then widens them into CameraInfo (:110); the
// topics and frame_rate are parsed and unused by the front end (frame_rate is unused by the
// reference as well, src/vo_node.cpp:98).
struct CameraConfig {
    float focal_length = 0.f, cx = 0.f, cy = 0.f, baseline = 0.f;
    std::string left_topic, right_topic;
    float frame_rate = 0.f;
    CameraInfo info() const {
        CameraInfo ci;
        ci.focal = focal_length; ci.cx = cx; ci.cy = cy; ci.baseline = baseline;
        return ci;
    }
};

// Parse the flat `key: value` YAML the reference ships. Throws std::runtime_error when one of
// focal_length / cx / cy / baseline is missing (the reference's ros::param::get leaves the
// variable unset in that case; a headless harness cannot run on that).
CameraConfig parse_camera_yaml(const std::string& text);
CameraConfig load_camera_yaml(const std::string& path);

}  // namespace svo
