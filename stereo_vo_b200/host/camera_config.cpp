// Reader for the reference's camera YAML files (config/kitti00.yaml etc.: seven flat
// `key: value` lines). Numbers are narrowed to float exactly as vo_node does when it copies the
// ROS parameters into float variables (src/vo_node.cpp:82-98).
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <stdexcept>

#include "stereo_vo/camera_info.hpp"

namespace svo {

static std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n\"'");
    if (a == std::string::npos) return "";
    size_t b = s.find_last_not_of(" \t\r\n\"'");
    return s.substr(a, b - a + 1);
}

CameraConfig parse_camera_yaml(const std::string& text) {
    CameraConfig c;
    unsigned seen = 0;
    std::istringstream in(text);
    std::string line;
    while (std::getline(in, line)) {
        size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        size_t colon = line.find(':');
        if (colon == std::string::npos) continue;
        const std::string key = trim(line.substr(0, colon));
        const std::string val = trim(line.substr(colon + 1));
        if (key.empty() || val.empty()) continue;
        auto num = [&](float& dst, unsigned bit) {
            char* end = nullptr;
            double d = std::strtod(val.c_str(), &end);
            if (end == val.c_str()) throw std::runtime_error("camera yaml: '" + key + "' is not a number");
            dst = (float)d;
            seen |= bit;
        };
        if (key == "focal_length") num(c.focal_length, 1u);
        else if (key == "cx") num(c.cx, 2u);
        else if (key == "cy") num(c.cy, 4u);
        else if (key == "baseline") num(c.baseline, 8u);
        else if (key == "frame_rate") num(c.frame_rate, 16u);
        else if (key == "left_topic") c.left_topic = val;
        else if (key == "right_topic") c.right_topic = val;
        // unknown keys are ignored, like parameters nobody asks the ROS master for
    }
    if ((seen & 15u) != 15u)
        throw std::runtime_error("camera yaml: focal_length, cx, cy and baseline are required");
    return c;
}

CameraConfig load_camera_yaml(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("camera yaml: cannot open " + path);
    std::stringstream ss;
    ss << f.rdbuf();
    return parse_camera_yaml(ss.str());
}

}  // namespace svo
