// sensor_msgs/Image -> mono8 and the stereo-pair queue of the reference's node
// (see include/stereo_vo/image_msg.hpp). Host code only: these are the copies the reference
// makes on the CPU before process() sees a frame (src/vo_node.cpp:70-73).
#include "stereo_vo/image_msg.hpp"

#include <cstring>
#include <stdexcept>

namespace svo {

namespace {

struct Layout {
    int channels = 0;     // 1, 3 or 4
    int depth = 0;        // bits per channel: 8 or 16
    int red = 0;          // index of the red channel (0: rgb order, 2: bgr order)
};

// the sensor_msgs::image_encodings names cv_bridge treats as colour formats of 8 / 16 bits
Layout layout_of(const std::string& enc) {
    static const struct { const char* name; Layout l; } table[] = {
        {"mono8", {1, 8, 0}},   {"mono16", {1, 16, 0}}, {"rgb8", {3, 8, 0}},   {"bgr8", {3, 8, 2}},
        {"rgba8", {4, 8, 0}},   {"bgra8", {4, 8, 2}},   {"rgb16", {3, 16, 0}}, {"bgr16", {3, 16, 2}},
        {"rgba16", {4, 16, 0}}, {"bgra16", {4, 16, 2}},
    };
    for (const auto& e : table)
        if (enc == e.name) return e.l;
    // cv_bridge: "[8UC1] is not a color format. but [mono8] is. The conversion does not make
    // sense" for the generic type encodings; Bayer / YUV422 sources are not supported here
    throw std::invalid_argument("to_mono8: cannot convert encoding [" + enc + "] to [mono8]");
}

inline bool host_is_bigendian() {
    const uint16_t probe = 0x0102;
    return *reinterpret_cast<const uint8_t*>(&probe) == 0x01;
}

// cv::cvtColor(RGB2GRAY / BGR2GRAY), 8- and 16-bit: 15-bit fixed point, rounded
inline uint32_t gray15(uint32_t r, uint32_t g, uint32_t b) {
    return (r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15;
}
// convertTo(CV_8U, 255. / 65535.): round(v / 257), never a tie
inline uint8_t narrow16(uint32_t v) { return (uint8_t)((v + 128u) / 257u); }

}  // namespace

void to_mono8(const ImageMsg& msg, uint8_t* out, size_t pitch) {
    const Layout l = layout_of(msg.encoding);
    if (msg.width == 0 || msg.height == 0) throw std::invalid_argument("to_mono8: empty image");
    if (!msg.data || !out) throw std::invalid_argument("to_mono8: null buffer");
    const size_t px = (size_t)l.channels * (size_t)(l.depth / 8);
    if ((size_t)msg.step < (size_t)msg.width * px)
        throw std::invalid_argument("to_mono8: step is smaller than a row of [" + msg.encoding + "]");
    if (msg.size < (size_t)msg.step * msg.height)
        throw std::invalid_argument("to_mono8: data is shorter than step * height");
    if (pitch < msg.width) throw std::invalid_argument("to_mono8: pitch < width");
    const bool swap = l.depth == 16 && (msg.is_bigendian != 0) != host_is_bigendian();
    auto rd16 = [swap](const uint8_t* p) -> uint32_t {
        uint16_t v;
        std::memcpy(&v, p, 2);
        return swap ? (uint32_t)(uint16_t)((v << 8) | (v >> 8)) : (uint32_t)v;
    };
    for (uint32_t y = 0; y < msg.height; ++y) {
        const uint8_t* s = msg.data + (size_t)y * msg.step;
        uint8_t* o = out + (size_t)y * pitch;
        if (l.channels == 1 && l.depth == 8) {
            std::memcpy(o, s, msg.width);
        } else if (l.channels == 1) {
            for (uint32_t x = 0; x < msg.width; ++x) o[x] = narrow16(rd16(s + 2 * (size_t)x));
        } else if (l.depth == 8) {
            for (uint32_t x = 0; x < msg.width; ++x) {
                const uint8_t* p = s + (size_t)x * px;
                o[x] = (uint8_t)gray15(p[l.red], p[1], p[2 - l.red]);
            }
        } else {
            // cvtColor on the 16-bit image first, then the depth change (cv_bridge's order)
            for (uint32_t x = 0; x < msg.width; ++x) {
                const uint8_t* p = s + (size_t)x * px;
                o[x] = narrow16(gray15(rd16(p + 2 * l.red), rd16(p + 2), rd16(p + 2 * (2 - l.red))));
            }
        }
    }
}

cv::Mat to_mono8(const ImageMsg& msg) {
    layout_of(msg.encoding);                       // refuse the encoding before allocating
    if (msg.width == 0 || msg.height == 0) throw std::invalid_argument("to_mono8: empty image");
    if (msg.width > 65536u || msg.height > 65536u) throw std::invalid_argument("to_mono8: implausible image size");
    cv::Mat m((int)msg.height, (int)msg.width, CV_8UC1);
    to_mono8(msg, m.data, m.step);
    return m;
}

}  // namespace svo

bool handle_images::operator()(const svo::ImageMsg& left_msg, const svo::ImageMsg& right_msg) {
    const double time = left_msg.stamp;            // left_msg->header.stamp.toSec(), :64
    if (!image_queue->empty() && time - image_queue->back().t < drop_time) return false;   // :65-67
    cv::Mat left = svo::to_mono8(left_msg);        // cv_bridge::toCvCopy(..., MONO8), :70-71
    cv::Mat right = svo::to_mono8(right_msg);
    image_queue->push(StereoPair(left, right, time));                                       // :73
    return true;
}
