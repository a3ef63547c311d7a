// image_msg.hpp - the ROS side of the reference's frame ingest without ROS
// (src/vo_node.cpp:38-76 and :139-143; SURVEY.md 8f rank 4 "ROS sensor_msgs/Image mono8"):
//
//   * svo::ImageMsg            the fields of sensor_msgs/Image that cv_bridge reads
//   * svo::to_mono8            what cv_bridge::toCvCopy(msg, "mono8") returns (vo_node.cpp:70-71)
//   * handle_images            the reference's synchronised-pair callback (:38-76): pairs closer
//                              than drop_time to the last queued one are dropped, the rest are
//                              converted to mono8 and queued as StereoPair
//   * svo::process_queue       the node loop's drain (:140-143)
//
// cv_bridge is a third-party dependency of the reference (package.xml, not vendored, unpinned;
// ROS 1 vision_opencv). Its published conversion rule is restated here: same encoding = deep
// copy; colour -> grey through cv::cvtColor (OpenCV's 8-bit fixed point: (R*9798 + G*19235 +
// B*3735 + 2^14) >> 15); 16 -> 8 bit through convertTo(..., 255 / 65535), i.e. round(v / 257);
// 16-bit pixels are byte-swapped when the message's endianness differs from the host's;
// encodings that are not colour formats (8UC1, 32FC1, ...) cannot become mono8 and throw.
// tests/test_image_msg.py pins every conversion against cv2 4.13 bit for bit. (OpenCV 3.x, the
// reference's era, weighted colours with 14-bit coefficients 4899 / 9617 / 1868: a colour source
// can come out one grey level apart there. Both of the reference's cameras publish mono8, which
// is a plain copy in every version.)
//
// The ApproximateTime synchroniser and the subscribers themselves are ROS machinery (out of
// scope): the caller hands over pairs that already belong together.
#pragma once
#include <cstddef>
#include <cstdint>
#include <memory>
#include <queue>
#include <string>

#include "cv_compat.hpp"
#include "image_processor.hpp"

namespace svo {

struct ImageMsg {
    double stamp = 0.0;            // header.stamp.toSec()
    uint32_t height = 0, width = 0;
    std::string encoding;          // sensor_msgs::image_encodings name
    uint8_t is_bigendian = 0;
    uint32_t step = 0;             // bytes per row
    const uint8_t* data = nullptr; // step * height bytes, not owned
    size_t size = 0;
};

// cv_bridge::toCvCopy(msg, sensor_msgs::image_encodings::MONO8): an owned CV_8UC1 image.
// Accepted source encodings: mono8, mono16, rgb8, bgr8, rgba8, bgra8, rgb16, bgr16, rgba16,
// bgra16. Throws std::invalid_argument on a malformed message (step too small, data too short,
// empty image) and on an encoding cv_bridge cannot convert to mono8.
cv::Mat to_mono8(const ImageMsg& msg);
// the same into caller memory (height rows, `pitch` >= width bytes apart)
void to_mono8(const ImageMsg& msg, uint8_t* out, size_t pitch);

// the node loop of src/vo_node.cpp:140-143: every queued pair goes through `process`, oldest
// first, and is popped after it returned (a pair whose processing throws stays at the front).
// Returns the number of pairs processed.
template <class F>
size_t drain_queue(std::queue<StereoPair>& image_queue, F&& process) {
    size_t n = 0;
    while (!image_queue.empty()) {
        process(image_queue.front());
        image_queue.pop();
        ++n;
    }
    return n;
}
inline size_t process_queue(std::queue<StereoPair>& image_queue, ImageProcessor& image_processor) {
    return drain_queue(image_queue, [&](const StereoPair& p) { image_processor.process(p); });
}

}  // namespace svo

// src/vo_node.cpp:38-76, spelled like the reference
class handle_images {
  public:
    handle_images(std::shared_ptr<std::queue<StereoPair>> image_queue, double drop_time)
        : image_queue(std::move(image_queue)), drop_time(drop_time) {}

    // returns true when the pair was queued (the reference's operator returns void)
    bool operator()(const svo::ImageMsg& left_msg, const svo::ImageMsg& right_msg);

  private:
    std::shared_ptr<std::queue<StereoPair>> image_queue;
    double drop_time;
};
