// kitti_source.hpp - the frame source of the reference's kitti_node (src/kitti_node.cpp:16-68)
// without ROS: KITTI odometry stereo pairs `<data>/<seq>/image_0/%06d.png`, `image_1/%06d.png`
// (8-bit grey PNG, what cv::imread(..., CV_LOAD_IMAGE_GRAYSCALE) at :67-68 returns) and the
// ground-truth poses `<data>/data_odometry_poses/dataset/poses/<seq>.txt`, one 3x4 row-major
// [R | t] per line (:37,47-50). SURVEY.md 8f rank 4: the I/O side of the front end.
//
// Decoding stays on the CPU (zlib inflate + PNG un-filtering) but off the critical path: a
// Prefetcher thread decodes frames ahead into caller-provided (pinned) stereo slots, so the
// engine's H2D copy of frame t overlaps the decode of frame t + depth.
#pragma once
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace svo {

struct KittiPose {
    double R[9];   // row major
    double t[3];
};

// 8-bit PNG -> 8-bit grey, rows `pitch` bytes apart. Colour types 0 (grey), 2 (RGB), 4 (grey +
// alpha), 6 (RGBA), non-interlaced; RGB is converted the way cv::imread's grey mode does it
// (libpng's png_set_rgb_to_gray(0.299, 0.587): 15-bit weights 9797 / 19234 / 3737, truncated). Throws std::runtime_error
// with the file name on anything else - never returns a partly decoded image.
void read_png_size(const std::string& path, int& width, int& height);
void read_png_gray(const std::string& path, uint8_t* out, int pitch, int width, int height);

class KittiSource {
  public:
    // data_path as in src/kitti_node.cpp:16 (ends with '/'), sequence "00".."21"
    KittiSource(const std::string& data_path, const std::string& sequence);
    size_t size() const { return poses_.size(); }        // frames with a ground-truth pose (:47)
    int width() const { return width_; }
    int height() const { return height_; }
    const KittiPose& pose(size_t i) const { return poses_.at(i); }
    std::string left_path(size_t i) const { return image_path(0, i); }
    std::string right_path(size_t i) const { return image_path(1, i); }
    // decode frame i into caller memory (left and right, `pitch` bytes per row)
    void read(size_t i, uint8_t* left, uint8_t* right, int pitch) const;

  private:
    std::string image_path(int cam, size_t i) const;
    std::string root_, seq_;
    std::vector<KittiPose> poses_;
    int width_ = 0, height_ = 0;
};

// Decodes frames first, first + 1, ... on `threads` worker threads into a ring of `depth` >= 3
// stereo slots carved from `slab` (depth * 2 * pitch * height bytes; pinned memory for
// asynchronous H2D), up to depth - 2 frames ahead. Frames are handed out in order whatever
// order the workers finish in: next() blocks until the next frame is decoded and returns its
// slot; a slot stays valid until two further next() calls (two steps may be in flight).
// One thread decodes ~90 KITTI-sized stereo pairs per second (11 ms per pair, zlib inflate is
// 3/4 of it; measured 91 / 184 / 365 / 687 pairs/s with 1 / 2 / 4 / 8 threads on 8 cores).
class KittiPrefetcher {
  public:
    KittiPrefetcher(const KittiSource& src, uint8_t* slab, int depth, int pitch, size_t first = 0,
                    int threads = 1);
    ~KittiPrefetcher();
    // false at the end of the sequence
    bool next(size_t& index, const uint8_t*& left, const uint8_t*& right);

  private:
    void run();
    const KittiSource& src_;
    uint8_t* slab_;
    int depth_, pitch_;
    size_t claimed_, consumed_, end_;      // next frame a worker takes / next frame handed out
    std::vector<size_t> ready_;            // per slot: index of the decoded frame it holds + 1
    std::string error_;                    // what() of the first (lowest-index) frame that failed
    size_t error_index_ = 0;
    bool stop_ = false;
    std::mutex mu_;
    std::condition_variable cv_;
    std::vector<std::thread> workers_;
};

}  // namespace svo
