// feature_tracker.hpp - the reference's FeatureTracker interface (src/feature_tracker.hpp:11-63)
// on top of libsvo.so. Public member functions, argument meaning and side effects are the
// reference's; the pixel work (two pyramidal LK passes, forward/backward filter, statistics:
// src/feature_tracker.cpp:18-67) runs in the CUDA library, the images live in device memory as
// cached 4-level pyramids, and host copies are kept only for the debug overlay.
#pragma once
#include <cstddef>
#include <memory>
#include <unordered_map>
#include <vector>

#include "cv_compat.hpp"

struct svo_engine;

namespace svo {
// Shared ownership of one single-sequence engine (created on the first image, when the frame
// size is known). FeatureTracker creates it; ImageProcessor borrows it for the fused calls.
struct EngineHandle {
    svo_engine* e = nullptr;
    int width = 0, height = 0, device = 0;
    ~EngineHandle();
};
}  // namespace svo

class FeatureTracker {
  public:
    FeatureTracker() = default;                       // as in the reference (GPU 0)
    explicit FeatureTracker(int cuda_device) : device_(cuda_device) {}

    // start tracking `features` (with their landmark ids) from `image`
    void init(const cv::Mat& image, const std::vector<cv::Point2f>& features,
              const std::vector<size_t>& ids);

    // track the current feature set into `image`; av_parallax / percent_lost as defined at
    // src/feature_tracker.cpp:63-64
    void track_features(float& av_parallax, float& percent_lost, const cv::Mat& image,
                        bool flow_back);

    void get_tracked_features(std::vector<cv::Point2f>& features, std::vector<size_t>& ids);

    // debug overlay (src/feature_tracker.cpp:74-91), CPU-side: arrows from the initial to the
    // current position of every live feature, drawn on the initial image
    void draw_track();
    cv::Mat get_drawing();

  private:
    friend class ImageProcessor;
    std::shared_ptr<svo::EngineHandle> engine(const cv::Mat& image);
    // the fused svo_step() already ran track_features on the device: take over its result
    void adopt_tracked(const cv::Mat& image, const cv::Point2f* pts, const size_t* ids, int n);
    // the device tracker was re-initialised by svo_tracker_set(): mirror it
    void adopt_init(const cv::Mat& image, const std::vector<cv::Point2f>& features,
                    const std::vector<size_t>& ids);

    int device_ = 0;
    std::shared_ptr<svo::EngineHandle> eng_;
    cv::Mat initial_image;
    cv::Mat track_drawing;
    cv::Mat last_image;
    std::unordered_map<size_t, cv::Point2f> initial_features;
    std::vector<cv::Point2f> feature_set;
    std::vector<size_t> feature_ids;
};
