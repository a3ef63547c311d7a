// image_processor.hpp - the reference's ImageProcessor interface (src/image_processor.hpp:9-83)
// on top of libsvo.so. process() keeps the reference's control flow
// (src/image_processor.cpp:18-163); detection, tracking, dedupe, block matching and
// reprojection are calls into the CUDA library. The pose step (src/image_processor.cpp:75-92):
// built with OpenCV it is cv::solvePnPRansac + cv::Rodrigues on the host, as in the reference;
// built without, it runs on the GPU (svo_pnp_ransac) unless the caller installs a host solver
// (set_pnp_solver).
#pragma once
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "bundle_adjuster.hpp"
#include "cv_compat.hpp"
#include "feature_tracker.hpp"

struct StereoPair {
    cv::Mat left;
    cv::Mat right;
    double t;
    StereoPair(const cv::Mat& l, const cv::Mat& r, double stamp) : left(l), right(r), t(stamp) {}
};

namespace svo {

// Host PnP step: world points, their image positions, K (3x3 CV_32F). rvec/tvec (3x1 CV_64F)
// enter as the extrinsic guess and leave as the solution; rmat receives the 3x3 rotation
// (row major, double); inliers receives indices into the input. Returns false on failure.
using PnpSolver = std::function<bool(const std::vector<cv::Point3f>& world,
                                     const std::vector<cv::Point2f>& image, const cv::Mat& K,
                                     cv::Mat& rvec, cv::Mat& tvec, double rmat[9],
                                     std::vector<int>& inliers)>;

// What process() did with the last pair (taps for the harness and the parity tests; the
// reference exposes none of this).
struct FrameRecord {
    enum Decision { NONE = 0, TOO_FEW = 1, FIRST = 2, SKIP = 3, KEYFRAME = 4 };
    Decision decision = NONE;
    std::vector<cv::Point2f> detected;
    float av_parallax = 0.f, percent_lost = 0.f;
    std::vector<cv::Point2f> tracked;          // after track_features
    std::vector<size_t> tracked_ids;
    std::vector<int> inliers;                  // PnP inliers (indices into tracked)
    float hmat[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};   // camera -> world
    std::vector<cv::Point2f> new_features;     // detections that survived the dedupe
    std::vector<int> valid_idx;                // indices into new_features with disparity > 0
    std::vector<short> disp16;                 // StereoBM 1/16 px value at every new feature
    std::vector<cv::Point2f> new_2d;
    std::vector<cv::Point3f> new_3d;
    std::vector<size_t> new_ids;
    std::vector<cv::Point2f> tracker_pts;      // tracker set after this frame
    std::vector<size_t> tracker_ids;
};

}  // namespace svo

class ImageProcessor {
  public:
    ImageProcessor(cv::Mat cam_mat, std::shared_ptr<FeatureTracker> tracker,
                   std::shared_ptr<BundleAdjuster> adjuster, float bline,
                   float min_feature_distance, float parallax_thresh);

    void process(const StereoPair& stereo_pair);

    // ---- additions (do not exist in the reference)
    void set_pnp_solver(svo::PnpSolver solver) { pnp_ = std::move(solver); }
    const svo::FrameRecord& last_record() const { return rec_; }

  private:
    std::shared_ptr<BundleAdjuster> bundle_adjuster;
    std::shared_ptr<FeatureTracker> feature_tracker;
    cv::Mat camera_matrix;
    float baseline;
    float min_feature_distance;
    float parallax_thresh;
    cv::Mat rvec;
    cv::Mat tvec;
    svo::PnpSolver pnp_;
    svo::FrameRecord rec_;

    // stand-alone form of the reference's helper (src/image_processor.cpp:165-208); process()
    // itself uses the fused svo_keyframe() call, which does the same work on resident frames
    void triangulate_stereo(std::vector<cv::Point3f>& features_3d,
                            std::vector<cv::Point2f>& valid_features_2d,
                            const std::vector<cv::Point2f>& left_features_2d, const cv::Mat& left,
                            const cv::Mat& right, const cv::Mat& camera_pose);
};
