// bundle_adjuster.hpp - the consumer side of the front end, reduced to what the front end
// needs from it. Same names and call signatures as the reference (src/bundle_adjuster.hpp:22-46
// Keyframe, :77-151 BundleAdjuster) so ImageProcessor compiles against either:
//   * a build with Ceres + Eigen links the reference's own bundle_adjuster.cpp unchanged and
//     never includes this file;
//   * this header is the Ceres-free build: the id/refcount/world-point bookkeeping of
//     src/bundle_adjuster.cpp:60-135,159-163 without the optimisation (bundle_adjust() is a
//     no-op), which is all the headless harness and the parity tests can exercise here.
#pragma once
#include <cstddef>
#include <memory>
#include <queue>
#include <stack>
#include <utility>
#include <vector>

#include "camera_info.hpp"
#include "cv_compat.hpp"

namespace svo {
// Stand-ins for Eigen::Vector3f / Eigen::Quaternionf in the Keyframe aggregate (Eigen is not
// available in this build): same accessor spelling as the reference uses on them.
struct Vec3f {
    float v[3] = {0.f, 0.f, 0.f};
    float& operator()(int i) { return v[i]; }
    float operator()(int i) const { return v[i]; }
};
struct Quatf {
    float q[4] = {1.f, 0.f, 0.f, 0.f};   // w x y z
    float& w() { return q[0]; } float& x() { return q[1]; } float& y() { return q[2]; } float& z() { return q[3]; }
    float w() const { return q[0]; } float x() const { return q[1]; } float y() const { return q[2]; } float z() const { return q[3]; }
    // rotation matrix (row major 3x3) -> unit quaternion
    static Quatf from_matrix(const float R[9]);
};
}  // namespace svo

// camera pose of a keyframe (world frame expressed in the camera frame), its image, the
// features carried over from the previous keyframe and the newly triangulated ones
struct Keyframe {
    svo::Vec3f position;
    svo::Quatf orientation;
    cv::Mat image;
    std::vector<cv::Point2f> tracked_features_2d;
    std::vector<size_t> tracked_ids;
    std::vector<cv::Point2f> new_features_2d;
    std::vector<cv::Point3f> new_features_3d;
    std::vector<size_t> new_ids;

    Keyframe(svo::Vec3f position_, svo::Quatf orientation_, cv::Mat image_,
             std::vector<cv::Point2f> tracked_2d, std::vector<size_t> tracked_ids_,
             std::vector<cv::Point2f> new_2d, std::vector<cv::Point3f> new_3d)
        : position(position_), orientation(orientation_), image(std::move(image_)),
          tracked_features_2d(std::move(tracked_2d)), tracked_ids(std::move(tracked_ids_)),
          new_features_2d(std::move(new_2d)), new_features_3d(std::move(new_3d)) {}
};

// per-keyframe feature limit (src/bundle_adjuster.hpp:75)
static const size_t max_features = 400;

class BundleAdjuster {
  public:
    BundleAdjuster(size_t window_size, CameraInfo info);

    std::shared_ptr<Keyframe> get_last_keyframe() { return last_keyframe_; }

    // Gives ids to the keyframe's new features (appending them to keyframe->new_ids), clips
    // them to the feature limit, slides the window. src/bundle_adjuster.cpp:60-135.
    void add_keyframe(std::shared_ptr<Keyframe> keyframe);

    // no optimiser in this build: keeps the poses PnP produced
    void bundle_adjust();

    // appends the stored world point of every id (src/bundle_adjuster.cpp:159-163)
    void get_world_points(std::vector<cv::Point3f>& world_points, const std::vector<size_t>& ids);

    size_t num_features() const { return features_.size(); }

  private:
    struct Feature {
        double position[3] = {0, 0, 0};
        size_t refcount = 0;
    };
    void remove_oldest_pose();

    std::vector<Feature> features_;
    std::stack<size_t> avail_ids_;
    std::queue<std::vector<size_t>> window_;     // observed feature ids per windowed keyframe
    size_t window_size_;
    std::shared_ptr<Keyframe> last_keyframe_;
    CameraInfo camera_info_;
    bool new_frame_added_ = false;
};
