// Ceres-free BundleAdjuster: id assignment, refcounts, sliding window and the world-point store
// of the reference (src/bundle_adjuster.cpp:41-58 remove_oldest_pose, :60-135 add_keyframe,
// :159-163 get_world_points). The factor graph itself is not built: there is nothing to solve
// with in this build, bundle_adjust() leaves the PnP poses untouched.
#include "stereo_vo/bundle_adjuster.hpp"

#include <cmath>

namespace svo {

Quatf Quatf::from_matrix(const float R[9]) {
    // branch on the largest of (trace, r00, r11, r22) to keep the divisor away from zero
    Quatf o;
    const float r00 = R[0], r11 = R[4], r22 = R[8];
    const float tr = r00 + r11 + r22;
    if (tr > 0.f) {
        float s = std::sqrt(tr + 1.f) * 2.f;
        o.q[0] = 0.25f * s;
        o.q[1] = (R[7] - R[5]) / s;
        o.q[2] = (R[2] - R[6]) / s;
        o.q[3] = (R[3] - R[1]) / s;
    } else if (r00 > r11 && r00 > r22) {
        float s = std::sqrt(1.f + r00 - r11 - r22) * 2.f;
        o.q[0] = (R[7] - R[5]) / s;
        o.q[1] = 0.25f * s;
        o.q[2] = (R[1] + R[3]) / s;
        o.q[3] = (R[2] + R[6]) / s;
    } else if (r11 > r22) {
        float s = std::sqrt(1.f + r11 - r00 - r22) * 2.f;
        o.q[0] = (R[2] - R[6]) / s;
        o.q[1] = (R[1] + R[3]) / s;
        o.q[2] = 0.25f * s;
        o.q[3] = (R[5] + R[7]) / s;
    } else {
        float s = std::sqrt(1.f + r22 - r00 - r11) * 2.f;
        o.q[0] = (R[3] - R[1]) / s;
        o.q[1] = (R[2] + R[6]) / s;
        o.q[2] = (R[5] + R[7]) / s;
        o.q[3] = 0.25f * s;
    }
    return o;
}

}  // namespace svo

BundleAdjuster::BundleAdjuster(size_t window_size, CameraInfo info)
    : window_size_(window_size), camera_info_(info) {}

void BundleAdjuster::remove_oldest_pose() {
    // every observation of the oldest keyframe releases its feature; a feature nobody in the
    // window observes any more frees its id for reuse (src/bundle_adjuster.cpp:43-53)
    for (size_t id : window_.front()) {
        if (--features_[id].refcount == 0) avail_ids_.push(id);
    }
    window_.pop();
}

void BundleAdjuster::add_keyframe(std::shared_ptr<Keyframe> keyframe) {
    std::vector<size_t> observed;
    const size_t num_tracked = keyframe->tracked_ids.size();
    observed.reserve(num_tracked + keyframe->new_features_2d.size());
    for (size_t i = 0; i < num_tracked; ++i) {
        const size_t id = keyframe->tracked_ids[i];
        features_.at(id).refcount++;
        observed.push_back(id);
    }
    // Feature limit. The reference also resizes the (still empty) new_ids vector here, which
    // leaves max_new zero ids in front of the real ones and breaks FeatureTracker::init
    // (latent bug); this build clips the new features and keeps ids aligned.
    const size_t max_new = max_features > num_tracked ? max_features - num_tracked : 0;
    if (keyframe->new_features_2d.size() > max_new) {
        keyframe->new_features_2d.resize(max_new);
        keyframe->new_features_3d.resize(max_new);
    }
    const size_t num_new = keyframe->new_features_2d.size();
    for (size_t i = 0; i < num_new; ++i) {
        size_t id;
        if (!avail_ids_.empty()) {
            id = avail_ids_.top();
            avail_ids_.pop();
        } else {
            features_.emplace_back();
            id = features_.size() - 1;
        }
        Feature& f = features_[id];
        f.position[0] = keyframe->new_features_3d[i].x;
        f.position[1] = keyframe->new_features_3d[i].y;
        f.position[2] = keyframe->new_features_3d[i].z;
        f.refcount = 1;          // the reference starts at 1 and then counts the observation
        f.refcount++;            // as well, so a feature id is never recycled
        keyframe->new_ids.push_back(id);
        observed.push_back(id);
    }
    window_.push(std::move(observed));
    if (window_.size() > window_size_) remove_oldest_pose();
    last_keyframe_ = keyframe;
    new_frame_added_ = true;
}

void BundleAdjuster::bundle_adjust() { new_frame_added_ = false; }

void BundleAdjuster::get_world_points(std::vector<cv::Point3f>& world_points,
                                      const std::vector<size_t>& ids) {
    for (size_t id : ids) {
        const Feature& f = features_.at(id);
        world_points.push_back(cv::Point3f((float)f.position[0], (float)f.position[1],
                                           (float)f.position[2]));
    }
}
