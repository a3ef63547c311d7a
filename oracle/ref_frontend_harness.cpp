// TEST INFRASTRUCTURE (never linked into the product). C entry points over the REFERENCE'S OWN
// ImageProcessor / FeatureTracker / BundleAdjuster, compiled by oracle/build_ref.py from
// /root/reference/src/{image_processor,feature_tracker,bundle_adjuster,reprojection_factor}.cpp
// where they lie, against the stand-in headers under oracle/ref_stubs/. The OpenCV algorithms
// the reference calls are served by the callbacks the test process installs (real OpenCV,
// through Python cv2); everything between those calls - process(), track_features(),
// triangulate_stereo(), add_keyframe() - is the reference's compiled code. This is the pin for
// the CONTROL FLOW of oracle/cv2_frontend.py (tests/test_reference_frontend.py).
#include <cstdint>
#include <cstring>
#include <string>

#include "image_processor.hpp"

namespace {
RefCvCallbacks g_callbacks;
std::string g_error;

struct Vo {
    std::shared_ptr<BundleAdjuster> adjuster;
    std::shared_ptr<FeatureTracker> tracker;
    std::unique_ptr<ImageProcessor> processor;
    std::shared_ptr<Keyframe> last_seen;
};
}  // namespace

extern "C" {

const RefCvCallbacks* ref_cv_callbacks() { return &g_callbacks; }
void ref_cv_set_callbacks(const RefCvCallbacks* cb) { g_callbacks = *cb; }
const char* ref_vo_last_error() { return g_error.c_str(); }

// the construction sequence of vo_node.cpp:104-119
void* ref_vo_create(float focal_length, float cx, float cy, float baseline, float min_feature_distance,
                    float parallax_thresh, uint64_t sliding_window_size) {
    cv::Mat camera_matrix = cv::Mat::eye(3, 3, CV_32F);
    camera_matrix.at<float>(0, 0) = focal_length;
    camera_matrix.at<float>(0, 2) = cx;
    camera_matrix.at<float>(1, 1) = focal_length;
    camera_matrix.at<float>(1, 2) = cy;
    CameraInfo info = {focal_length, cx, cy, 0, 0, 0, 0, baseline};
    Vo* v = new Vo;
    v->adjuster = std::make_shared<BundleAdjuster>((size_t)sliding_window_size, info);
    v->tracker = std::make_shared<FeatureTracker>();
    v->processor.reset(new ImageProcessor(camera_matrix, v->tracker, v->adjuster, baseline,
                                          min_feature_distance, parallax_thresh));
    return v;
}

void ref_vo_destroy(void* hv) { delete static_cast<Vo*>(hv); }

// ImageProcessor::process(StereoPair(left, right, t)) followed by bundle_adjust(), as the node
// loop does (vo_node.cpp:140-146). Returns 1 when the call made a keyframe, 0 when it did not,
// -1 when the reference threw (message in ref_vo_last_error()).
int ref_vo_process(void* hv, const uint8_t* left, const uint8_t* right, int width, int height, int pitch, double t) {
    Vo* v = static_cast<Vo*>(hv);
    try {
        cv::Mat l(height, width, CV_8UC1), r(height, width, CV_8UC1);
        for (int y = 0; y < height; ++y) {
            std::memcpy(l.ptr(y), left + (size_t)y * pitch, (size_t)width);
            std::memcpy(r.ptr(y), right + (size_t)y * pitch, (size_t)width);
        }
        v->processor->process(StereoPair(l, r, t));
        if (v->adjuster->get_last_keyframe() != nullptr) v->adjuster->bundle_adjust();
        std::shared_ptr<Keyframe> now = v->adjuster->get_last_keyframe();
        const bool made = now != v->last_seen;
        v->last_seen = now;
        return made ? 1 : 0;
    } catch (const std::exception& ex) {
        g_error = ex.what();
        return -1;
    }
}

// FeatureTracker::get_tracked_features
int64_t ref_vo_tracker(void* hv, float* xy, uint64_t* ids, int64_t cap) {
    std::vector<cv::Point2f> f;
    std::vector<size_t> id;
    static_cast<Vo*>(hv)->tracker->get_tracked_features(f, id);
    for (size_t i = 0; i < f.size() && (int64_t)i < cap; ++i) {
        xy[2 * i] = f[i].x; xy[2 * i + 1] = f[i].y;
        ids[i] = (uint64_t)id[i];
    }
    return (int64_t)f.size();
}

// sizes of get_last_keyframe(): tracked, new 2-D, new 3-D, new ids ; 0 when there is none
int ref_vo_keyframe_sizes(void* hv, int64_t* out4) {
    std::shared_ptr<Keyframe> kf = static_cast<Vo*>(hv)->adjuster->get_last_keyframe();
    if (!kf) return 0;
    out4[0] = (int64_t)kf->tracked_ids.size();
    out4[1] = (int64_t)kf->new_features_2d.size();
    out4[2] = (int64_t)kf->new_features_3d.size();
    out4[3] = (int64_t)kf->new_ids.size();
    return 1;
}

// contents of get_last_keyframe() into arrays sized by ref_vo_keyframe_sizes; pose7 = qw qx qy qz tx ty tz
void ref_vo_keyframe(void* hv, float* pose7, uint64_t* tracked_ids, float* tracked_2d, float* new_2d,
                     float* new_3d, uint64_t* new_ids) {
    std::shared_ptr<Keyframe> kf = static_cast<Vo*>(hv)->adjuster->get_last_keyframe();
    pose7[0] = kf->orientation.w(); pose7[1] = kf->orientation.x();
    pose7[2] = kf->orientation.y(); pose7[3] = kf->orientation.z();
    pose7[4] = kf->position(0); pose7[5] = kf->position(1); pose7[6] = kf->position(2);
    for (size_t i = 0; i < kf->tracked_ids.size(); ++i) {
        tracked_ids[i] = (uint64_t)kf->tracked_ids[i];
        tracked_2d[2 * i] = kf->tracked_features_2d[i].x;
        tracked_2d[2 * i + 1] = kf->tracked_features_2d[i].y;
    }
    for (size_t i = 0; i < kf->new_features_2d.size(); ++i) {
        new_2d[2 * i] = kf->new_features_2d[i].x;
        new_2d[2 * i + 1] = kf->new_features_2d[i].y;
    }
    for (size_t i = 0; i < kf->new_features_3d.size(); ++i) {
        new_3d[3 * i] = kf->new_features_3d[i].x;
        new_3d[3 * i + 1] = kf->new_features_3d[i].y;
        new_3d[3 * i + 2] = kf->new_features_3d[i].z;
    }
    for (size_t i = 0; i < kf->new_ids.size(); ++i) new_ids[i] = (uint64_t)kf->new_ids[i];
}

}  // extern "C"
