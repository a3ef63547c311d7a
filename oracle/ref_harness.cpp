// TEST INFRASTRUCTURE (never linked into the product). C entry points over the REFERENCE'S OWN
// BundleAdjuster and ReprojectionFactor, compiled by oracle/build_ref.py from
// /root/reference/src/bundle_adjuster.cpp and reprojection_factor.cpp where they lie, against the
// stand-in headers under oracle/ref_stubs/ (Eigen, Ceres and OpenCV are not in this image; the
// stand-in Ceres records the problem and does not solve). What runs behind these calls is the
// reference's add_keyframe / remove_oldest_pose / get_world_points / bundle_adjust /
// ReprojectionFactor::Evaluate, instruction for instruction - the pin for
// oracle/cv2_frontend.py:BundleAdjusterBookkeeping, stereo_vo_b200/host/bundle_adjuster.cpp and
// oracle/reprojection_factor.py (tests/golden/make_golden_ba.py, tests/test_reference_ba.py).
#include <cstdint>
#include <cstring>

#include "bundle_adjuster.hpp"
#include "reprojection_factor.hpp"

namespace {
struct Handle {
    BundleAdjuster* ba;
    ceres::Problem* problem;       // the one the adjuster's constructor made (stub registry)
    // the observation each cost function was constructed with (ReprojectionFactor keeps it
    // private): the harness knows it because add_keyframe adds the blocks of a keyframe in the
    // order tracked features, then new features (bundle_adjuster.cpp:73-85,94-125)
    std::map<const ceres::CostFunction*, std::pair<double, double>> obs;
};
}  // namespace

extern "C" {

void* ref_ba_create(uint64_t window, double focal, double cx, double cy, double baseline) {
    CameraInfo info = {focal, cx, cy, 0, 0, 0, 0, baseline};      // vo_node.cpp:110
    Handle* h = new Handle;
    h->ba = new BundleAdjuster((size_t)window, info);
    h->problem = ceres::last_problem();
    return h;
}

void ref_ba_destroy(void* hv) {
    Handle* h = static_cast<Handle*>(hv);
    delete h->ba;
    delete h;
}

uint64_t ref_ba_max_features() { return (uint64_t)max_features; }

// BundleAdjuster::add_keyframe on a Keyframe built like image_processor.cpp:41-47,136-144 builds
// it. pose7 = qw qx qy qz tx ty tz. Returns new_ids.size() after the call and copies
// min(that, cap) ids; n_new_after = new_features_2d.size() after the call (the truncation).
int64_t ref_ba_add_keyframe(void* hv, const float* pose7, const uint64_t* tracked_ids,
                            const float* tracked_2d, int64_t n_tracked, const float* new_2d,
                            const float* new_3d, int64_t n_new, uint64_t* out_new_ids, int64_t cap,
                            int64_t* n_new_after) {
    Handle* h = static_cast<Handle*>(hv);
    std::vector<cv::Point2f> t2d, n2d;
    std::vector<size_t> tids;
    std::vector<cv::Point3f> n3d;
    for (int64_t i = 0; i < n_tracked; ++i) {
        t2d.push_back(cv::Point2f(tracked_2d[2 * i], tracked_2d[2 * i + 1]));
        tids.push_back((size_t)tracked_ids[i]);
    }
    for (int64_t i = 0; i < n_new; ++i) {
        n2d.push_back(cv::Point2f(new_2d[2 * i], new_2d[2 * i + 1]));
        n3d.push_back(cv::Point3f(new_3d[3 * i], new_3d[3 * i + 1], new_3d[3 * i + 2]));
    }
    std::shared_ptr<Keyframe> kf = std::make_shared<Keyframe>(
        Vector3f(pose7[4], pose7[5], pose7[6]), Quaternionf(pose7[0], pose7[1], pose7[2], pose7[3]),
        cv::Mat(), t2d, tids, n2d, n3d);
    h->ba->add_keyframe(kf);
    {
        // this keyframe's blocks are the newest n_tracked + new_features_2d.size() of the problem
        // (the oldest pose's blocks, if the window slid, left from the front)
        const std::vector<ceres::ResidualBlock*>& blocks = h->problem->blocks();
        const size_t n_obs = (size_t)n_tracked + kf->new_features_2d.size();
        const size_t first = blocks.size() - n_obs;
        for (size_t i = 0; i < n_obs; ++i) {
            const cv::Point2f& o = i < (size_t)n_tracked ? t2d[i] : kf->new_features_2d[i - (size_t)n_tracked];
            h->obs[blocks[first + i]->cost] = std::make_pair((double)o.x, (double)o.y);
        }
    }
    const int64_t n_ids = (int64_t)kf->new_ids.size();
    for (int64_t i = 0; i < n_ids && i < cap; ++i) out_new_ids[i] = (uint64_t)kf->new_ids[(size_t)i];
    if (n_new_after) *n_new_after = (int64_t)kf->new_features_2d.size();
    return n_ids;
}

void ref_ba_bundle_adjust(void* hv) { static_cast<Handle*>(hv)->ba->bundle_adjust(); }

// pose of get_last_keyframe() (qw qx qy qz tx ty tz); returns 0 when there is none yet
int ref_ba_last_pose(void* hv, float* pose7) {
    std::shared_ptr<Keyframe> kf = static_cast<Handle*>(hv)->ba->get_last_keyframe();
    if (!kf) return 0;
    pose7[0] = kf->orientation.w(); pose7[1] = kf->orientation.x();
    pose7[2] = kf->orientation.y(); pose7[3] = kf->orientation.z();
    pose7[4] = kf->position(0); pose7[5] = kf->position(1); pose7[6] = kf->position(2);
    return 1;
}

void ref_ba_world_points(void* hv, const uint64_t* ids, int64_t n, float* out_xyz) {
    std::vector<size_t> v;
    for (int64_t i = 0; i < n; ++i) v.push_back((size_t)ids[i]);
    std::vector<cv::Point3f> pts;
    static_cast<Handle*>(hv)->ba->get_world_points(pts, v);
    for (size_t i = 0; i < pts.size(); ++i) {
        out_xyz[3 * i] = pts[i].x; out_xyz[3 * i + 1] = pts[i].y; out_xyz[3 * i + 2] = pts[i].z;
    }
}

// state of the problem the reference built: [0] live residual blocks, [1] parameter blocks,
// [2] API misuse count, [3] residual blocks that went away with a removed parameter block,
// [4] distinct pose blocks among the residual blocks, [5] of those, how many are held constant
void ref_ba_problem_stats(void* hv, int64_t* out6) {
    ceres::Problem* p = static_cast<Handle*>(hv)->problem;
    std::set<double*> poses;
    for (const ceres::ResidualBlock* b : p->blocks()) poses.insert(b->pose);
    int64_t n_const = 0;
    for (double* q : poses) n_const += p->is_constant(q) ? 1 : 0;
    out6[0] = (int64_t)p->blocks().size();
    out6[1] = (int64_t)p->num_parameter_blocks();
    out6[2] = p->errors;
    out6[3] = p->dropped_with_parameter;
    out6[4] = (int64_t)poses.size();
    out6[5] = n_const;
}

// every live residual block evaluated by the reference's ReprojectionFactor::Evaluate at the
// current parameter values, in the order the blocks were added: parameters (7 + 3), residual
// (2), d r / d pose (2 x 7 row major), d r / d point (2 x 3), observation (2) = 34 doubles per
// block.
// Returns the number of blocks written (<= cap).
int64_t ref_ba_evaluate_window(void* hv, double* out, int64_t cap) {
    ceres::Problem* p = static_cast<Handle*>(hv)->problem;
    int64_t n = 0;
    for (const ceres::ResidualBlock* b : p->blocks()) {
        if (n >= cap) break;
        double* o = out + 34 * n;
        std::memcpy(o, b->pose, 7 * sizeof(double));
        std::memcpy(o + 7, b->point, 3 * sizeof(double));
        const double* params[2] = {b->pose, b->point};
        double* jac[2] = {o + 12, o + 26};
        b->cost->Evaluate(params, o + 10, jac);
        const std::pair<double, double>& ob = static_cast<Handle*>(hv)->obs[b->cost];
        o[32] = ob.first;
        o[33] = ob.second;
        ++n;
    }
    return n;
}

// one ReprojectionFactor(ox, oy, info).Evaluate(pose7, point3)
void ref_factor_evaluate(double ox, double oy, double focal, double cx, double cy, const double* pose7,
                         const double* point3, double* resid2, double* jac_pose14, double* jac_point6) {
    CameraInfo info = {focal, cx, cy, 0, 0, 0, 0, 0};
    ReprojectionFactor f(ox, oy, info);
    const double* params[2] = {pose7, point3};
    double* jac[2] = {jac_pose14, jac_point6};
    f.Evaluate(params, resid2, jac);
}

}  // extern "C"
