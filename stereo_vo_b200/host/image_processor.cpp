// ImageProcessor facade over libsvo.so: the reference's process() control flow
// (src/image_processor.cpp:18-163) with every pixel stage replaced by a C-ABI call.
//   :22      goodFeaturesToTrack            \  one svo_step(): upload the pair, pyramid, detect,
//   :62      track_features (fwd+bwd LK)    /  and - gated on >= 4 corners like :23-25 - track
//   :113-142 dedupe + triangulate_stereo       one svo_keyframe()
//   :51,162  FeatureTracker::init              one svo_tracker_set() (frame already resident)
// PnP-RANSAC (:75-92) and the BundleAdjuster calls stay on the host, as in the reference.
#include "stereo_vo/image_processor.hpp"

#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>

#include "svo_c.h"

#if SVO_HAVE_OPENCV
#include <opencv2/calib3d.hpp>
#endif

namespace svo {
void check(svo_engine* e, int rc, const char* what);
const uint8_t* image_bytes(const cv::Mat& m, int w, int h);
}  // namespace svo

ImageProcessor::ImageProcessor(cv::Mat cam_mat, std::shared_ptr<FeatureTracker> tracker,
                               std::shared_ptr<BundleAdjuster> adjuster, float bline,
                               float min_feature_distance_, float parallax_thresh_)
    : bundle_adjuster(std::move(adjuster)), feature_tracker(std::move(tracker)),
      camera_matrix(std::move(cam_mat)), baseline(bline),
      min_feature_distance(min_feature_distance_), parallax_thresh(parallax_thresh_) {
#if SVO_HAVE_OPENCV
    // the reference's own host step (src/image_processor.cpp:76-85)
    pnp_ = [](const std::vector<cv::Point3f>& world, const std::vector<cv::Point2f>& image,
              const cv::Mat& K, cv::Mat& rv, cv::Mat& tv, double rmat[9],
              std::vector<int>& inliers) {
        cv::Mat inl;
        cv::solvePnPRansac(world, image, K, cv::Mat::zeros(4, 1, CV_32F), rv, tv, true, 100, 8.0,
                           0.99, inl);
        inliers.clear();
        for (int i = 0; i < inl.rows; ++i) inliers.push_back(inl.at<int>(i, 0));
        cv::Mat R;
        cv::Rodrigues(rv, R);
        R.convertTo(R, CV_64F);
        for (int i = 0; i < 9; ++i) rmat[i] = R.at<double>(i / 3, i % 3);
        return true;
    };
#endif
}

namespace {

struct Intrinsics {
    float focal, cx, cy;
};
// src/image_processor.cpp:178-180: read back from the CV_32F camera matrix
Intrinsics intrinsics(const cv::Mat& K) {
    if (K.empty() || K.rows != 3 || K.cols != 3 || K.type() != CV_32F)
        throw std::invalid_argument("ImageProcessor: camera matrix must be 3x3 CV_32F");
    return {K.at<float>(0, 0), K.at<float>(0, 2), K.at<float>(1, 2)};
}

// hmat = [R^T | -R^T t] as float (src/image_processor.cpp:130-134); R row major, double
void pose_inverse(const double R[9], const cv::Mat& tvec, float hmat[16]) {
    double t[3];
    for (int i = 0; i < 3; ++i)
        t[i] = tvec.type() == CV_64F ? tvec.at<double>(i) : (double)tvec.at<float>(i);
    for (int i = 0; i < 16; ++i) hmat[i] = (i % 5 == 0) ? 1.f : 0.f;
    for (int r = 0; r < 3; ++r) {
        double acc = 0.0;
        for (int c = 0; c < 3; ++c) {
            hmat[r * 4 + c] = (float)R[c * 3 + r];
            acc += R[c * 3 + r] * t[c];
        }
        hmat[r * 4 + 3] = (float)(-acc);
    }
}

}  // namespace

void ImageProcessor::process(const StereoPair& stereo_pair) {
    svo::FrameRecord& rec = rec_;
    rec = svo::FrameRecord();
    const cv::Mat& left = stereo_pair.left;
    const cv::Mat& right = stereo_pair.right;
    auto h = feature_tracker->engine(left);
    svo_engine* e = h->e;
    svo_params p;
    svo_get_params(e, &p);
    if (p.min_distance != min_feature_distance) {
        // the constructor argument is the detector's minDistance (src/image_processor.cpp:22);
        // the engine was created by the tracker with the default: rebuild it once
        if (feature_tracker->feature_ids.size() || bundle_adjuster->get_last_keyframe())
            throw std::logic_error("ImageProcessor: min_feature_distance changed mid-sequence");
        svo_destroy(e);
        h->e = nullptr;
        p.min_distance = min_feature_distance;
        int rc = svo_create(&p, h->device, 1, &h->e);
        if (rc != SVO_OK) throw std::runtime_error(std::string("svo_create failed: ") + svo_last_error(nullptr));
        e = h->e;
    }
    if (right.step != left.step)
        throw std::invalid_argument("ImageProcessor: left and right images must share one pitch");
    const uint8_t* lp = svo::image_bytes(left, h->width, h->height);
    const uint8_t* rp = svo::image_bytes(right, h->width, h->height);

    // ---- detect (+ track when the tracker holds a feature set) on the device
    const int32_t seq = 0;
    std::vector<svo_float2> corners((size_t)p.max_corners), tracked((size_t)p.max_features);
    std::vector<uint64_t> tracked_ids((size_t)p.max_features);
    int32_t n_corners = 0;
    svo_track_stats st;
    svo::check(e, svo_step(e, 1, &seq, &lp, &rp, (int)left.step, corners.data(), &n_corners, &st,
                           tracked.data(), tracked_ids.data()), "svo_step");
    rec.detected.resize((size_t)n_corners);
    for (int i = 0; i < n_corners; ++i) rec.detected[i] = cv::Point2f(corners[i].x, corners[i].y);
    if (n_corners < 4) {                                   // src/image_processor.cpp:23-25
        rec.decision = svo::FrameRecord::TOO_FEW;
        return;
    }
    const Intrinsics in = intrinsics(camera_matrix);
    const int mc = p.max_corners;
    std::vector<svo_float2> new_feats((size_t)mc), valid_2d((size_t)mc);
    std::vector<svo_float3> pts_3d((size_t)mc);
    std::vector<int32_t> valid_idx((size_t)mc);
    std::vector<int16_t> disp16((size_t)mc);
    int32_t n_new = 0, n_valid = 0;

    auto fill_new = [&](Keyframe& kf) {
        rec.new_features.resize((size_t)n_new);
        rec.disp16.assign(disp16.begin(), disp16.begin() + n_new);
        for (int i = 0; i < n_new; ++i) rec.new_features[i] = cv::Point2f(new_feats[i].x, new_feats[i].y);
        rec.valid_idx.assign(valid_idx.begin(), valid_idx.begin() + n_valid);
        for (int i = 0; i < n_valid; ++i) {
            kf.new_features_2d.push_back(cv::Point2f(valid_2d[i].x, valid_2d[i].y));
            kf.new_features_3d.push_back(cv::Point3f(pts_3d[i].x, pts_3d[i].y, pts_3d[i].z));
        }
    };
    auto set_tracker = [&](const std::vector<cv::Point2f>& pts, const std::vector<size_t>& ids) {
        if (pts.size() > (size_t)p.max_features)
            throw std::length_error("ImageProcessor: tracker set exceeds max_features");
        std::vector<uint64_t> ids64(ids.begin(), ids.end());
        const int32_t count = (int32_t)pts.size();
        std::vector<svo_float2> buf((size_t)p.max_features);
        std::vector<uint64_t> idbuf((size_t)p.max_features);
        for (int i = 0; i < count; ++i) { buf[i].x = pts[i].x; buf[i].y = pts[i].y; idbuf[i] = ids64[i]; }
        svo::check(e, svo_tracker_set(e, 1, &seq, buf.data(), idbuf.data(), &count), "svo_tracker_set");
        feature_tracker->adopt_init(left, pts, ids);
        rec.tracker_pts = pts;
        rec.tracker_ids = ids;
    };

    std::shared_ptr<Keyframe> last_keyframe = bundle_adjuster->get_last_keyframe();
    if (last_keyframe == nullptr) {                        // :30-58 first keyframe
        const float eye[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        const int32_t zero = 0;
        svo::check(e, svo_keyframe(e, 1, &seq, nullptr, &zero, eye, in.focal, in.cx, in.cy, baseline,
                                   new_feats.data(), &n_new, valid_2d.data(), pts_3d.data(),
                                   valid_idx.data(), disp16.data(), &n_valid), "svo_keyframe");
        auto kf = std::make_shared<Keyframe>(svo::Vec3f(), svo::Quatf(), left,
                                             std::vector<cv::Point2f>(), std::vector<size_t>(),
                                             std::vector<cv::Point2f>(), std::vector<cv::Point3f>());
        fill_new(*kf);
        bundle_adjuster->add_keyframe(kf);
        set_tracker(kf->new_features_2d, kf->new_ids);
        tvec = cv::Mat::zeros(3, 1, CV_64F);
        rvec = cv::Mat::zeros(3, 1, CV_64F);
        rec.decision = svo::FrameRecord::FIRST;
        rec.new_2d = kf->new_features_2d;
        rec.new_3d = kf->new_features_3d;
        rec.new_ids = kf->new_ids;
        return;
    }

    // ---- track_features already ran inside svo_step (:62)
    const float av_parallax = st.av_parallax, percent_lost = st.percent_lost;
    std::vector<cv::Point2f> tracked_features((size_t)st.n_kept);
    std::vector<size_t> ids((size_t)st.n_kept);
    for (int i = 0; i < st.n_kept; ++i) {
        tracked_features[i] = cv::Point2f(tracked[i].x, tracked[i].y);
        ids[i] = (size_t)tracked_ids[i];
    }
    feature_tracker->adopt_tracked(left, tracked_features.data(), ids.data(), st.n_kept);
    rec.av_parallax = av_parallax;
    rec.percent_lost = percent_lost;
    rec.tracked = tracked_features;
    rec.tracked_ids = ids;
    if (av_parallax <= parallax_thresh && percent_lost < 0.4) {   // :63-65 (NaN falls through)
        rec.decision = svo::FrameRecord::SKIP;
        return;
    }

    std::vector<cv::Point3f> world;
    bundle_adjuster->get_world_points(world, ids);          // :72
    double rmat[9];
    std::vector<int> inliers;
    if (pnp_) {
        if (!pnp_(world, tracked_features, camera_matrix, rvec, tvec, rmat, inliers))
            throw std::runtime_error("ImageProcessor: PnP solver failed");
    } else {
        // no host solver installed: the pose step runs on the GPU with the reference's
        // solvePnPRansac arguments (:80: extrinsic guess, 100 iterations, 8 px, 0.99)
        const int32_t n_pts = (int32_t)world.size();
        std::vector<svo_float3> w((size_t)p.max_features);
        std::vector<svo_float2> im((size_t)p.max_features);
        std::vector<int32_t> inl((size_t)p.max_features);
        for (int32_t i = 0; i < n_pts; ++i) {
            w[(size_t)i] = {world[(size_t)i].x, world[(size_t)i].y, world[(size_t)i].z};
            im[(size_t)i] = {tracked_features[(size_t)i].x, tracked_features[(size_t)i].y};
        }
        double guess[6];
        for (int i = 0; i < 3; ++i) {
            guess[i] = rvec.type() == CV_64F ? rvec.at<double>(i) : (double)rvec.at<float>(i);
            guess[3 + i] = tvec.type() == CV_64F ? tvec.at<double>(i) : (double)tvec.at<float>(i);
        }
        svo_pnp_result r;
        svo::check(e, svo_pnp_ransac(e, 1, w.data(), im.data(), &n_pts, in.focal, in.cx, in.cy, 100,
                                     8.f, 0.99, guess, &r, inl.data()), "svo_pnp_ransac");
        rvec = cv::Mat::zeros(3, 1, CV_64F);
        tvec = cv::Mat::zeros(3, 1, CV_64F);
        for (int i = 0; i < 3; ++i) {
            rvec.at<double>(i) = r.rvec[i];
            tvec.at<double>(i) = r.tvec[i];
        }
        for (int i = 0; i < 9; ++i) rmat[i] = r.rmat[i];
        inliers.assign(inl.begin(), inl.begin() + (r.ok ? r.n_inliers : 0));
    }
    const size_t num_inliers = inliers.size();
    // the solver is caller-supplied: a list longer than the tracker set (duplicates) must not
    // run past inl_pts before svo_keyframe gets to reject it
    if (num_inliers > (size_t)p.max_features || num_inliers > tracked_features.size())
        throw std::length_error("ImageProcessor: PnP solver returned more inliers than tracked features");
    float rmat_f[9];
    for (int i = 0; i < 9; ++i) rmat_f[i] = (float)rmat[i];
    svo::Vec3f position;
    for (int i = 0; i < 3; ++i)
        position(i) = tvec.type() == CV_64F ? (float)tvec.at<double>(i) : tvec.at<float>(i);
    auto kf = std::make_shared<Keyframe>(position, svo::Quatf::from_matrix(rmat_f), left,
                                         std::vector<cv::Point2f>(num_inliers),
                                         std::vector<size_t>(num_inliers),
                                         std::vector<cv::Point2f>(), std::vector<cv::Point3f>());
    std::vector<svo_float2> inl_pts((size_t)p.max_features);
    for (size_t i = 0; i < num_inliers; ++i) {             // :103-108
        const size_t idx = (size_t)inliers[i];
        kf->tracked_ids[i] = ids.at(idx);
        kf->tracked_features_2d[i] = tracked_features.at(idx);
        inl_pts[i].x = tracked_features[idx].x;
        inl_pts[i].y = tracked_features[idx].y;
    }
    pose_inverse(rmat, tvec, rec.hmat);                     // :130-134
    const int32_t n_inl = (int32_t)num_inliers;
    svo::check(e, svo_keyframe(e, 1, &seq, inl_pts.data(), &n_inl, rec.hmat, in.focal, in.cx, in.cy,
                               baseline, new_feats.data(), &n_new, valid_2d.data(), pts_3d.data(),
                               valid_idx.data(), disp16.data(), &n_valid), "svo_keyframe");
    fill_new(*kf);
    bundle_adjuster->add_keyframe(kf);                      // :144
    feature_tracker->draw_track();                          // :146

    std::vector<cv::Point2f> pts_for_tracker(kf->tracked_features_2d);   // :148-162
    pts_for_tracker.insert(pts_for_tracker.end(), kf->new_features_2d.begin(), kf->new_features_2d.end());
    std::vector<size_t> ids_for_tracker(kf->tracked_ids);
    ids_for_tracker.insert(ids_for_tracker.end(), kf->new_ids.begin(), kf->new_ids.end());
    set_tracker(pts_for_tracker, ids_for_tracker);
    rec.decision = svo::FrameRecord::KEYFRAME;
    rec.inliers = inliers;
    rec.new_2d = kf->new_features_2d;
    rec.new_3d = kf->new_features_3d;
    rec.new_ids = kf->new_ids;
}

void ImageProcessor::triangulate_stereo(std::vector<cv::Point3f>& features_3d,
                                        std::vector<cv::Point2f>& valid_features_2d,
                                        const std::vector<cv::Point2f>& left_features_2d,
                                        const cv::Mat& left, const cv::Mat& right,
                                        const cv::Mat& camera_pose) {
    auto h = feature_tracker->engine(left);
    const Intrinsics in = intrinsics(camera_matrix);
    if (camera_pose.rows != 4 || camera_pose.cols != 4 || camera_pose.type() != CV_32F)
        throw std::invalid_argument("triangulate_stereo: camera_pose must be 4x4 CV_32F");
    float pose[16];
    for (int i = 0; i < 16; ++i) pose[i] = camera_pose.at<float>(i / 4, i % 4);
    const int n = (int)left_features_2d.size();
    std::vector<svo_float3> p3((size_t)n + 1);
    std::vector<svo_float2> v2((size_t)n + 1);
    std::vector<int32_t> vi((size_t)n + 1);
    std::vector<int16_t> d16((size_t)n + 1);
    int32_t nv = 0;
    svo::check(h->e,
               svo_triangulate_stereo(h->e, 0, svo::image_bytes(left, h->width, h->height),
                                      svo::image_bytes(right, h->width, h->height), (int)left.step,
                                      reinterpret_cast<const svo_float2*>(left_features_2d.data()),
                                      n, pose, in.focal, in.cx, in.cy, baseline, p3.data(),
                                      v2.data(), vi.data(), d16.data(), &nv),
               "svo_triangulate_stereo");
    for (int i = 0; i < nv; ++i) {
        valid_features_2d.push_back(cv::Point2f(v2[i].x, v2[i].y));
        features_3d.push_back(cv::Point3f(p3[i].x, p3[i].y, p3[i].z));
    }
}
