// C entry points of libsvo_host.so (include/svo_host_c.h): thin, exception-tight wrappers
// around the C++ host layer.
#include "svo_host_c.h"

#include <cstring>
#include <exception>
#include <memory>
#include <string>
#include <vector>

#include "stereo_vo/batch_front_end.hpp"
#include "stereo_vo/camera_info.hpp"
#include "stereo_vo/image_msg.hpp"
#include "stereo_vo/image_processor.hpp"
#include "stereo_vo/kitti_source.hpp"

namespace {
thread_local std::string g_create_error;

template <typename F>
int guarded(std::string& err, F&& f) {
    try {
        f();
        return SVO_OK;
    } catch (const std::invalid_argument& ex) {
        err = ex.what();
        return SVO_ERR_INVALID;
    } catch (const std::logic_error& ex) {
        err = ex.what();
        return SVO_ERR_STATE;
    } catch (const std::bad_alloc&) {
        err = "out of host memory";
        return SVO_ERR_NOMEM;
    } catch (const std::exception& ex) {
        err = ex.what();
        return SVO_ERR_CUDA;
    } catch (...) {
        err = "unknown exception";
        return SVO_ERR_CUDA;
    }
}
}  // namespace

struct svo_vo {
    std::shared_ptr<BundleAdjuster> adjuster;
    std::shared_ptr<FeatureTracker> tracker;
    std::unique_ptr<ImageProcessor> processor;
    std::string error;
    // flat copies of the last record (size_t ids widened to uint64_t)
    std::vector<uint64_t> tracked_ids, new_ids, tracker_ids;
    std::vector<int32_t> inliers, valid_idx;
};

struct svo_batch {
    std::unique_ptr<svo::BatchFrontEnd> fe;
    std::string error;
};

struct svo_ba {
    std::unique_ptr<BundleAdjuster> ba;
    std::string error;
};

struct svo_stereo_queue {
    std::shared_ptr<std::queue<StereoPair>> pairs = std::make_shared<std::queue<StereoPair>>();
    std::unique_ptr<handle_images> handler;
    std::string error;
};

struct svo_kitti {
    std::unique_ptr<svo::KittiSource> src;
    std::unique_ptr<svo::KittiPrefetcher> pf;     // destroyed before src
    std::string error;
    ~svo_kitti() { pf.reset(); }
};

extern "C" {

int svo_camera_from_yaml(const char* yaml_text, float* focal, float* cx, float* cy,
                         float* baseline, float* frame_rate) {
    if (!yaml_text) return SVO_ERR_INVALID;
    return guarded(g_create_error, [&] {
        svo::CameraConfig c = svo::parse_camera_yaml(yaml_text);
        if (focal) *focal = c.focal_length;
        if (cx) *cx = c.cx;
        if (cy) *cy = c.cy;
        if (baseline) *baseline = c.baseline;
        if (frame_rate) *frame_rate = c.frame_rate;
    });
}

int svo_vo_create(float focal, float cx, float cy, float baseline, float min_feature_distance,
                  float parallax_thresh, int32_t device, svo_vo** out) {
    if (!out) return SVO_ERR_INVALID;
    *out = nullptr;
    return guarded(g_create_error, [&] {
        auto v = std::make_unique<svo_vo>();
        cv::Mat K = cv::Mat::eye(3, 3, CV_32F);                      // src/vo_node.cpp:104-108
        K.at<float>(0, 0) = focal; K.at<float>(0, 2) = cx;
        K.at<float>(1, 1) = focal; K.at<float>(1, 2) = cy;
        CameraInfo info;                                             // :110
        info.focal = focal; info.cx = cx; info.cy = cy; info.baseline = baseline;
        v->adjuster = std::make_shared<BundleAdjuster>(5, info);     // window 5, vo_node.cpp:36
        v->tracker = std::make_shared<FeatureTracker>(device);
        v->processor = std::make_unique<ImageProcessor>(K, v->tracker, v->adjuster, baseline,
                                                        min_feature_distance, parallax_thresh);
        *out = v.release();
    });
}

void svo_vo_destroy(svo_vo* v) { delete v; }

const char* svo_vo_last_error(const svo_vo* v) { return v ? v->error.c_str() : g_create_error.c_str(); }

int svo_vo_set_pnp(svo_vo* v, svo_pnp_fn fn, void* user) {
    if (!v || !fn) return SVO_ERR_INVALID;
    v->processor->set_pnp_solver(
        [fn, user](const std::vector<cv::Point3f>& world, const std::vector<cv::Point2f>& image,
                   const cv::Mat& K, cv::Mat& rvec, cv::Mat& tvec, double rmat[9],
                   std::vector<int>& inliers) {
            const int32_t n = (int32_t)world.size();
            float k[9];
            for (int i = 0; i < 9; ++i) k[i] = K.at<float>(i / 3, i % 3);
            double rv[3], tv[3];
            for (int i = 0; i < 3; ++i) {
                rv[i] = rvec.empty() ? 0.0 : rvec.at<double>(i);
                tv[i] = tvec.empty() ? 0.0 : tvec.at<double>(i);
            }
            std::vector<int32_t> inl((size_t)n + 1);
            int32_t n_inl = 0;
            static_assert(sizeof(cv::Point3f) == 12 && sizeof(cv::Point2f) == 8, "packed points");
            int rc = fn(user, n, reinterpret_cast<const float*>(world.data()),
                        reinterpret_cast<const float*>(image.data()), k, rv, tv, rmat, inl.data(),
                        &n_inl);
            if (rc != 0) return false;
            rvec = cv::Mat::zeros(3, 1, CV_64F);
            tvec = cv::Mat::zeros(3, 1, CV_64F);
            for (int i = 0; i < 3; ++i) { rvec.at<double>(i) = rv[i]; tvec.at<double>(i) = tv[i]; }
            inliers.assign(inl.begin(), inl.begin() + n_inl);
            return true;
        });
    return SVO_OK;
}

int svo_vo_process(svo_vo* v, const uint8_t* left, const uint8_t* right, int32_t width,
                   int32_t height, int32_t pitch, double t) {
    if (!v || !left || !right || width <= 0 || height <= 0 || pitch < width) return SVO_ERR_INVALID;
    return guarded(v->error, [&] {
        // owned copies, like the cv_bridge::toCvCopy images vo_node queues (vo_node.cpp:70-73):
        // the Keyframe keeps a shallow reference to the left image past this call
        cv::Mat l = cv::Mat(height, width, CV_8UC1, const_cast<uint8_t*>(left), (size_t)pitch).clone();
        cv::Mat r = cv::Mat(height, width, CV_8UC1, const_cast<uint8_t*>(right), (size_t)pitch).clone();
        v->processor->process(StereoPair(l, r, t));
    });
}

int svo_vo_last_record(svo_vo* v, svo_vo_record* out) {
    if (!v || !out) return SVO_ERR_INVALID;
    const svo::FrameRecord& r = v->processor->last_record();
    std::memset(out, 0, sizeof *out);
    auto f2 = [](const std::vector<cv::Point2f>& a) { return reinterpret_cast<const svo_float2*>(a.data()); };
    v->tracked_ids.assign(r.tracked_ids.begin(), r.tracked_ids.end());
    v->new_ids.assign(r.new_ids.begin(), r.new_ids.end());
    v->tracker_ids.assign(r.tracker_ids.begin(), r.tracker_ids.end());
    v->inliers.assign(r.inliers.begin(), r.inliers.end());
    v->valid_idx.assign(r.valid_idx.begin(), r.valid_idx.end());
    out->decision = (int32_t)r.decision;
    out->n_detected = (int32_t)r.detected.size();
    out->detected = f2(r.detected);
    out->av_parallax = r.av_parallax;
    out->percent_lost = r.percent_lost;
    out->n_tracked = (int32_t)r.tracked.size();
    out->tracked = f2(r.tracked);
    out->tracked_ids = v->tracked_ids.data();
    out->n_inliers = (int32_t)r.inliers.size();
    out->inliers = v->inliers.data();
    std::memcpy(out->hmat, r.hmat, sizeof out->hmat);
    out->n_new_features = (int32_t)r.new_features.size();
    out->new_features = f2(r.new_features);
    out->disp16 = r.disp16.data();
    out->n_new = (int32_t)r.new_2d.size();
    out->valid_idx = v->valid_idx.data();
    out->new_2d = f2(r.new_2d);
    out->new_3d = reinterpret_cast<const svo_float3*>(r.new_3d.data());
    out->new_ids = v->new_ids.data();
    out->n_tracker = (int32_t)r.tracker_pts.size();
    out->tracker_pts = f2(r.tracker_pts);
    out->tracker_ids = v->tracker_ids.data();
    return SVO_OK;
}

int svo_vo_get_drawing(svo_vo* v, uint8_t* out_rgb, int32_t out_pitch, int32_t* width, int32_t* height) {
    if (!v) return SVO_ERR_INVALID;
    return guarded(v->error, [&] {
        cv::Mat d = v->tracker->get_drawing();
        if (d.empty()) throw std::logic_error("svo_vo_get_drawing: no frame processed yet");
        if (width) *width = d.cols;
        if (height) *height = d.rows;
        if (out_rgb) {
            if (out_pitch < 3 * d.cols) throw std::invalid_argument("svo_vo_get_drawing: out_pitch < 3 * width");
            for (int r = 0; r < d.rows; ++r) std::memcpy(out_rgb + (size_t)r * out_pitch, d.ptr(r), (size_t)3 * d.cols);
        }
    });
}

// ------------------------------------------------------------------------------- batch
int svo_batch_create(const svo_params* p, int32_t device, int32_t n_seq, int32_t n_groups,
                     float focal, float cx, float cy, float baseline, float parallax_thresh,
                     svo_batch** out) {
    if (!p || !out) return SVO_ERR_INVALID;
    *out = nullptr;
    return guarded(g_create_error, [&] {
        svo::BatchOptions o;
        o.params = *p;
        o.device = device;
        o.n_seq = n_seq;
        o.n_groups = n_groups;
        o.focal = focal; o.cx = cx; o.cy = cy; o.baseline = baseline;
        o.parallax_thresh = parallax_thresh;
        auto b = std::make_unique<svo_batch>();
        b->fe = std::make_unique<svo::BatchFrontEnd>(o);
        *out = b.release();
    });
}

void svo_batch_destroy(svo_batch* b) { delete b; }

const char* svo_batch_last_error(const svo_batch* b) { return b ? b->error.c_str() : g_create_error.c_str(); }

int svo_batch_step(svo_batch* b, const uint8_t* const* left, const uint8_t* const* right,
                   int32_t pitch, int32_t on_device, const float* poses,
                   const uint8_t* const* next_left, const uint8_t* const* next_right,
                   svo_batch_summary* summary) {
    if (!b || !left) return SVO_ERR_INVALID;
    return guarded(b->error, [&] {
        svo::BatchSummary s = b->fe->step(left, right, pitch, on_device != 0, poses, next_left, next_right);
        if (summary) {
            summary->frames = s.frames; summary->keyframes = s.keyframes;
            summary->corners = s.corners; summary->tracked = s.tracked;
            summary->new_points = s.new_points;
            summary->right_images = s.right_images;
        }
    });
}

int svo_batch_submit(svo_batch* b, const uint8_t* const* left, const uint8_t* const* right,
                     int32_t pitch, int32_t on_device, const float* poses,
                     const uint8_t* const* next_left, const uint8_t* const* next_right) {
    if (!b || !left) return SVO_ERR_INVALID;
    return guarded(b->error, [&] {
        b->fe->submit(left, right, pitch, on_device != 0, poses, next_left, next_right);
    });
}

int svo_batch_wait(svo_batch* b, svo_batch_summary* summary) {
    if (!b) return SVO_ERR_INVALID;
    return guarded(b->error, [&] {
        svo::BatchSummary s = b->fe->wait();
        if (summary) {
            summary->frames = s.frames; summary->keyframes = s.keyframes;
            summary->corners = s.corners; summary->tracked = s.tracked;
            summary->new_points = s.new_points;
            summary->right_images = s.right_images;
        }
    });
}

int svo_batch_get(svo_batch* b, int32_t seq, int32_t* decision, int32_t* n_corners,
                  svo_float2* corners, svo_track_stats* stats, int32_t* n_tracker,
                  svo_float2* tracker_pts, uint64_t* tracker_ids, int32_t* n_new,
                  svo_float2* new_2d, svo_float3* new_3d) {
    if (!b || seq < 0 || seq >= b->fe->n_seq()) return SVO_ERR_INVALID;
    const svo::BatchFrontEnd& fe = *b->fe;
    if (decision) *decision = fe.decision(seq);
    if (n_corners) *n_corners = fe.n_corners(seq);
    if (corners) std::memcpy(corners, fe.corners(seq), (size_t)fe.n_corners(seq) * sizeof(svo_float2));
    if (stats) *stats = fe.stats(seq);
    if (n_tracker) *n_tracker = fe.tracker_count(seq);
    if (tracker_pts) std::memcpy(tracker_pts, fe.tracker_points(seq), (size_t)fe.tracker_count(seq) * sizeof(svo_float2));
    if (tracker_ids) std::memcpy(tracker_ids, fe.tracker_ids(seq), (size_t)fe.tracker_count(seq) * sizeof(uint64_t));
    if (n_new) *n_new = fe.n_new(seq);
    if (new_2d) std::memcpy(new_2d, fe.new_2d(seq), (size_t)fe.n_new(seq) * sizeof(svo_float2));
    if (new_3d) std::memcpy(new_3d, fe.new_3d(seq), (size_t)fe.n_new(seq) * sizeof(svo_float3));
    return SVO_OK;
}

int svo_batch_get_keyframe(svo_batch* b, int32_t seq, int32_t* n_new_features,
                           svo_float2* new_features, int16_t* disp16, int32_t* n_valid,
                           int32_t* valid_idx) {
    if (!b || seq < 0 || seq >= b->fe->n_seq()) return SVO_ERR_INVALID;
    const svo::BatchFrontEnd& fe = *b->fe;
    const size_t nf = (size_t)fe.n_new_features(seq), nv = (size_t)fe.n_valid(seq);
    if (n_new_features) *n_new_features = (int32_t)nf;
    if (n_valid) *n_valid = (int32_t)nv;
    if (new_features) std::memcpy(new_features, fe.new_features(seq), nf * sizeof(svo_float2));
    if (disp16) std::memcpy(disp16, fe.disp16(seq), nf * sizeof(int16_t));
    if (valid_idx) std::memcpy(valid_idx, fe.valid_idx(seq), nv * sizeof(int32_t));
    return SVO_OK;
}

int svo_batch_set_closed_loop(svo_batch* b, int32_t on, int32_t iterations, float reproj_error,
                              double confidence) {
    if (!b) return SVO_ERR_INVALID;
    return guarded(b->error, [&] { b->fe->set_closed_loop(on != 0, iterations, reproj_error, confidence); });
}

int svo_batch_get_pose(svo_batch* b, int32_t seq, float hmat[16], int32_t* n_inliers) {
    if (!b || seq < 0 || seq >= b->fe->n_seq()) return SVO_ERR_INVALID;
    if (hmat) std::memcpy(hmat, b->fe->pose(seq), 16 * sizeof(float));
    if (n_inliers) *n_inliers = b->fe->n_inliers(seq);
    return SVO_OK;
}

int svo_batch_set_blocking_sync(svo_batch* b, int32_t on) {
    if (!b) return SVO_ERR_INVALID;
    b->fe->set_blocking_sync(on != 0);
    return SVO_OK;
}

int64_t svo_batch_kernel_launches(const svo_batch* b) { return b ? b->fe->kernel_launches() : 0; }

int svo_batch_timer_begin(svo_batch* b) {
    if (!b) return SVO_ERR_INVALID;
    b->fe->timer_begin();
    return SVO_OK;
}

int svo_batch_timer_end(svo_batch* b, float* ms) {
    if (!b || !ms) return SVO_ERR_INVALID;
    *ms = b->fe->timer_end();
    return SVO_OK;
}

int svo_batch_set_profiling(svo_batch* b, int32_t on) {
    if (!b) return SVO_ERR_INVALID;
    b->fe->set_profiling(on != 0);
    return SVO_OK;
}

int svo_batch_get_profile(const svo_batch* b, double* ms, int64_t* counts) {
    if (!b || !ms || !counts) return SVO_ERR_INVALID;
    std::vector<double> m;
    std::vector<int64_t> c;
    b->fe->get_profile(m, c);
    std::memcpy(ms, m.data(), m.size() * sizeof(double));
    std::memcpy(counts, c.data(), c.size() * sizeof(int64_t));
    return SVO_OK;
}

svo_engine* svo_batch_engine(svo_batch* b, int32_t group) {
    if (!b || group < 0 || group >= b->fe->n_groups()) return nullptr;
    return b->fe->engine(group);
}

// ------------------------------------------------------------------------------- KITTI source
int svo_kitti_open(const char* data_path, const char* sequence, svo_kitti** out) {
    if (!data_path || !sequence || !out) return SVO_ERR_INVALID;
    *out = nullptr;
    return guarded(g_create_error, [&] {
        auto k = std::make_unique<svo_kitti>();
        k->src = std::make_unique<svo::KittiSource>(data_path, sequence);
        *out = k.release();
    });
}

void svo_kitti_close(svo_kitti* k) { delete k; }

const char* svo_kitti_last_error(const svo_kitti* k) { return k ? k->error.c_str() : g_create_error.c_str(); }

int svo_kitti_info(const svo_kitti* k, int64_t* n_frames, int32_t* width, int32_t* height) {
    if (!k) return SVO_ERR_INVALID;
    if (n_frames) *n_frames = (int64_t)k->src->size();
    if (width) *width = k->src->width();
    if (height) *height = k->src->height();
    return SVO_OK;
}

static void copy_pose(const svo::KittiPose& p, double* pose12) {
    if (!pose12) return;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) pose12[4 * r + c] = p.R[3 * r + c];
        pose12[4 * r + 3] = p.t[r];
    }
}

int svo_kitti_read(svo_kitti* k, int64_t i, uint8_t* left, uint8_t* right, int32_t pitch,
                   double* pose12) {
    if (!k || !left || !right || i < 0) return SVO_ERR_INVALID;
    return guarded(k->error, [&] {
        k->src->read((size_t)i, left, right, pitch);
        copy_pose(k->src->pose((size_t)i), pose12);
    });
}

int svo_kitti_prefetch_start_threads(svo_kitti* k, uint8_t* slab, int32_t depth, int32_t pitch,
                                     int64_t first, int32_t threads) {
    if (!k || first < 0) return SVO_ERR_INVALID;
    return guarded(k->error, [&] {
        k->pf.reset();
        k->pf = std::make_unique<svo::KittiPrefetcher>(*k->src, slab, depth, pitch, (size_t)first, threads);
    });
}

int svo_kitti_prefetch_start(svo_kitti* k, uint8_t* slab, int32_t depth, int32_t pitch, int64_t first) {
    return svo_kitti_prefetch_start_threads(k, slab, depth, pitch, first, 1);
}

int svo_kitti_next(svo_kitti* k, int64_t* index, const uint8_t** left, const uint8_t** right,
                   double* pose12) {
    if (!k || !k->pf || !index || !left || !right) return SVO_ERR_INVALID;
    int got = 0;
    const int rc = guarded(k->error, [&] {
        size_t i = 0;
        if (k->pf->next(i, *left, *right)) {
            *index = (int64_t)i;
            copy_pose(k->src->pose(i), pose12);
            got = 1;
        }
    });
    return rc != SVO_OK ? rc : got;
}

// ------------------------------------------------------------- BundleAdjuster bookkeeping alone
int svo_ba_create(int32_t window_size, svo_ba** out) {
    if (!out || window_size <= 0) return SVO_ERR_INVALID;
    *out = nullptr;
    return guarded(g_create_error, [&] {
        auto b = std::make_unique<svo_ba>();
        CameraInfo info;
        b->ba = std::make_unique<BundleAdjuster>((size_t)window_size, info);
        *out = b.release();
    });
}

void svo_ba_destroy(svo_ba* b) { delete b; }

const char* svo_ba_last_error(const svo_ba* b) { return b ? b->error.c_str() : g_create_error.c_str(); }

int svo_ba_add_keyframe(svo_ba* b, const uint64_t* tracked_ids, int32_t n_tracked,
                        const svo_float3* new_3d, int32_t n_new, uint64_t* new_ids, int32_t* n_kept) {
    if (!b || n_tracked < 0 || n_new < 0 || (n_tracked > 0 && !tracked_ids) ||
        (n_new > 0 && (!new_3d || !new_ids)))
        return SVO_ERR_INVALID;
    if (n_kept) *n_kept = 0;
    return guarded(b->error, [&] {
        std::vector<size_t> tids(tracked_ids, tracked_ids + n_tracked);
        for (size_t id : tids)
            if (id >= b->ba->num_features()) throw std::invalid_argument("svo_ba_add_keyframe: unknown tracked id");
        std::vector<cv::Point2f> t2d((size_t)n_tracked), n2d((size_t)n_new);
        std::vector<cv::Point3f> n3d;
        for (int32_t i = 0; i < n_new; ++i) n3d.push_back(cv::Point3f(new_3d[i].x, new_3d[i].y, new_3d[i].z));
        auto kf = std::make_shared<Keyframe>(svo::Vec3f(), svo::Quatf(), cv::Mat(), t2d, tids, n2d, n3d);
        b->ba->add_keyframe(kf);
        for (size_t i = 0; i < kf->new_ids.size(); ++i) new_ids[i] = (uint64_t)kf->new_ids[i];
        if (n_kept) *n_kept = (int32_t)kf->new_ids.size();
    });
}

int svo_ba_get_world_points(svo_ba* b, const uint64_t* ids, int32_t n, svo_float3* out) {
    if (!b || n < 0 || (n > 0 && (!ids || !out))) return SVO_ERR_INVALID;
    return guarded(b->error, [&] {
        std::vector<size_t> v(ids, ids + n);
        for (size_t id : v)
            if (id >= b->ba->num_features()) throw std::invalid_argument("svo_ba_get_world_points: unknown id");
        std::vector<cv::Point3f> pts;
        b->ba->get_world_points(pts, v);
        for (size_t i = 0; i < pts.size(); ++i) out[i] = svo_float3{pts[i].x, pts[i].y, pts[i].z};
    });
}

int64_t svo_ba_num_features(const svo_ba* b) { return b ? (int64_t)b->ba->num_features() : (int64_t)SVO_ERR_INVALID; }

// ------------------------------------------------- image messages, handle_images, drain loop
namespace {
svo::ImageMsg msg_of(const svo_image_msg* m) {
    if (!m || !m->encoding) throw std::invalid_argument("image message or its encoding is null");
    svo::ImageMsg o;
    o.stamp = m->stamp;
    o.height = m->height;
    o.width = m->width;
    o.encoding = m->encoding;
    o.is_bigendian = m->is_bigendian;
    o.step = m->step;
    o.data = m->data;
    o.size = m->size;
    return o;
}
}  // namespace

int svo_image_to_mono8(const svo_image_msg* msg, uint8_t* out, int32_t pitch) {
    return guarded(g_create_error, [&] {
        if (pitch <= 0) throw std::invalid_argument("svo_image_to_mono8: pitch must be positive");
        svo::to_mono8(msg_of(msg), out, (size_t)pitch);
    });
}

const char* svo_image_last_error(void) { return g_create_error.c_str(); }

int svo_stereo_queue_create(double drop_time, svo_stereo_queue** out) {
    if (!out) return SVO_ERR_INVALID;
    *out = nullptr;
    return guarded(g_create_error, [&] {
        auto q = std::make_unique<svo_stereo_queue>();
        q->handler = std::make_unique<handle_images>(q->pairs, drop_time);
        *out = q.release();
    });
}

void svo_stereo_queue_destroy(svo_stereo_queue* q) { delete q; }

const char* svo_stereo_queue_last_error(const svo_stereo_queue* q) {
    return q ? q->error.c_str() : g_create_error.c_str();
}

int svo_stereo_queue_push(svo_stereo_queue* q, const svo_image_msg* left, const svo_image_msg* right,
                          int32_t* queued) {
    if (!q) return SVO_ERR_INVALID;
    if (queued) *queued = 0;
    return guarded(q->error, [&] {
        const bool took = (*q->handler)(msg_of(left), msg_of(right));
        if (queued) *queued = took ? 1 : 0;
    });
}

int32_t svo_stereo_queue_size(const svo_stereo_queue* q) { return q ? (int32_t)q->pairs->size() : SVO_ERR_INVALID; }

int svo_stereo_queue_drain(svo_stereo_queue* q, svo_stereo_pair_fn fn, void* user, int32_t* n_processed) {
    if (!q || !fn) return SVO_ERR_INVALID;
    if (n_processed) *n_processed = 0;
    return guarded(q->error, [&] {
        int32_t done = 0;
        try {
            svo::drain_queue(*q->pairs, [&](const StereoPair& p) {
                const int rc = fn(user, p.left.data, p.right.data, p.left.cols, p.left.rows,
                                  (int32_t)p.left.step, (int32_t)p.right.step, p.t);
                if (rc != 0) throw std::logic_error("svo_stereo_queue_drain: the consumer returned " + std::to_string(rc));
                ++done;
            });
        } catch (...) {
            if (n_processed) *n_processed = done;
            throw;
        }
        if (n_processed) *n_processed = done;
    });
}

int svo_vo_process_queue(svo_vo* v, svo_stereo_queue* q, int32_t* n_processed) {
    if (!v || !q) return SVO_ERR_INVALID;
    if (n_processed) *n_processed = 0;
    return guarded(v->error, [&] {
        int32_t done = 0;
        try {
            svo::drain_queue(*q->pairs, [&](const StereoPair& p) {       // src/vo_node.cpp:140-143
                v->processor->process(p);
                ++done;
            });
        } catch (...) {
            if (n_processed) *n_processed = done;
            throw;
        }
        if (n_processed) *n_processed = done;
    });
}

}  // extern "C"
