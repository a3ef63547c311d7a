// FeatureTracker facade over libsvo.so. The tracker state of the reference (last_image,
// feature_set, feature_ids, initial_features: src/feature_tracker.hpp:57-62) lives on the
// device inside a single-sequence svo_engine; this class mirrors the small host-visible part
// (feature set, ids, initial positions) and forwards every pixel operation:
//   init            -> svo_tracker_init     (src/feature_tracker.cpp:3-16)
//   track_features  -> svo_tracker_track    (src/feature_tracker.cpp:18-67)
// Errors of the C ABI surface as std::runtime_error, the way cv::Exception would in the
// reference. There is no CPU implementation of LK in here.
#include "stereo_vo/feature_tracker.hpp"

#include <cmath>
#include <cstdlib>
#include <stdexcept>
#include <string>

#include "svo_c.h"

namespace svo {

EngineHandle::~EngineHandle() {
    if (e) svo_destroy(e);
}

void check(svo_engine* e, int rc, const char* what) {
    if (rc == SVO_OK) return;
    const char* msg = svo_last_error(e);
    throw std::runtime_error(std::string(what) + " failed (" + std::to_string(rc) + "): " +
                             (msg ? msg : ""));
}

// the reference hands CV_8UC1 images around; anything else is a programming error there too
const uint8_t* image_bytes(const cv::Mat& m, int w, int h) {
    if (m.empty() || m.type() != CV_8UC1 || m.cols != w || m.rows != h)
        throw std::invalid_argument("stereo_vo: expected a " + std::to_string(w) + "x" +
                                    std::to_string(h) + " CV_8UC1 image");
    return m.ptr(0);
}

}  // namespace svo

std::shared_ptr<svo::EngineHandle> FeatureTracker::engine(const cv::Mat& image) {
    if (image.empty() || image.type() != CV_8UC1)
        throw std::invalid_argument("FeatureTracker: expected a CV_8UC1 image");
    if (eng_ && (eng_->width != image.cols || eng_->height != image.rows))
        throw std::invalid_argument("FeatureTracker: the image size changed mid-sequence");
    if (!eng_) {
        svo_params p;
        svo_default_params(&p, image.cols, image.rows);
        auto h = std::make_shared<svo::EngineHandle>();
        int rc = svo_create(&p, device_, 1, &h->e);
        if (rc != SVO_OK) {
            const char* msg = svo_last_error(nullptr);
            throw std::runtime_error(std::string("svo_create failed: ") + (msg ? msg : ""));
        }
        h->width = image.cols;
        h->height = image.rows;
        h->device = device_;
        eng_ = h;
    }
    return eng_;
}

void FeatureTracker::adopt_init(const cv::Mat& image, const std::vector<cv::Point2f>& features,
                                const std::vector<size_t>& ids) {
    feature_ids = ids;
    feature_set = features;
    initial_features.clear();
    for (size_t i = 0; i < ids.size(); ++i)
        initial_features.insert(std::make_pair(ids[i], features[i]));   // first value wins
    last_image = image.clone();
    initial_image = last_image;
}

void FeatureTracker::init(const cv::Mat& image, const std::vector<cv::Point2f>& features,
                          const std::vector<size_t>& ids) {
    if (ids.size() != features.size())
        throw std::invalid_argument("FeatureTracker::init: features and ids differ in length");
    auto h = engine(image);
    static_assert(sizeof(cv::Point2f) == sizeof(svo_float2), "Point2f must be two floats");
    std::vector<uint64_t> ids64(ids.begin(), ids.end());
    svo::check(h->e,
               svo_tracker_init(h->e, 0, svo::image_bytes(image, h->width, h->height),
                                (int)image.step,
                                reinterpret_cast<const svo_float2*>(features.data()),
                                ids64.data(), (int)features.size()),
               "svo_tracker_init");
    adopt_init(image, features, ids);
}

void FeatureTracker::adopt_tracked(const cv::Mat& image, const cv::Point2f* pts, const size_t* ids,
                                   int n) {
    feature_set.assign(pts, pts + n);
    feature_ids.assign(ids, ids + n);
    last_image = image.clone();
}

void FeatureTracker::track_features(float& av_parallax, float& percent_lost, const cv::Mat& image,
                                    bool flow_back) {
    auto h = engine(image);
    svo::check(h->e,
               svo_tracker_track(h->e, 0, svo::image_bytes(image, h->width, h->height),
                                 (int)image.step, flow_back ? 1 : 0, &av_parallax, &percent_lost),
               "svo_tracker_track");
    svo_params p;
    svo_get_params(h->e, &p);
    std::vector<svo_float2> pts((size_t)p.max_features);
    std::vector<uint64_t> ids((size_t)p.max_features);
    int n = 0;
    svo::check(h->e, svo_tracker_get(h->e, 0, pts.data(), ids.data(), &n), "svo_tracker_get");
    feature_set.resize((size_t)n);
    feature_ids.resize((size_t)n);
    for (int i = 0; i < n; ++i) {
        feature_set[i] = cv::Point2f(pts[i].x, pts[i].y);
        feature_ids[i] = (size_t)ids[i];
    }
    last_image = image.clone();
}

void FeatureTracker::get_tracked_features(std::vector<cv::Point2f>& features,
                                          std::vector<size_t>& ids) {
    features = feature_set;
    ids = feature_ids;
}

// ---- debug overlay (src/feature_tracker.cpp:74-91): initial_image as RGB with a green
// cv::arrowedLine (thickness 4) from every feature's initial position to its current one,
// rasterised on the GPU (svo_draw_track)
static cv::Mat render_overlay(svo_engine* e, const cv::Mat& gray, const std::vector<svo_float2>& from,
                              const std::vector<svo_float2>& to) {
    cv::Mat rgb(gray.rows, gray.cols, CV_8UC3);
    svo::check(e, svo_draw_track(e, gray.ptr(0), (int32_t)gray.step, from.data(), to.data(),
                                 (int32_t)from.size(), rgb.ptr(0), (int32_t)rgb.step), "svo_draw_track");
    return rgb;
}

void FeatureTracker::draw_track() {
    if (initial_image.empty()) return;
    std::vector<svo_float2> from(feature_set.size()), to(feature_set.size());
    for (size_t i = 0; i < feature_set.size(); ++i) {
        const cv::Point2f a = initial_features.at(feature_ids[i]);
        from[i] = {a.x, a.y};
        to[i] = {feature_set[i].x, feature_set[i].y};
    }
    track_drawing = render_overlay(engine(initial_image)->e, initial_image, from, to);
}

cv::Mat FeatureTracker::get_drawing() {
    if (track_drawing.empty() && !initial_image.empty())      // :86-89: the plain RGB copy
        track_drawing = render_overlay(engine(initial_image)->e, initial_image, {}, {});
    return track_drawing;
}
