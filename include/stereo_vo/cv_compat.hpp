// cv_compat.hpp - the OpenCV types that appear in the reference's ImageProcessor /
// FeatureTracker / Keyframe signatures (src/image_processor.hpp:9-82,
// src/feature_tracker.hpp:20-62, src/bundle_adjuster.hpp:22-46).
//
// With an OpenCV SDK on the include path the real headers are used and the facade is a
// source-level drop-in for the reference's classes. Without one (this repository's build
// container has no OpenCV C++ SDK) a minimal stand-in with the same spelling is defined: a
// ref-counted row-major cv::Mat restricted to the element types the front end touches
// (CV_8UC1 images, the CV_8UC3 debug overlay, CV_32F / CV_64F small matrices), cv::Point2f,
// cv::Point3f, cv::Size.
// No image processing lives here - every pixel operation is a call into libsvo.so.
#pragma once

#if defined(SVO_USE_OPENCV) || (defined(__has_include) && __has_include(<opencv2/core.hpp>))
#include <opencv2/core.hpp>
#define SVO_HAVE_OPENCV 1
#else
#define SVO_HAVE_OPENCV 0

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>

#ifndef CV_8UC1
#define CV_8U 0
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_32F 5
#define CV_32FC1 5
#define CV_64F 6
#define CV_64FC1 6
#endif

namespace cv {

struct Point2f {
    float x = 0.f, y = 0.f;
    Point2f() = default;
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};
inline Point2f operator-(const Point2f& a, const Point2f& b) { return {a.x - b.x, a.y - b.y}; }
inline Point2f operator+(const Point2f& a, const Point2f& b) { return {a.x + b.x, a.y + b.y}; }

struct Point3f {
    float x = 0.f, y = 0.f, z = 0.f;
    Point3f() = default;
    Point3f(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};

struct Size {
    int width = 0, height = 0;
    Size() = default;
    Size(int w, int h) : width(w), height(h) {}
};

// Dense 2-D matrix with shared ownership (copying a Mat shares the pixels, clone() copies
// them - the two behaviours the reference relies on: vo_node.cpp:70-73, feature_tracker.cpp:14,66).
class Mat {
  public:
    int rows = 0, cols = 0;
    size_t step = 0;          // bytes per row
    unsigned char* data = nullptr;

    Mat() = default;
    Mat(int r, int c, int type) { create(r, c, type); }
    // wrap caller memory (not owned), like cv::Mat(rows, cols, type, void*, step)
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0)
        : rows(r), cols(c), step(ext_step ? ext_step : (size_t)c * elem(type)),
          data(static_cast<unsigned char*>(ext)), type_(type) {}

    void create(int r, int c, int type) {
        rows = r; cols = c; type_ = type;
        step = (size_t)c * elem(type);
        owner_.reset(new unsigned char[step * (size_t)(r > 0 ? r : 0) + 1],
                     std::default_delete<unsigned char[]>());
        data = owner_.get();
    }
    static Mat zeros(int r, int c, int type) {
        Mat m(r, c, type);
        std::memset(m.data, 0, m.step * (size_t)r);
        return m;
    }
    static Mat eye(int r, int c, int type) {
        Mat m = zeros(r, c, type);
        for (int i = 0; i < r && i < c; ++i) {
            if (type == CV_32F) m.at<float>(i, i) = 1.f;
            else if (type == CV_64F) m.at<double>(i, i) = 1.0;
            else m.at<unsigned char>(i, i) = 1;
        }
        return m;
    }
    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(m.ptr(r), ptr(r), m.step);
        return m;
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t elemSize() const { return elem(type_); }
    bool isContinuous() const { return step == (size_t)cols * elem(type_); }
    Size size() const { return Size(cols, rows); }
    unsigned char* ptr(int r = 0) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T& at(int r, int c) {
        return *reinterpret_cast<T*>(data + (size_t)r * step + (size_t)c * sizeof(T));
    }
    template <typename T> const T& at(int r, int c) const {
        return *reinterpret_cast<const T*>(data + (size_t)r * step + (size_t)c * sizeof(T));
    }
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }

  private:
    static size_t elem(int type) {
        switch (type) {
            case CV_8UC1: return 1;
            case CV_8UC3: return 3;
            case CV_32F: return 4;
            case CV_64F: return 8;
            default: throw std::invalid_argument("cv_compat: unsupported Mat type");
        }
    }
    int type_ = CV_8UC1;
    std::shared_ptr<unsigned char> owner_;
};

}  // namespace cv
#endif  // OpenCV or stand-in
