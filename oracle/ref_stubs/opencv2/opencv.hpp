// TEST INFRASTRUCTURE - a stand-in for <opencv2/opencv.hpp>, written for this repository, so that
// oracle/build_ref.py can compile the reference's sources WHERE THEY LIE without an OpenCV SDK:
//
//   * the VALUE TYPES the reference's code manipulates itself are implemented here with OpenCV's
//     semantics: cv::Mat (shared ownership, clone(), ROI headers, copyTo() that writes in place
//     only when size and type match, convertTo, t(), unary minus, at<T>()), Point2f / Point3f
//     (+, -, cv::norm in double), Size, Rect, Scalar, TermCriteria, Ptr;
//   * every OpenCV ALGORITHM the reference calls - goodFeaturesToTrack, calcOpticalFlowPyrLK,
//     StereoBM::compute, solvePnPRansac, Rodrigues and the matrix product - is NOT implemented:
//     the call is forwarded, with the arguments the reference passed, through a table of C
//     callbacks (RefCvCallbacks) that the test process fills with functions running the real
//     OpenCV (Python cv2). So the arithmetic is OpenCV's own and the control flow around it -
//     decisions, id plumbing, dedupe, filter, statistics, the reprojection loop - is the
//     reference's compiled code;
//   * drawing (cvtColor, arrowedLine) only keeps sizes right: the overlay is not compared here.
//
// Output types follow the OpenCV 3.x generation the reference was written against:
// solvePnPRansac with useExtrinsicGuess leaves rvec / tvec in the type they came in with
// (CV_32F, image_processor.cpp:54-55) and Rodrigues answers in the type of its input.
#pragma once
#include <math.h>

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <memory>
#include <queue>
#include <stdexcept>
#include <utility>
#include <vector>

#define CV_8U 0
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_GRAY2RGB 8
#define CV_RGB(r, g, b) cv::Scalar((b), (g), (r), 0)

extern "C" {
// filled by the test process (oracle/ref_frontend.py); every function runs the real OpenCV
struct RefCvCallbacks {
    // cv::goodFeaturesToTrack(image, corners, maxCorners, qualityLevel, minDistance) -> count
    int (*gftt)(const unsigned char* img, int rows, int cols, int step, int max_corners,
                double quality, double min_distance, float* out_xy, int cap);
    // cv::calcOpticalFlowPyrLK(prev, next, prevPts, nextPts, status, err, winSize, maxLevel, criteria, flags, minEig)
    void (*lk)(const unsigned char* prev, int prev_step, const unsigned char* next, int next_step,
               int rows, int cols, const float* pts, int n, float* out_pts, unsigned char* status,
               float* err, int win_w, int win_h, int max_level, int crit_type, int crit_count,
               double crit_eps, int flags, double min_eig);
    // cv::StereoBM::create(numDisparities, blockSize)->compute(left, right, disparity CV_16S)
    void (*stereo_bm)(const unsigned char* left, int lstep, const unsigned char* right, int rstep,
                      int rows, int cols, int ndisp, int block, short* out);
    // cv::solvePnPRansac(world, image, K, dist = 0, rvec, tvec, useExtrinsicGuess, iterations,
    // reprojectionError, confidence, inliers): rvec / tvec in and out; returns the inlier count
    int (*pnp)(const float* world_xyz, const float* image_xy, int n, const float* K9, double* rvec3,
               double* tvec3, int depth, int use_guess, int iterations, float reproj_error,
               double confidence, int* inliers, int cap);
    // cv::Rodrigues(rvec (depth), rmat)
    void (*rodrigues)(const double* rvec3, int depth, double* rmat9);
    // cv::gemm(A (m x k), B (k x n), alpha, noArray(), 0, out), all of `depth`
    void (*gemm)(const double* a, const double* b, int depth, int m, int k, int n, double alpha, double* out);
};
const RefCvCallbacks* ref_cv_callbacks();
}

typedef unsigned char uchar;     // OpenCV's cvdef.h puts it in the global namespace

namespace cv {

using ::uchar;

template <typename T>
struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<float> Point2f;
template <typename T> Point_<T> operator+(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x + b.x, a.y + b.y); }
template <typename T> Point_<T> operator-(const Point_<T>& a, const Point_<T>& b) { return Point_<T>(a.x - b.x, a.y - b.y); }
// types.hpp: norm(Point_) = sqrt((double)x * x + (double)y * y)
template <typename T> double norm(const Point_<T>& p) { return std::sqrt((double)p.x * p.x + (double)p.y * p.y); }

template <typename T>
struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
};
typedef Point3_<float> Point3f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};
struct Rect {
    int x, y, width, height;
    Rect(int x_, int y_, int w, int h) : x(x_), y(y_), width(w), height(h) {}
};
struct Scalar {
    double v[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { v[0] = a; v[1] = b; v[2] = c; v[3] = d; }
};
struct TermCriteria {
    enum { COUNT = 1, MAX_ITER = 1, EPS = 2 };
    int type, maxCount;
    double epsilon;
    TermCriteria(int t, int c, double e) : type(t), maxCount(c), epsilon(e) {}
};
template <typename T> using Ptr = std::shared_ptr<T>;

class Mat {
  public:
    int rows, cols;
    size_t step;
    uchar* data;

    Mat() : rows(0), cols(0), step(0), data(nullptr), type_(CV_8UC1) {}
    Mat(int r, int c, int type) : rows(0), cols(0), step(0), data(nullptr), type_(type) { create(r, c, type); }

    static size_t elem(int type) {
        switch (type) {
            case CV_8UC1: return 1;
            case CV_8UC3: return 3;
            case CV_16S: return 2;
            case CV_32S: case CV_32F: return 4;
            case CV_64F: return 8;
            default: throw std::invalid_argument("ref_stubs cv::Mat: unsupported type");
        }
    }
    // Mat::create: keeps the buffer when size and type already match, else allocates a new one
    void create(int r, int c, int type) {
        if (data && rows == r && cols == c && type_ == type) return;
        rows = r; cols = c; type_ = type;
        step = (size_t)c * elem(type);
        owner_.reset(new uchar[step * (size_t)(r > 0 ? r : 0) + 8](), std::default_delete<uchar[]>());
        data = owner_.get();
    }
    static Mat zeros(int r, int c, int type) { return Mat(r, c, type); }   // create() zero-fills
    static Mat eye(int r, int c, int type) {
        Mat m(r, c, type);
        for (int i = 0; i < r && i < c; ++i) m.set(i, i, 1.0);
        return m;
    }
    int type() const { return type_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    uchar* ptr(int r) { return data + (size_t)r * step; }
    const uchar* ptr(int r) const { return data + (size_t)r * step; }
    template <typename T> T& at(int r, int c) { return *reinterpret_cast<T*>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *reinterpret_cast<const T*>(data + (size_t)r * step + (size_t)c * sizeof(T)); }
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : at<T>(i, 0); }

    Mat clone() const {
        Mat m;
        if (empty()) return m;
        m.create(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * elem(type_));
        return m;
    }
    // header on a sub-rectangle of the same pixels
    Mat operator()(const Rect& roi) const {
        Mat m(*this);
        m.rows = roi.height; m.cols = roi.width;
        m.data = data + (size_t)roi.y * step + (size_t)roi.x * elem(type_);
        return m;
    }
    // Mat::copyTo(OutputArray): dst.create(size, type) first - a destination header of the same
    // size and type is written in place (that is how image_processor.cpp:132,134 fill hmat
    // through a ROI), any other destination gets a buffer of its own
    void copyTo(Mat&& dst) const {
        dst.create(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(dst.ptr(r), ptr(r), (size_t)cols * elem(type_));
    }
    void copyTo(Mat& dst) const {
        dst.create(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(dst.ptr(r), ptr(r), (size_t)cols * elem(type_));
    }
    // convertTo(dst, rtype, alpha): saturate_cast<rtype>(src * alpha); only the conversions the
    // reference performs (CV_16S disparity -> CV_32F, exact for alpha = 1/16)
    void convertTo(Mat& dst, int rtype, double alpha = 1.0) const {
        Mat out(rows, cols, rtype);
        if (type_ == CV_16S && rtype == CV_32F) {          // the disparity map: a whole image
            for (int r = 0; r < rows; ++r) {
                const short* s = reinterpret_cast<const short*>(ptr(r));
                float* d = reinterpret_cast<float*>(out.ptr(r));
                for (int c = 0; c < cols; ++c) d[c] = (float)(s[c] * alpha);
            }
        } else {
            for (int r = 0; r < rows; ++r)
                for (int c = 0; c < cols; ++c) out.set(r, c, get(r, c) * alpha);
        }
        dst = out;
    }
    Mat t() const {
        Mat m(cols, rows, type_);
        for (int r = 0; r < rows; ++r) for (int c = 0; c < cols; ++c) m.set(c, r, get(r, c));
        return m;
    }
    Mat operator-() const {
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; ++r) for (int c = 0; c < cols; ++c) m.set(r, c, -get(r, c));
        return m;
    }
    double get(int r, int c) const {
        switch (type_) {
            case CV_8UC1: return at<uchar>(r, c);
            case CV_16S: return at<short>(r, c);
            case CV_32S: return at<int>(r, c);
            case CV_32F: return at<float>(r, c);
            case CV_64F: return at<double>(r, c);
            default: throw std::invalid_argument("ref_stubs cv::Mat: element access on this type");
        }
    }
    void set(int r, int c, double v) {
        switch (type_) {
            case CV_8UC1: at<uchar>(r, c) = (uchar)v; break;
            case CV_16S: at<short>(r, c) = (short)v; break;
            case CV_32S: at<int>(r, c) = (int)v; break;
            case CV_32F: at<float>(r, c) = (float)v; break;
            case CV_64F: at<double>(r, c) = v; break;
            default: throw std::invalid_argument("ref_stubs cv::Mat: element access on this type");
        }
    }

  private:
    int type_;
    std::shared_ptr<uchar> owner_;
};

// A * B: OpenCV evaluates the expression with cv::gemm - forwarded to the real one
inline Mat operator*(const Mat& a, const Mat& b) {
    if (a.cols != b.rows || a.type() != b.type() || (a.type() != CV_32F && a.type() != CV_64F))
        throw std::invalid_argument("ref_stubs cv::Mat: product of incompatible matrices");
    std::vector<double> va((size_t)a.rows * a.cols), vb((size_t)b.rows * b.cols), vo((size_t)a.rows * b.cols);
    for (int r = 0; r < a.rows; ++r) for (int c = 0; c < a.cols; ++c) va[(size_t)r * a.cols + c] = a.get(r, c);
    for (int r = 0; r < b.rows; ++r) for (int c = 0; c < b.cols; ++c) vb[(size_t)r * b.cols + c] = b.get(r, c);
    ref_cv_callbacks()->gemm(va.data(), vb.data(), a.type(), a.rows, a.cols, b.cols, 1.0, vo.data());
    Mat o(a.rows, b.cols, a.type());
    for (int r = 0; r < o.rows; ++r) for (int c = 0; c < o.cols; ++c) o.set(r, c, vo[(size_t)r * o.cols + c]);
    return o;
}

inline void goodFeaturesToTrack(const Mat& image, std::vector<Point2f>& corners, int max_corners,
                                double quality, double min_distance) {
    std::vector<float> xy((size_t)2 * (max_corners > 0 ? max_corners : 1));
    const int n = ref_cv_callbacks()->gftt(image.data, image.rows, image.cols, (int)image.step, max_corners,
                                           quality, min_distance, xy.data(), max_corners);
    corners.clear();
    for (int i = 0; i < n; ++i) corners.push_back(Point2f(xy[2 * (size_t)i], xy[2 * (size_t)i + 1]));
}

inline void calcOpticalFlowPyrLK(const Mat& prev, const Mat& next, const std::vector<Point2f>& prev_pts,
                                 std::vector<Point2f>& next_pts, std::vector<uchar>& status,
                                 std::vector<float>& err, Size win, int max_level, TermCriteria crit,
                                 int flags, double min_eig) {
    static_assert(sizeof(Point2f) == 8, "packed points");
    const int n = (int)prev_pts.size();
    next_pts.assign((size_t)n, Point2f());
    status.assign((size_t)n, 0);
    err.assign((size_t)n, 0.f);
    if (n == 0) return;
    ref_cv_callbacks()->lk(prev.data, (int)prev.step, next.data, (int)next.step, prev.rows, prev.cols,
                           reinterpret_cast<const float*>(prev_pts.data()), n,
                           reinterpret_cast<float*>(next_pts.data()), status.data(), err.data(),
                           win.width, win.height, max_level, crit.type, crit.maxCount, crit.epsilon,
                           flags, min_eig);
}

class StereoBM {
  public:
    static Ptr<StereoBM> create(int ndisp, int block) { return Ptr<StereoBM>(new StereoBM(ndisp, block)); }
    void compute(const Mat& left, const Mat& right, Mat& disparity) {
        disparity.create(left.rows, left.cols, CV_16S);
        ref_cv_callbacks()->stereo_bm(left.data, (int)left.step, right.data, (int)right.step, left.rows,
                                      left.cols, ndisp_, block_, reinterpret_cast<short*>(disparity.data));
    }
  private:
    StereoBM(int n, int b) : ndisp_(n), block_(b) {}
    int ndisp_, block_;
};

inline bool solvePnPRansac(const std::vector<Point3f>& world, const std::vector<Point2f>& image,
                           const Mat& K, const Mat& dist, Mat& rvec, Mat& tvec, bool use_guess,
                           int iterations, float reproj_error, double confidence, Mat& inliers) {
    static_assert(sizeof(Point3f) == 12, "packed points");
    for (int r = 0; r < dist.rows; ++r)
        for (int c = 0; c < dist.cols; ++c)
            if (dist.get(r, c) != 0.0) throw std::invalid_argument("ref_stubs solvePnPRansac: distortion");
    float k9[9];
    for (int i = 0; i < 9; ++i) k9[i] = (float)K.get(i / 3, i % 3);
    // OpenCV 3.x: with an extrinsic guess rvec / tvec keep the type they came in with
    const int depth = (use_guess && !rvec.empty()) ? rvec.type() : CV_64F;
    double rv[3] = {0, 0, 0}, tv[3] = {0, 0, 0};
    if (use_guess && !rvec.empty())
        for (int i = 0; i < 3; ++i) { rv[i] = rvec.get(i, 0); tv[i] = tvec.get(i, 0); }
    const int n = (int)world.size();
    std::vector<int> idx((size_t)(n > 0 ? n : 1));
    const int n_inl = ref_cv_callbacks()->pnp(reinterpret_cast<const float*>(world.data()),
                                              reinterpret_cast<const float*>(image.data()), n, k9, rv, tv,
                                              depth, use_guess ? 1 : 0, iterations, reproj_error,
                                              confidence, idx.data(), n);
    rvec.create(3, 1, depth);
    tvec.create(3, 1, depth);
    for (int i = 0; i < 3; ++i) { rvec.set(i, 0, rv[i]); tvec.set(i, 0, tv[i]); }
    const int m = n_inl > 0 ? n_inl : 0;
    inliers = Mat(m, m > 0 ? 1 : 0, CV_32S);
    for (int i = 0; i < m; ++i) inliers.at<int>(i, 0) = idx[(size_t)i];
    return n_inl > 0;
}

inline void Rodrigues(const Mat& src, Mat& dst) {
    double rv[3], rm[9];
    for (int i = 0; i < 3; ++i) rv[i] = src.get(i, 0);
    ref_cv_callbacks()->rodrigues(rv, src.type(), rm);
    dst = Mat(3, 3, src.type());
    for (int i = 0; i < 9; ++i) dst.set(i / 3, i % 3, rm[i]);
}

// drawing is outside what this build checks: sizes only
inline void cvtColor(const Mat& src, Mat& dst, int) {
    Mat out(src.rows, src.cols, CV_8UC3);
    dst = out;
}
inline void arrowedLine(Mat&, Point2f, Point2f, const Scalar&, int) {}

}  // namespace cv
