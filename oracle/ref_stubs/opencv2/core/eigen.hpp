// TEST INFRASTRUCTURE - stand-in for <opencv2/core/eigen.hpp>: cv::cv2eigen for the two fixed-size
// conversions image_processor.cpp:87-88 performs. See ../opencv.hpp.
#pragma once
#include <Eigen/Dense>

#include "../opencv.hpp"

namespace cv {
template <typename T, int R, int C, int O>
void cv2eigen(const Mat& src, Eigen::Matrix<T, R, C, O>& dst) {
    if (src.rows != R || src.cols != C) throw std::invalid_argument("ref_stubs cv2eigen: size");
    for (int r = 0; r < R; ++r)
        for (int c = 0; c < C; ++c) dst(r, c) = (T)src.get(r, c);
}
}  // namespace cv
