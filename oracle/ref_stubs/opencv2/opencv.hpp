// TEST INFRASTRUCTURE - stand-in for <opencv2/opencv.hpp> with the three value types the
// reference's bundle_adjuster.hpp names (cv::Point2f, cv::Point3f, cv::Mat as an opaque image
// handle); see oracle/build_ref.py. No image processing.
#pragma once
#include <iostream>
#include <memory>
#include <queue>
#include <utility>
#include <vector>

namespace cv {
struct Point2f {
    float x, y;
    Point2f() : x(0), y(0) {}
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};
struct Point3f {
    float x, y, z;
    Point3f() : x(0), y(0), z(0) {}
    Point3f(float x_, float y_, float z_) : x(x_), y(y_), z(z_) {}
};
struct Mat {};
}  // namespace cv
