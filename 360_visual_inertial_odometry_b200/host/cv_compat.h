// cv_compat.h -- the OpenCV vocabulary the reference's front-end interface speaks
// (cv::Mat, cv::Point2f, uchar).  With OpenCV installed the real headers are used; on
// machines without them (this build container) a minimal look-alike with the same member
// names keeps the shim compilable and testable.  Only what crosses the interface exists.
#pragma once

#if defined(VIO360_WITH_OPENCV) || __has_include(<opencv2/core.hpp>)
#include <opencv2/core.hpp>
#else
#include <cstddef>
#include <cstdint>
typedef unsigned char uchar;
#ifndef CV_8UC1
#define CV_8UC1 0
#endif
namespace cv {
struct Point2f {
    float x, y;
    Point2f() : x(0), y(0) {}
    Point2f(float x_, float y_) : x(x_), y(y_) {}
};
// non-owning view with cv::Mat's field names (rows, cols, data, step)
struct Mat {
    int rows = 0, cols = 0;
    uchar* data = nullptr;
    size_t step = 0;
    Mat() {}
    Mat(int rows_, int cols_, int /*type*/, void* data_, size_t step_ = 0)
        : rows(rows_), cols(cols_), data(static_cast<uchar*>(data_)), step(step_ ? step_ : (size_t)cols_) {}
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return CV_8UC1; }
};
}  // namespace cv
#endif
