// TEST INFRASTRUCTURE ONLY (oracle/): minimal stand-in for OpenCV's cv::Mat.
//
// OpenCV core is an un-vendored dependency of the reference (README.md:82,
// CMakeLists.txt find_package(OpenCV)) and its C++ headers are not installed in
// this image.  The reference's hot path uses cv::Mat purely as storage
// (ctor(rows, cols, CV_64F), .data, .rows, .cols, .type(),
// .at<double>(r,c), .row(i)); no OpenCV arithmetic is involved.  This header
// provides that storage so the reference headers compile UNMODIFIED.
#pragma once
// standard headers the real OpenCV core pulls in transitively and the reference relies on
#include <chrono>
#include <memory>
#include <string>
#include <cfloat>
#include <climits>
#include <functional>
#include <mutex>
#include <iostream>
#include <vector>
#include <cstring>
#include <cstddef>

#ifndef CV_64F
#define CV_64F 6
#endif

// opencv2/core/cvdef.h defines these; GCRANSAC.h:287 uses MIN
#ifndef MIN
#define MIN(a, b) ((a) > (b) ? (b) : (a))
#endif
#ifndef MAX
#define MAX(a, b) ((a) < (b) ? (b) : (a))
#endif

typedef unsigned char uchar;

namespace cv
{
class Mat
{
public:
    int rows = 0;
    int cols = 0;
    uchar* data = nullptr;

    Mat() {}
    Mat(int rows_, int cols_, int type_) : rows(rows_), cols(cols_), type_id(type_)
    {
        const size_t n = static_cast<size_t>(rows_) * static_cast<size_t>(cols_) * sizeof(double);
        store = std::make_shared<std::vector<uchar>>(n ? n : 1, uchar(0));
        data = store->data();
    }
    // size_t overload: the reference passes size_t row counts (gcransac_python.cpp:285)
    Mat(size_t rows_, size_t cols_, int type_) : Mat(static_cast<int>(rows_), static_cast<int>(cols_), type_) {}
    Mat(size_t rows_, int cols_, int type_) : Mat(static_cast<int>(rows_), cols_, type_) {}
    Mat(int rows_, size_t cols_, int type_) : Mat(rows_, static_cast<int>(cols_), type_) {}

    int type() const { return type_id; }

    template <typename T> T& at(size_t r, size_t c)
    {
        return reinterpret_cast<T*>(data)[r * static_cast<size_t>(cols) + c];
    }
    template <typename T> const T& at(size_t r, size_t c) const
    {
        return reinterpret_cast<const T*>(data)[r * static_cast<size_t>(cols) + c];
    }
    // header onto one row, sharing the buffer (as cv::Mat::row does)
    Mat row(size_t i) const
    {
        Mat m;
        m.rows = 1;
        m.cols = cols;
        m.type_id = type_id;
        m.store = store;
        m.data = data + i * static_cast<size_t>(cols) * sizeof(double);
        return m;
    }

private:
    int type_id = CV_64F;
    std::shared_ptr<std::vector<uchar>> store;
};
} // namespace cv
