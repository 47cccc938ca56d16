// cv_compat.h -- the handful of OpenCV types the drop-in classes need.
//
// In an ORB-SLAM2 tree OpenCV is present and this header only includes it.  Where OpenCV's C++
// headers are absent (this build container), layout-identical stand-ins are defined so the same
// ORBextractor.cc / ORBmatcher.cc compile and run: cv::KeyPoint is the 28-byte POD of
// include/orbgpu.h, cv::Mat is a ref-counted 2-D byte matrix with ROI support, and
// InputArray / OutputArray are the usual thin proxies.
#pragma once

#if defined(ORBGPU_USE_OPENCV) || (defined(__has_include) && __has_include(<opencv2/core/core.hpp>) && !defined(ORBGPU_NO_OPENCV))
#include <opencv2/core/core.hpp>
#define ORBGPU_HAVE_OPENCV 1
#else
#define ORBGPU_HAVE_OPENCV 0
#include <cstddef>
#include <cstring>
#include <memory>
#include <vector>

#ifndef CV_8U
#define CV_8U 0
#define CV_8UC1 0
#define CV_8UC3 16
#define CV_8UC4 24
#endif

namespace cv {

template <typename T> struct Point_ { T x, y; Point_() : x(0), y(0) {} Point_(T a, T b) : x(a), y(b) {} };
typedef Point_<int> Point;
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Rect { int x, y, width, height; Rect() : x(0), y(0), width(0), height(0) {} Rect(int a, int b, int w, int h) : x(a), y(b), width(w), height(h) {} };

struct KeyPoint {                 // == cv::KeyPoint: 5 floats + 2 ints, 28 bytes
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : size(0), angle(-1), response(0), octave(0), class_id(-1) {}
};

class Mat {
public:
    int rows, cols;
    unsigned char* data;
    size_t step;                  // bytes between rows
    Mat() : rows(0), cols(0), data(0), step(0), type_(CV_8UC1) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(0), step(0), type_(type) { create(r, c, type); }
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0)   // wraps caller memory, no ownership
        : rows(r), cols(c), data((unsigned char*)ext), step(ext_step ? ext_step : (size_t)c * ((type >> 3) + 1)), type_(type) {}
    void create(int r, int c, int type)
    {
        const size_t rb = (size_t)c * ((type >> 3) + 1);
        if (r == rows && c == cols && type == type_ && buf_ && step == rb) return;
        rows = r; cols = c; type_ = type; step = rb;
        buf_.reset(new std::vector<unsigned char>((size_t)r * rb + 1));
        data = buf_->data();
    }
    void release() { buf_.reset(); data = 0; rows = cols = 0; step = 0; }
    bool empty() const { return data == 0 || rows == 0 || cols == 0; }
    int type() const { return type_; }
    int channels() const { return (type_ >> 3) + 1; }
    bool isContinuous() const { return step == (size_t)cols * channels(); }
    unsigned char* ptr(int r = 0) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step); }
    template <typename T> T& at(int r, int c) { return ((T*)(data + (size_t)r * step))[c]; }
    Mat row(int r) const { Mat m(*this); m.data = data + (size_t)r * step; m.rows = 1; return m; }
    Mat operator()(const Rect& q) const
    {
        Mat m(*this);
        m.data = data + (size_t)q.y * step + (size_t)q.x * channels(); m.rows = q.height; m.cols = q.width;
        return m;
    }
    Mat clone() const
    {
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; ++r) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * channels());
        return m;
    }
private:
    int type_;
    std::shared_ptr<std::vector<unsigned char> > buf_;
};

class _InputArray {
public:
    _InputArray() : m_(0) {}
    _InputArray(const Mat& m) : m_(&m) {}
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
private:
    const Mat* m_;
};
class _OutputArray {
public:
    _OutputArray(Mat& m) : m_(&m) {}
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { m_->release(); }
    Mat getMat() const { return *m_; }
private:
    Mat* m_;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

}  // namespace cv
#endif
