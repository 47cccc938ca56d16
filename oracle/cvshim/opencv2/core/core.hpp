// oracle/cvshim/opencv2/core/core.hpp -- TEST INFRASTRUCTURE, not product code.
//
// A minimal stand-in for the OpenCV C++ API surface that /root/reference/src/ORBextractor.cc
// uses, so that file can be compiled UNCHANGED (read from the reference mount, never copied)
// into oracle/_ref/liborbref.so.  OpenCV C++ headers/libs are not installed in this image.
// The pixel arithmetic behind resize / GaussianBlur / FAST / fastAtan2 / copyMakeBorder is in
// oracle/orb_prims.c, pinned byte-for-byte against cv2 4.13.0.
// Everything (types + functions) lives in this one header; the other headers forward here.
#ifndef ORBO_CVSHIM_CORE_HPP
#define ORBO_CVSHIM_CORE_HPP

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <vector>

#define CV_PI 3.1415926535897932384626433832795
#define CV_8U 0
#define CV_8UC1 0

typedef unsigned char uchar;

inline int cvRound(double v) { return (int)std::lrint(v); }
inline int cvRound(float v) { return (int)std::lrintf(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { return (int)std::floor(v); }
inline int cvCeil(double v) { return (int)std::ceil(v); }

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

struct Rect {
    int x, y, width, height;
    Rect() : x(0), y(0), width(0), height(0) {}
    Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {}
};

// cv::KeyPoint: 28-byte POD (pt.x, pt.y, size, angle, response, octave, class_id)
struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0,
             int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

// 8-bit single-channel matrix header over shared, malloc-backed storage (never operator new, so
// the oracle's bump arena for the octree node allocations is not disturbed by pixel buffers).
class Mat {
public:
    struct Step {
        size_t v;
        Step() : v(0) {}
        operator size_t() const { return v; }
    };
    struct ZerosExpr;
    int rows, cols;
    uchar* data;
    Step step;

    Mat() : rows(0), cols(0), data(0), blk(0) {}
    Mat(const Mat& o);
    Mat& operator=(const Mat& o);
    ~Mat();
    Mat(Size sz, int type);
    Mat(int r, int c, int type);
    Mat operator()(const Rect& r) const;
    Mat rowRange(int a, int b) const;
    Mat colRange(int a, int b) const;
    Mat clone() const;
    static ZerosExpr zeros(int r, int c, int type);
    int type() const { return CV_8UC1; }
    bool empty() const { return data == 0 || rows == 0 || cols == 0; }
    size_t step1() const { return step.v; }
    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step.v; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step.v; }
    void create(int r, int c, int type);
    void release();
    // Mat::zeros(...) is an expression in OpenCV: assigning it to a Mat of matching size fills the
    // EXISTING storage (src/ORBextractor.cc:1039 relies on that to write through a rowRange view).
    struct ZerosExpr { int r, c; operator Mat() const; };
    Mat& operator=(const ZerosExpr& e);
private:
    int* blk;   // malloc'ed block: [refcount, pad to 64 B][pixels]; 0 for an empty header
};

class _InputArray {
public:
    _InputArray() : m(0) {}
    _InputArray(const Mat& _m) : m(&_m) {}
    Mat getMat() const { return m ? *m : Mat(); }
    bool empty() const { return !m || m->empty(); }
private:
    const Mat* m;
};
typedef const _InputArray& InputArray;

class _OutputArray {
public:
    _OutputArray(Mat& _m) : m(&_m) {}
    void create(int r, int c, int type) const { m->create(r, c, type); }
    void release() const { m->release(); }
    Mat getMat() const { return *m; }
private:
    Mat* m;
};
typedef const _OutputArray& OutputArray;

inline _InputArray noArray() { return _InputArray(); }

struct KeyPointsFilter {
    static void retainBest(std::vector<KeyPoint>& kps, int n);   // only used by dead code
};

enum { INTER_LINEAR = 1 };
enum { BORDER_REFLECT_101 = 4, BORDER_ISOLATED = 16 };

void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interp = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sx, double sy = 0, int borderType = BORDER_REFLECT_101);
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
float fastAtan2(float y, float x);

}  // namespace cv

#endif
