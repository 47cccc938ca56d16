// oracle/cvshim2/opencv2/core/core.hpp -- TEST INFRASTRUCTURE, not product code.
//
// A second, wider stand-in for the OpenCV C++ API: the surface that the reference's src/Frame.cc, src/ORBmatcher.cc
// and src/ORBextractor.cc (plus the DBoW2 headers they pull in) use, so those files compile UNCHANGED, read from the
// reference mount, into oracle/_ref/libframeref.so (recipe: oracle/Makefile).  OpenCV C++ headers are not installed.
//
// What has to be arithmetic-exact, and how it is pinned:
//   * typed Mat (CV_8U / CV_32F / CV_64F, channels), views, clone, convertTo(8U -> 32F), reshape;
//   * the lazy matrix expressions of the tested paths, evaluated the way cv::MatExpr evaluates them:
//       A*B (+C)            -> cv::gemm(A, B, alpha, C, beta, flags)   (small-matrix float case / generic double case)
//       -A.t()*B            -> gemm with GEMM_1_T and alpha = -1  (the generic path: double accumulation)
//       A - s*Mat::ones()   -> element-wise float subtract
//     cv_gemm32f is checked against cv2 4.13.0 golden vectors (tests/golden/gemm_golden.npz) through
//     frameref_gemm3 (oracle/ref_frame_wrap.cc);
//   * norm(A, B, NORM_L1) / norm(A): double accumulation, as cv::norm;
//   * undistortPoints -> oracle/orb_prims.c (pinned to cv2 golden vectors); resize / GaussianBlur / FAST / fastAtan2
//     -> oracle/orb_prims.c as in oracle/cvshim.
// Everything else (FileStorage, inv(), ...) is declared so the translation units parse; paths that would need it are
// never executed by the oracle and abort if they are.
#ifndef ORBO_CVSHIM2_CORE_HPP
#define ORBO_CVSHIM2_CORE_HPP

#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <list>
#include <map>
#include <set>
#include <sstream>
#include <string>
#include <vector>

#define CV_PI 3.1415926535897932384626433832795
#define CV_CN_SHIFT 3
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8U 0
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)

typedef unsigned char uchar;

inline int cvRound(double v) { return (int)std::lrint(v); }
inline int cvRound(float v) { return (int)std::lrintf(v); }
inline int cvRound(int v) { return v; }
inline int cvFloor(double v) { return (int)std::floor(v); }
inline int cvCeil(double v) { return (int)std::ceil(v); }

namespace cv {

typedef std::string String;

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
};
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
template <typename T> struct Point3_ {
    T x, y, z;
    Point3_() : x(0), y(0), z(0) {}
    Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;

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

struct Scalar {
    double val[4];
    Scalar(double a = 0, double b = 0, double c = 0, double d = 0) { val[0] = a; val[1] = b; val[2] = c; val[3] = d; }
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

enum { GEMM_1_T = 1, GEMM_2_T = 2, GEMM_3_T = 4 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
enum { DECOMP_LU = 0, DECOMP_SVD = 1 };

class Mat;
class MatExpr;

// Matrix header over shared, malloc-backed storage (never operator new: oracle/ref_wrap.cc's bump arena must only see
// the octree node allocations) or over caller-owned memory.
class Mat {
public:
    struct Step {
        size_t v;
        Step() : v(0) {}
        operator size_t() const { return v; }
    };
    int flags_type;   // CV_MAKETYPE value
    int rows, cols;
    uchar* data;
    Step step;

    Mat() : flags_type(0), rows(0), cols(0), data(0), blk(0) {}
    Mat(const Mat& o);
    Mat& operator=(const Mat& o);
    ~Mat();
    Mat(Size sz, int type);
    Mat(int r, int c, int type);
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0);   // caller-owned pixels
    Mat& operator=(const MatExpr& e);
    Mat& operator=(const Scalar& s);

    Mat operator()(const Rect& r) const;
    Mat rowRange(int a, int b) const;
    Mat colRange(int a, int b) const;
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }
    Mat clone() const;
    void copyTo(const class _OutputArray& dst) const;
    void convertTo(Mat& dst, int rtype, double alpha = 1, double beta = 0) const;
    Mat reshape(int cn, int rows = 0) const;
    void create(int r, int c, int type);
    void release();

    static MatExpr zeros(int r, int c, int type);
    static MatExpr ones(int r, int c, int type);
    static MatExpr eye(int r, int c, int type);
    MatExpr t() const;
    MatExpr inv(int method = DECOMP_LU) const;
    MatExpr mul(const Mat& m, double scale = 1) const;
    double dot(const Mat& m) const;
    Mat cross(const Mat& m) const;

    int type() const { return flags_type; }
    int depth() const { return flags_type & 7; }
    int channels() const { return (flags_type >> CV_CN_SHIFT) + 1; }
    size_t elemSize1() const { static const int s[8] = {1, 1, 2, 2, 4, 4, 8, 2}; return (size_t)s[depth()]; }
    size_t elemSize() const { return elemSize1() * (size_t)channels(); }
    bool empty() const { return data == 0 || rows == 0 || cols == 0; }
    size_t total() const { return (size_t)rows * (size_t)cols; }
    bool isContinuous() const { return step.v == (size_t)cols * elemSize() || rows == 1; }
    size_t step1() const { return step.v / elemSize1(); }
    Size size() const { return Size(cols, rows); }

    template <typename T> T& at(int r, int c) { return *(T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    template <typename T> const T& at(int r, int c) const { return *(const T*)(data + (size_t)r * step.v + (size_t)c * sizeof(T)); }
    // single index: element i of a row or column vector (cv::Mat::at(int i0))
    template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : (cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols)); }
    template <typename T> const T& at(int i) const { return rows == 1 ? at<T>(0, i) : (cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols)); }
    uchar* ptr(int r = 0) { return data + (size_t)r * step.v; }
    const uchar* ptr(int r = 0) const { return data + (size_t)r * step.v; }
    template <typename T> T* ptr(int r = 0) { return (T*)(data + (size_t)r * step.v); }
    template <typename T> const T* ptr(int r = 0) const { return (const T*)(data + (size_t)r * step.v); }

private:
    int* blk;   // malloc'ed block: [refcount, pad to 64 B][pixels]; 0 for an empty header or caller-owned pixels
};

// The few cv::MatExpr shapes the tested code forms.  Evaluation (operator Mat) follows cv::MatOp_*::assign.
class MatExpr {
public:
    enum Kind { CONST_FILL, EYE, TRANSPOSE, GEMM, SCALED, INVERT, MULT };
    Kind kind;
    Mat a, b, c;        // GEMM: alpha * op(a) * b + beta * c
    double alpha, beta;
    int gemm_flags;
    int r, cN, type;    // CONST_FILL / EYE
    MatExpr() : kind(SCALED), alpha(1), beta(0), gemm_flags(0), r(0), cN(0), type(0) {}
    Mat eval() const;
    operator Mat() const { return eval(); }
    template <typename T> T at(int i, int j) const { return eval().at<T>(i, j); }
    template <typename T> T at(int i) const { return eval().at<T>(i); }
    MatExpr t() const;
    Mat row(int y) const { return eval().row(y); }
    Mat col(int x) const { return eval().col(x); }
    double dot(const Mat& m) const { return eval().dot(m); }
};

MatExpr operator-(const MatExpr& e);
MatExpr operator-(const Mat& m);
MatExpr operator*(const Mat& a, const Mat& b);
MatExpr operator*(const MatExpr& a, const Mat& b);
MatExpr operator*(const Mat& a, const MatExpr& b);
MatExpr operator*(const MatExpr& a, const MatExpr& b);
MatExpr operator*(double s, const Mat& m);
MatExpr operator*(const Mat& m, double s);
MatExpr operator*(double s, const MatExpr& e);
MatExpr operator*(const MatExpr& e, double s);
MatExpr operator/(const Mat& m, double s);
MatExpr operator/(const MatExpr& e, double s);
MatExpr operator+(const MatExpr& e, const Mat& m);
MatExpr operator+(const Mat& m, const MatExpr& e);
MatExpr operator+(const MatExpr& e, const MatExpr& f);
Mat operator+(const Mat& a, const Mat& b);
Mat operator-(const Mat& a, const Mat& b);
Mat operator-(const Mat& a, const MatExpr& e);
Mat operator-(const MatExpr& e, const Mat& b);
Mat operator-(const MatExpr& e, const MatExpr& f);

// cv::Mat_<T>(r, c) << a, b, c: the comma initialiser
template <typename T> struct MatDepth;
template <> struct MatDepth<float> { enum { value = CV_32F }; };
template <> struct MatDepth<double> { enum { value = CV_64F }; };
template <> struct MatDepth<uchar> { enum { value = CV_8U }; };
template <> struct MatDepth<int> { enum { value = CV_32S }; };

template <typename T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c, MatDepth<T>::value) {}
    Mat_(const Mat& m) : Mat(m) {}
    T& operator()(int r, int c) { return this->template at<T>(r, c); }
};
template <typename T> class MatCommaInitializer_ {
public:
    MatCommaInitializer_(Mat_<T>* _m) : m(_m), i(0) {}
    MatCommaInitializer_& operator,(T v) { m->template at<T>(i / m->cols, i % m->cols) = v; ++i; return *this; }
    operator Mat() const { return *m; }
    operator Mat_<T>() const { return *m; }
    Mat_<T>* m;
    int i;
};
template <typename T> inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, T v)
{
    MatCommaInitializer_<T> ci(const_cast<Mat_<T>*>(&m));
    return (ci, v);
}

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
    _OutputArray(const Mat& _m) : m(const_cast<Mat*>(&_m)) {}   // cv: lets a temporary view be a destination
    void create(int r, int c, int type) const { m->create(r, c, type); }
    void release() const { m->release(); }
    Mat getMat() const { return *m; }
    Mat& getMatRef() const { return *m; }
private:
    Mat* m;
};
typedef const _OutputArray& OutputArray;

inline _InputArray noArray() { return _InputArray(); }

double norm(const Mat& a, int normType = NORM_L2);
double norm(const Mat& a, const Mat& b, int normType = NORM_L2);
inline double norm(const MatExpr& e, int normType = NORM_L2) { return norm(e.eval(), normType); }
// cv::gemm for CV_32F (exported for the golden-vector check)
void cv_gemm32f(const Mat& A, const Mat& B, double alpha, const Mat* C, double beta, Mat& D, int flags);

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
void undistortPoints(InputArray src, OutputArray dst, InputArray K, InputArray dist, InputArray R = noArray(),
                     InputArray P = noArray());

// cv::FileStorage / FileNode: declarations only, so DBoW2's TemplatedVocabulary.h parses.  Never executed.
class FileNode;
class FileNodeIterator;
class FileNode {
public:
    FileNode operator[](const char*) const;
    FileNode operator[](const std::string&) const;
    FileNode operator[](int) const;
    size_t size() const;
    bool empty() const;
    operator int() const;
    operator float() const;
    operator double() const;
    operator std::string() const;
    FileNodeIterator begin() const;
    FileNodeIterator end() const;
};
class FileNodeIterator {
public:
    FileNode operator*() const;
    FileNodeIterator& operator++();
    bool operator!=(const FileNodeIterator&) const;
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage();
    FileStorage(const std::string& filename, int flags);
    ~FileStorage();
    bool isOpened() const;
    void release();
    FileNode operator[](const char*) const;
    FileNode operator[](const std::string&) const;
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { std::abort(); return fs; }
template <typename T> inline void operator>>(const FileNode&, T&) { std::abort(); }

}  // namespace cv

#endif
