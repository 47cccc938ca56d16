// oracle/cvshim/cvshim.cc -- TEST INFRASTRUCTURE, not product code.
// Implementation of the OpenCV stand-in declared in opencv2/core/core.hpp, on top of the
// pinned primitives in oracle/orb_prims.c.
#include "opencv2/core/core.hpp"
#include "../orb_prims.h"
#include <cstdlib>

namespace cv {

static const size_t HDR = 64;   // refcount header in front of the pixels

static int* blk_alloc(size_t bytes)
{
    int* b = (int*)std::malloc(HDR + bytes + 64);
    *b = 1;
    return b;
}

Mat::Mat(Size sz, int) : rows(0), cols(0), data(0), blk(0) { create(sz.height, sz.width, CV_8UC1); }
Mat::Mat(int r, int c, int) : rows(0), cols(0), data(0), blk(0) { create(r, c, CV_8UC1); }

Mat::Mat(const Mat& o) : rows(o.rows), cols(o.cols), data(o.data), step(o.step), blk(o.blk)
{
    if (blk) __atomic_add_fetch(blk, 1, __ATOMIC_RELAXED);
}

Mat& Mat::operator=(const Mat& o)
{
    if (this == &o) return *this;
    if (o.blk) __atomic_add_fetch(o.blk, 1, __ATOMIC_RELAXED);
    release();
    rows = o.rows; cols = o.cols; data = o.data; step = o.step; blk = o.blk;
    return *this;
}

Mat::~Mat() { release(); }

void Mat::release()
{
    if (blk && __atomic_sub_fetch(blk, 1, __ATOMIC_ACQ_REL) == 0) std::free(blk);
    blk = 0; data = 0; rows = cols = 0; step.v = 0;
}

void Mat::create(int r, int c, int)
{
    if (data && rows == r && cols == c) return;   // OpenCV keeps matching storage (views stay valid)
    release();
    if (r <= 0 || c <= 0) return;
    blk = blk_alloc((size_t)r * (size_t)c);
    data = (uchar*)blk + HDR;
    rows = r; cols = c; step.v = (size_t)c;
}

Mat Mat::operator()(const Rect& r) const
{
    Mat m(*this);
    m.data = data + (size_t)r.y * step.v + (size_t)r.x;
    m.rows = r.height; m.cols = r.width;
    return m;
}

Mat Mat::rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
Mat Mat::colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }

Mat Mat::clone() const
{
    Mat m;
    m.create(rows, cols, CV_8UC1);
    for (int y = 0; y < rows; ++y) std::memcpy(m.ptr(y), ptr(y), (size_t)cols);
    return m;
}

Mat::ZerosExpr Mat::zeros(int r, int c, int) { ZerosExpr e; e.r = r; e.c = c; return e; }

Mat::ZerosExpr::operator Mat() const
{
    Mat m;
    m = *this;
    return m;
}

Mat& Mat::operator=(const ZerosExpr& e)
{
    create(e.r, e.c, CV_8UC1);
    for (int y = 0; y < rows; ++y) std::memset(ptr(y), 0, (size_t)cols);
    return *this;
}

void KeyPointsFilter::retainBest(std::vector<KeyPoint>&, int) { std::abort(); }   // dead code path only

void resize(InputArray _src, OutputArray _dst, Size dsize, double, double, int)
{
    Mat src = _src.getMat();
    _dst.create(dsize.height, dsize.width, CV_8UC1);
    Mat dst = _dst.getMat();
    orbo_resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int)
{
    // The extractor always passes top=bottom=left=right and REFLECT_101; src may be the interior
    // view of dst (src/ORBextractor.cc:1127), which is safe here: only border cells are written
    // from interior cells, and the interior itself is copied with memmove.
    Mat src = _src.getMat();
    _dst.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
    Mat dst = _dst.getMat();
    const int w = src.cols, h = src.rows;
    for (int y = 0; y < h; ++y) std::memmove(dst.ptr(y + top) + left, src.ptr(y), (size_t)w);
    // fill the frame from a private copy of the interior (malloc: keep pixel data out of the arena)
    Mat inner = dst(Rect(left, top, w, h));
    (void)right; (void)bottom;
    uchar* tmp = (uchar*)std::malloc((size_t)w * (size_t)h);
    for (int y = 0; y < h; ++y) std::memcpy(tmp + (size_t)y * w, inner.ptr(y), (size_t)w);
    orbo_border_reflect101_u8(tmp, w, h, (size_t)w, dst.data, dst.step, top);
    std::free(tmp);
}

void GaussianBlur(InputArray _src, OutputArray _dst, Size, double, double, int)
{
    Mat src = _src.getMat();
    _dst.create(src.rows, src.cols, CV_8UC1);
    Mat dst = _dst.getMat();
    orbo_gauss7_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.step);
}

void FAST(InputArray _img, std::vector<KeyPoint>& keypoints, int threshold, bool nonmax)
{
    Mat img = _img.getMat();
    keypoints.clear();
    const int cap = ((img.cols + 1) / 2) * ((img.rows + 1) / 2) + 16;
    int* buf = (int*)std::malloc(sizeof(int) * 3 * (size_t)cap);
    int n = orbo_fast9_16(img.data, img.cols, img.rows, img.step, threshold, nonmax ? 1 : 0,
                          buf, buf + cap, buf + 2 * cap, cap);
    if (n > cap) n = cap;
    for (int i = 0; i < n; ++i)
        keypoints.push_back(KeyPoint((float)buf[i], (float)buf[cap + i], 7.f, -1.f, (float)buf[2 * cap + i]));
    std::free(buf);
}

float fastAtan2(float y, float x) { return orbo_fast_atan2(y, x); }

}  // namespace cv
