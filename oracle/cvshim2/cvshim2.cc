// oracle/cvshim2/cvshim2.cc -- TEST INFRASTRUCTURE, not product code.
// Implementation of the OpenCV stand-in declared in oracle/cvshim2/opencv2/core/core.hpp.  Pixel primitives and
// undistortPoints come from oracle/orb_prims.c (pinned to cv2 4.13.0 golden vectors); cv::gemm / norm / MatExpr
// evaluation are restated here from OpenCV 4.13 (modules/core/src/matmul.simd.hpp gemmImpl + GEMMSingleMul,
// matrix_expressions.cpp MatOp_GEMM / MatOp_T / MatOp_AddEx, norm.cpp) and checked against cv2 golden vectors.
#include <cmath>
#include "opencv2/core/core.hpp"
#include "../orb_prims.h"
#include <cstdio>

namespace cv {

static const size_t HDR = 64;   // refcount header in front of the pixels

static void unreachable(const char* what)
{
    std::fprintf(stderr, "oracle/cvshim2: %s is declared for parsing only and must not run\n", what);
    std::abort();
}

static int* blk_alloc(size_t bytes)
{
    int* b = (int*)std::malloc(HDR + bytes + 64);
    *b = 1;
    return b;
}

Mat::Mat(Size sz, int type) : flags_type(0), rows(0), cols(0), data(0), blk(0) { create(sz.height, sz.width, type); }
Mat::Mat(int r, int c, int type) : flags_type(0), rows(0), cols(0), data(0), blk(0) { create(r, c, type); }
Mat::Mat(int r, int c, int type, void* ext, size_t ext_step)
    : flags_type(type), rows(r), cols(c), data((uchar*)ext), blk(0)
{
    step.v = ext_step ? ext_step : (size_t)c * elemSize();
}

Mat::Mat(const Mat& o) : flags_type(o.flags_type), rows(o.rows), cols(o.cols), data(o.data), step(o.step), blk(o.blk)
{
    if (blk) __atomic_add_fetch(blk, 1, __ATOMIC_RELAXED);
}

Mat& Mat::operator=(const Mat& o)
{
    if (this == &o) return *this;
    if (o.blk) __atomic_add_fetch(o.blk, 1, __ATOMIC_RELAXED);
    release();
    flags_type = o.flags_type; rows = o.rows; cols = o.cols; data = o.data; step = o.step; blk = o.blk;
    return *this;
}

Mat::~Mat() { release(); }

void Mat::release()
{
    if (blk && __atomic_sub_fetch(blk, 1, __ATOMIC_ACQ_REL) == 0) std::free(blk);
    blk = 0; data = 0; rows = cols = 0; step.v = 0;
}

void Mat::create(int r, int c, int type)
{
    if (data && rows == r && cols == c && flags_type == type) return;   // OpenCV keeps matching storage
    release();
    flags_type = type;
    if (r <= 0 || c <= 0) return;
    const size_t es = elemSize();
    blk = blk_alloc((size_t)r * (size_t)c * es);
    data = (uchar*)blk + HDR;
    rows = r; cols = c; step.v = (size_t)c * es;
}

Mat Mat::operator()(const Rect& r) const
{
    Mat m(*this);
    m.data = data + (size_t)r.y * step.v + (size_t)r.x * elemSize();
    m.rows = r.height; m.cols = r.width;
    return m;
}

Mat Mat::rowRange(int a, int b) const { return (*this)(Rect(0, a, cols, b - a)); }
Mat Mat::colRange(int a, int b) const { return (*this)(Rect(a, 0, b - a, rows)); }

void Mat::copyTo(OutputArray dst) const
{
    Mat& m = dst.getMatRef();
    if (empty()) { m.release(); return; }
    m.create(rows, cols, flags_type);   // a destination view of matching shape is written in place
    const size_t rb = (size_t)cols * elemSize();
    for (int y = 0; y < rows; ++y) std::memmove(m.ptr(y), ptr(y), rb);
}

Mat Mat::clone() const
{
    Mat m;
    copyTo(m);
    return m;
}

void Mat::convertTo(Mat& dst, int rtype, double alpha, double beta) const
{
    const int ddepth = rtype & 7;
    if (alpha != 1 || beta != 0 || channels() != 1) unreachable("convertTo with scale / channels");
    Mat src = *this;                 // keeps the pixels alive when dst is *this (src/Frame.cc:564)
    Mat out;
    out.create(rows, cols, ddepth);
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < cols; ++x) {
            double v;
            switch (src.depth()) {
                case CV_8U: v = src.at<uchar>(y, x); break;
                case CV_32F: v = src.at<float>(y, x); break;
                case CV_64F: v = src.at<double>(y, x); break;
                default: unreachable("convertTo source depth"); v = 0;
            }
            switch (ddepth) {
                case CV_32F: out.at<float>(y, x) = (float)v; break;
                case CV_64F: out.at<double>(y, x) = v; break;
                case CV_8U: { int iv = cvRound(v); out.at<uchar>(y, x) = (uchar)(iv < 0 ? 0 : iv > 255 ? 255 : iv); break; }
                default: unreachable("convertTo destination depth");
            }
        }
    dst = out;
}

Mat Mat::reshape(int cn, int new_rows) const
{
    if (new_rows != 0 || !isContinuous()) unreachable("reshape with a row count / on a view");
    Mat m(*this);
    const int total_ch = cols * channels();
    if (cn == 0) cn = channels();
    if (total_ch % cn) unreachable("reshape: channels do not divide the row");
    m.flags_type = CV_MAKETYPE(depth(), cn);
    m.cols = total_ch / cn;
    return m;
}

Mat& Mat::operator=(const Scalar& s)
{
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < cols * channels(); ++x) {
            switch (depth()) {
                case CV_8U: ptr<uchar>(y)[x] = (uchar)cvRound(s.val[0]); break;
                case CV_32F: ptr<float>(y)[x] = (float)s.val[0]; break;
                case CV_64F: ptr<double>(y)[x] = s.val[0]; break;
                default: unreachable("Scalar fill depth");
            }
        }
    return *this;
}

Mat& Mat::operator=(const MatExpr& e)
{
    // cv::MatOp_Initializer::assign: "m = Mat::zeros(...)" fills the EXISTING storage when the shape matches
    // (src/ORBextractor.cc:1039 relies on that to write through a rowRange view)
    if ((e.kind == MatExpr::CONST_FILL || e.kind == MatExpr::EYE) && data && rows == e.r && cols == e.cN &&
        flags_type == e.type) {
        *this = Scalar(e.kind == MatExpr::EYE ? 0.0 : e.alpha);
        if (e.kind == MatExpr::EYE)
            for (int i = 0; i < std::min(rows, cols); ++i) {
                if (depth() == CV_32F) at<float>(i, i) = (float)e.alpha;
                else if (depth() == CV_64F) at<double>(i, i) = e.alpha;
                else at<uchar>(i, i) = (uchar)e.alpha;
            }
        return *this;
    }
    return *this = e.eval();
}

static MatExpr make_fill(MatExpr::Kind k, int r, int c, int type, double v)
{
    MatExpr e;
    e.kind = k; e.r = r; e.cN = c; e.type = type; e.alpha = v;
    return e;
}
MatExpr Mat::zeros(int r, int c, int type) { return make_fill(MatExpr::CONST_FILL, r, c, type, 0.0); }
MatExpr Mat::ones(int r, int c, int type) { return make_fill(MatExpr::CONST_FILL, r, c, type, 1.0); }
MatExpr Mat::eye(int r, int c, int type) { return make_fill(MatExpr::EYE, r, c, type, 1.0); }

MatExpr Mat::t() const
{
    MatExpr e;
    e.kind = MatExpr::TRANSPOSE; e.a = *this; e.alpha = 1;
    return e;
}
MatExpr Mat::inv(int) const
{
    MatExpr e;
    e.kind = MatExpr::INVERT; e.a = *this;
    return e;
}
MatExpr Mat::mul(const Mat& m, double scale) const
{
    MatExpr e;
    e.kind = MatExpr::MULT; e.a = *this; e.b = m; e.alpha = scale;
    return e;
}

// cv::Mat::dot for CV_32F: dotProd_32f accumulates in double (blocks of float products widened to double)
double Mat::dot(const Mat& m) const
{
    if (depth() != CV_32F || m.depth() != CV_32F) unreachable("dot on non-float matrices");
    double s = 0;
    for (int y = 0; y < rows; ++y)
        for (int x = 0; x < cols; ++x) s += (double)at<float>(y, x) * (double)m.at<float>(y, x);
    return s;
}
Mat Mat::cross(const Mat&) const { unreachable("Mat::cross"); return Mat(); }

// ---- cv::gemm, CV_32F -------------------------------------------------------------------------------------------------
// OpenCV 4.13 gemmImpl: with flags == 0 and inner length 2..4 equal to one side of D, products and sums are formed in
// float, left to right, then d = (float)(t*alpha + c*beta) with alpha, beta as doubles.  Every other case goes through
// GEMMSingleMul<float,double>: s += double(a)*double(b) in k order, s *= alpha, d = float(s + double(c)*beta).
void cv_gemm32f(const Mat& A, const Mat& B, double alpha, const Mat* C, double beta, Mat& D, int flags)
{
    if (A.depth() != CV_32F || B.depth() != CV_32F || A.channels() != 1) unreachable("gemm on non-float matrices");
    const bool at = (flags & GEMM_1_T) != 0, bt = (flags & GEMM_2_T) != 0;
    if (flags & GEMM_3_T) unreachable("gemm with a transposed C");
    const int M = at ? A.cols : A.rows, K = at ? A.rows : A.cols;
    const int Kb = bt ? B.cols : B.rows, N = bt ? B.rows : B.cols;
    if (K != Kb) unreachable("gemm: inner sizes differ");
    if (C && C->empty()) C = 0;
    Mat out;
    out.create(M, N, CV_32F);
    const bool small = flags == 0 && K >= 2 && K <= 4 && (K == N || K == M);
    for (int i = 0; i < M; ++i)
        for (int j = 0; j < N; ++j) {
            const float cij = C ? C->at<float>(i, j) : 0.f;
            if (small) {
                float t = A.at<float>(i, 0) * B.at<float>(0, j);
                for (int k = 1; k < K; ++k) t = t + A.at<float>(i, k) * B.at<float>(k, j);
                out.at<float>(i, j) = (float)((double)t * alpha + (double)cij * (C ? beta : 0.0));
            } else {
                double s = 0;
                for (int k = 0; k < K; ++k) {
                    const float a = at ? A.at<float>(k, i) : A.at<float>(i, k);
                    const float b = bt ? B.at<float>(j, k) : B.at<float>(k, j);
                    s += (double)a * (double)b;
                }
                s = s * alpha;
                out.at<float>(i, j) = C ? (float)(s + (double)cij * beta) : (float)s;
            }
        }
    D = out;
}

static Mat transpose_of(const Mat& a)
{
    Mat o;
    o.create(a.cols, a.rows, a.type());
    const size_t es = a.elemSize();
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) std::memcpy(o.ptr(x) + (size_t)y * es, a.ptr(y) + (size_t)x * es, es);
    return o;
}

// element-wise m*alpha for CV_32F = cv::Mat::convertTo with a scale: OpenCV's cvtScale32f narrows the scale to float first
// and multiplies in float, v * (float)alpha (pinned to cv2 4.13.0: tests/golden/scale_golden.npz, made through
// cv2.normalize(NORM_INF), which ends in the same convertTo call; float(double(v) * alpha) differs in the last bit often)
static Mat scaled_of(const Mat& m, double alpha)
{
    if (alpha == 1) return m;
    if (m.depth() != CV_32F) unreachable("scaling a non-float matrix");
    Mat o;
    o.create(m.rows, m.cols, m.type());
    for (int y = 0; y < m.rows; ++y)
        for (int x = 0; x < m.cols; ++x) o.at<float>(y, x) = m.at<float>(y, x) * (float)alpha;
    return o;
}

Mat MatExpr::eval() const
{
    switch (kind) {
        case CONST_FILL:
        case EYE: {
            Mat m;
            m.create(r, cN, type);
            m = Scalar(kind == EYE ? 0.0 : alpha);
            if (kind == EYE)
                for (int i = 0; i < std::min(r, cN); ++i) {
                    if (m.depth() == CV_32F) m.at<float>(i, i) = (float)alpha;
                    else if (m.depth() == CV_64F) m.at<double>(i, i) = alpha;
                    else m.at<uchar>(i, i) = (uchar)alpha;
                }
            return m;
        }
        case TRANSPOSE:
            return scaled_of(transpose_of(a), alpha);      // MatOp_T::assign: transpose, then convertTo with alpha
        case SCALED:
            return scaled_of(a, alpha);
        case GEMM: {
            Mat d;
            cv_gemm32f(a, b, alpha, c.empty() ? 0 : &c, beta, d, gemm_flags);
            return d;
        }
        case INVERT: unreachable("Mat::inv");
        case MULT: unreachable("Mat::mul");
    }
    return Mat();
}

MatExpr MatExpr::t() const
{
    if (kind == TRANSPOSE) { MatExpr e; e.kind = SCALED; e.a = a; e.alpha = alpha; return e; }
    return eval().t();
}

// cv::operator-(const MatExpr&) = e.op->subtract(Scalar(0), e, res).  Only MatOp_AddEx overrides that (it flips alpha);
// for every other op -- MatOp_T included -- the default MatOp::subtract MATERIALISES e and returns AddEx(m, alpha = -1).
// So "-R.t()*t" (src/ORBmatcher.cc:1341, src/Frame.cc:266) is gemm(transposed copy, t, alpha = -1, flags = 0): the
// small-matrix float case, NOT a GEMM_1_T call.
MatExpr operator-(const MatExpr& e)
{
    MatExpr n;
    n.kind = MatExpr::SCALED;
    if (e.kind == MatExpr::SCALED) { n.a = e.a; n.alpha = -e.alpha; }
    else { n.a = e.eval(); n.alpha = -1; }
    return n;
}
MatExpr operator-(const Mat& m)
{
    MatExpr n;
    n.kind = MatExpr::SCALED; n.a = m; n.alpha = -1;
    return n;
}

// MatOp::matmul: a transposed / scaled left or right operand folds into the gemm flags and alpha
static void gemm_operand(const MatExpr& e, Mat& m, double& alpha, bool& transposed)
{
    if (e.kind == MatExpr::TRANSPOSE) { m = e.a; alpha = e.alpha; transposed = true; }
    else if (e.kind == MatExpr::SCALED) { m = e.a; alpha = e.alpha; transposed = false; }
    else { m = e.eval(); alpha = 1; transposed = false; }
}
static MatExpr make_gemm(const Mat& a, bool at, const Mat& b, bool bt, double alpha)
{
    MatExpr e;
    e.kind = MatExpr::GEMM; e.a = a; e.b = b; e.alpha = alpha; e.beta = 0;
    e.gemm_flags = (at ? GEMM_1_T : 0) | (bt ? GEMM_2_T : 0);
    return e;
}
MatExpr operator*(const Mat& a, const Mat& b) { return make_gemm(a, false, b, false, 1); }
MatExpr operator*(const MatExpr& ea, const Mat& b)
{
    Mat a; double al; bool at;
    gemm_operand(ea, a, al, at);
    return make_gemm(a, at, b, false, al);
}
MatExpr operator*(const Mat& a, const MatExpr& eb)
{
    Mat b; double al; bool bt;
    gemm_operand(eb, b, al, bt);
    return make_gemm(a, false, b, bt, al);
}
MatExpr operator*(const MatExpr& ea, const MatExpr& eb)
{
    Mat a, b; double al, bl; bool at, bt;
    gemm_operand(ea, a, al, at);
    gemm_operand(eb, b, bl, bt);
    return make_gemm(a, at, b, bt, al * bl);
}
MatExpr operator*(double s, const Mat& m)
{
    MatExpr e;
    e.kind = MatExpr::SCALED; e.a = m; e.alpha = s;
    return e;
}
MatExpr operator*(const Mat& m, double s) { return s * m; }
MatExpr operator*(double s, const MatExpr& e)
{
    if (e.kind == MatExpr::GEMM) { MatExpr n = e; n.alpha *= s; n.beta *= s; return n; }
    if (e.kind == MatExpr::INVERT || e.kind == MatExpr::MULT || e.kind == MatExpr::EYE) return s * e.eval();
    MatExpr n = e;
    n.alpha = e.alpha * s;
    return n;
}
MatExpr operator*(const MatExpr& e, double s) { return s * e; }
MatExpr operator/(const Mat& m, double s) { return (1.0 / s) * m; }
MatExpr operator/(const MatExpr& e, double s) { return (1.0 / s) * e; }

MatExpr operator+(const MatExpr& e, const Mat& m)
{
    if (e.kind == MatExpr::GEMM && e.c.empty()) {   // MatOp_GEMM::add: alpha*A*B + 1*C
        MatExpr n = e;
        n.c = m; n.beta = 1;
        return n;
    }
    if (e.kind == MatExpr::SCALED && e.alpha != 1 && e.alpha != -1 && !e.a.empty()) {
        // MatOp::add folds "alpha * A" and a plain matrix into MatOp_AddEx(A, alpha, M, 1); its assign() then calls
        // cv::scaleAdd(A, alpha, M): alpha narrowed to float, one fused multiply-add per element (OpenCV 4.13.0 as
        // installed: cv2.scaleAdd golden vectors, tests/golden/scale_add_golden.npz)
        if (e.a.depth() != CV_32F || m.depth() != CV_32F || e.a.rows != m.rows || e.a.cols != m.cols)
            unreachable("scaleAdd on mismatching matrices");
        Mat o;
        o.create(m.rows, m.cols, m.type());
        const float af = (float)e.alpha;
        for (int y = 0; y < m.rows; ++y)
            for (int x = 0; x < m.cols; ++x) o.at<float>(y, x) = std::fmaf(e.a.at<float>(y, x), af, m.at<float>(y, x));
        MatExpr n;
        n.kind = MatExpr::SCALED; n.a = o; n.alpha = 1;
        return n;
    }
    MatExpr n;
    n.kind = MatExpr::SCALED; n.a = e.eval() + m; n.alpha = 1;
    return n;
}
MatExpr operator+(const Mat& m, const MatExpr& e) { return e + m; }
MatExpr operator+(const MatExpr& e, const MatExpr& f) { return e + f.eval(); }

static Mat addsub(const Mat& a, const Mat& b, float sign)
{
    if (a.depth() != CV_32F || b.depth() != CV_32F || a.rows != b.rows || a.cols != b.cols)
        unreachable("element-wise +/- on mismatching matrices");
    Mat o;
    o.create(a.rows, a.cols, a.type());
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x)
            o.at<float>(y, x) = sign > 0 ? a.at<float>(y, x) + b.at<float>(y, x) : a.at<float>(y, x) - b.at<float>(y, x);
    return o;
}
Mat operator+(const Mat& a, const Mat& b) { return addsub(a, b, 1.f); }
Mat operator-(const Mat& a, const Mat& b) { return addsub(a, b, -1.f); }
Mat operator-(const Mat& a, const MatExpr& e) { return addsub(a, e.eval(), -1.f); }   // MatOp::subtract: e2 materialised
Mat operator-(const MatExpr& e, const Mat& b) { return addsub(e.eval(), b, -1.f); }
Mat operator-(const MatExpr& e, const MatExpr& f) { return addsub(e.eval(), f.eval(), -1.f); }

// cv::norm: CV_32F sums are accumulated in double (normL1_/normL2Sqr_ with ST = double)
double norm(const Mat& a, int normType)
{
    if (a.depth() != CV_32F) unreachable("norm on a non-float matrix");
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) {
            const double v = a.at<float>(y, x);
            s += normType == NORM_L1 ? std::fabs(v) : v * v;
        }
    return normType == NORM_L1 ? s : std::sqrt(s);
}
double norm(const Mat& a, const Mat& b, int normType)
{
    if (a.depth() != CV_32F || b.depth() != CV_32F) unreachable("norm on non-float matrices");
    double s = 0;
    for (int y = 0; y < a.rows; ++y)
        for (int x = 0; x < a.cols; ++x) {
            const double v = (double)(a.at<float>(y, x) - b.at<float>(y, x));   // normDiffL1_: abs(a - b) in float
            s += normType == NORM_L1 ? std::fabs(v) : v * v;
        }
    return normType == NORM_L1 ? s : std::sqrt(s);
}

void KeyPointsFilter::retainBest(std::vector<KeyPoint>&, int) { unreachable("KeyPointsFilter::retainBest"); }

void resize(InputArray _src, OutputArray _dst, Size dsize, double, double, int)
{
    Mat src = _src.getMat();
    _dst.create(dsize.height, dsize.width, CV_8UC1);
    Mat dst = _dst.getMat();
    orbo_resize_linear_u8(src.data, src.cols, src.rows, src.step, dst.data, dst.cols, dst.rows, dst.step);
}

void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int)
{
    Mat src = _src.getMat();
    _dst.create(src.rows + top + bottom, src.cols + left + right, CV_8UC1);
    Mat dst = _dst.getMat();
    const int w = src.cols, h = src.rows;
    for (int y = 0; y < h; ++y) std::memmove(dst.ptr(y + top) + left, src.ptr(y), (size_t)w);
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

// cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) as src/Frame.cc:422,448 call it: N x 1 CV_32FC2, in place
void undistortPoints(InputArray _src, OutputArray _dst, InputArray _K, InputArray _dist, InputArray _R, InputArray _P)
{
    Mat src = _src.getMat(), K = _K.getMat(), dist = _dist.getMat(), P = _P.getMat();
    if (!_R.empty()) unreachable("undistortPoints with a rectification");
    if (src.type() != CV_32FC2 || src.cols != 1 || K.depth() != CV_32F || dist.depth() != CV_32F || P.data != K.data)
        unreachable("undistortPoints in a form the extractor path does not use");
    const float K4[4] = {K.at<float>(0, 0), K.at<float>(1, 1), K.at<float>(0, 2), K.at<float>(1, 2)};
    const int nd = (int)dist.total();
    float d[8] = {0};
    for (int i = 0; i < nd && i < 8; ++i) d[i] = dist.at<float>(i);
    const int n = src.rows;
    std::vector<float> in((size_t)n * 2), out((size_t)n * 2);
    for (int i = 0; i < n; ++i) { in[2 * i] = src.ptr<float>(i)[0]; in[2 * i + 1] = src.ptr<float>(i)[1]; }
    orbo_undistort_points(in.data(), n, K4, d, nd, out.data());
    _dst.create(n, 1, CV_32FC2);
    Mat dst = _dst.getMat();
    for (int i = 0; i < n; ++i) { dst.ptr<float>(i)[0] = out[2 * i]; dst.ptr<float>(i)[1] = out[2 * i + 1]; }
}

// ---- parse-only API ---------------------------------------------------------------------------------------------------
FileNode FileNode::operator[](const char*) const { unreachable("FileNode"); return FileNode(); }
FileNode FileNode::operator[](const std::string&) const { unreachable("FileNode"); return FileNode(); }
FileNode FileNode::operator[](int) const { unreachable("FileNode"); return FileNode(); }
size_t FileNode::size() const { unreachable("FileNode"); return 0; }
bool FileNode::empty() const { unreachable("FileNode"); return true; }
FileNode::operator int() const { unreachable("FileNode"); return 0; }
FileNode::operator float() const { unreachable("FileNode"); return 0; }
FileNode::operator double() const { unreachable("FileNode"); return 0; }
FileNode::operator std::string() const { unreachable("FileNode"); return std::string(); }
FileNodeIterator FileNode::begin() const { unreachable("FileNode"); return FileNodeIterator(); }
FileNodeIterator FileNode::end() const { unreachable("FileNode"); return FileNodeIterator(); }
FileNode FileNodeIterator::operator*() const { unreachable("FileNodeIterator"); return FileNode(); }
FileNodeIterator& FileNodeIterator::operator++() { unreachable("FileNodeIterator"); return *this; }
bool FileNodeIterator::operator!=(const FileNodeIterator&) const { unreachable("FileNodeIterator"); return false; }
FileStorage::FileStorage() {}
FileStorage::FileStorage(const std::string&, int) { unreachable("FileStorage"); }
FileStorage::~FileStorage() {}
bool FileStorage::isOpened() const { return false; }
void FileStorage::release() {}
FileNode FileStorage::operator[](const char*) const { unreachable("FileStorage"); return FileNode(); }
FileNode FileStorage::operator[](const std::string&) const { unreachable("FileStorage"); return FileNode(); }

}  // namespace cv
