/* oracle/orb_prims.c -- TEST INFRASTRUCTURE, not product code.  See orb_prims.h.
 * Must be compiled with -ffp-contract=off: every float expression below is meant to be
 * evaluated as separate IEEE operations unless it calls fma() explicitly. */
#include "orb_prims.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

int orbo_round(double v) { return (int)lrint(v); }   /* default rounding mode = ties-to-even */
int orbo_roundf(float v) { return (int)lrintf(v); }

/* ------------------------------------------------------------------------------------------
 * Bilinear resize, 8U.  Follows OpenCV imgproc resize.cpp (INTER_LINEAR, fixed-point path):
 * coefficient scale 2^11, horizontal pass to int32, vertical pass with the >>4 / >>16 / +2>>2
 * cascade.  Reference call site: src/ORBextractor.cc:1123.
 * ---------------------------------------------------------------------------------------- */
void orbo_resize_axis_tables(int ssize, int dsize, int* ofs, short* a0, short* a1)
{
    double inv_scale = (double)dsize / ssize;
    double scale = 1.0 / inv_scale;
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (s < 0) { s = 0; f = 0.f; }
        if (s >= ssize - 1) { s = ssize - 1; f = 0.f; }
        float c0 = 1.f - f;
        ofs[d] = s;
        a0[d] = (short)orbo_roundf(c0 * 2048);
        a1[d] = (short)orbo_roundf(f * 2048);
    }
}

void orbo_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride,
                           uint8_t* dst, int dw, int dh, size_t dstride)
{
    int* xofs = (int*)malloc(sizeof(int) * (size_t)dw);
    short* xa0 = (short*)malloc(sizeof(short) * (size_t)dw);
    short* xa1 = (short*)malloc(sizeof(short) * (size_t)dw);
    int* row0 = (int*)malloc(sizeof(int) * (size_t)dw);
    int* row1 = (int*)malloc(sizeof(int) * (size_t)dw);
    orbo_resize_axis_tables(sw, dw, xofs, xa0, xa1);

    double scale_y = 1.0 / ((double)dh / sh);
    for (int dy = 0; dy < dh; ++dy) {
        /* the vertical axis keeps its fraction and clips the two row indices instead */
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)floorf(fy);
        fy -= (float)sy;
        short b0 = (short)orbo_roundf((1.f - fy) * 2048);
        short b1 = (short)orbo_roundf(fy * 2048);
        int y0 = sy < 0 ? 0 : (sy < sh ? sy : sh - 1);
        int y1 = sy + 1 < 0 ? 0 : (sy + 1 < sh ? sy + 1 : sh - 1);
        const uint8_t* S0 = src + (size_t)y0 * sstride;
        const uint8_t* S1 = src + (size_t)y1 * sstride;
        for (int dx = 0; dx < dw; ++dx) {
            int s = xofs[dx];
            int s1 = s + 1 < sw ? s + 1 : sw - 1;
            row0[dx] = S0[s] * xa0[dx] + S0[s1] * xa1[dx];
            row1[dx] = S1[s] * xa0[dx] + S1[s1] * xa1[dx];
        }
        uint8_t* D = dst + (size_t)dy * dstride;
        for (int dx = 0; dx < dw; ++dx) {
            int v = (((b0 * (row0[dx] >> 4)) >> 16) + ((b1 * (row1[dx] >> 4)) >> 16) + 2) >> 2;
            D[dx] = (uint8_t)v;
        }
    }
    free(xofs); free(xa0); free(xa1); free(row0); free(row1);
}

/* ------------------------------------------------------------------------------------------
 * 7x7 Gaussian, sigma 2, 8U, REFLECT_101.  OpenCV >= 4 evaluates it in unsigned fixed point:
 * 8.8 taps [18,34,48,56,48,34,18], 16-bit horizontal sums, 32-bit vertical sums, round by
 * +2^15 >> 16.  Reference call site: src/ORBextractor.cc:1088.
 * ---------------------------------------------------------------------------------------- */
static inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * (n - 1) - i;
    }
    return i;
}

void orbo_gauss7_u8(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride)
{
    /* horizontal pass over a REFLECT_101-padded copy of each row, then a 7-row vertical pass;
     * straight loops so the compiler can vectorise (this is also the timed CPU baseline) */
    uint16_t* H = (uint16_t*)malloc(sizeof(uint16_t) * (size_t)w * (size_t)h);
    uint8_t* P = (uint8_t*)malloc((size_t)w + 6);
    for (int y = 0; y < h; ++y) {
        const uint8_t* S = src + (size_t)y * sstride;
        for (int k = 0; k < 3; ++k) { P[k] = S[reflect101(k - 3, w)]; P[w + 3 + k] = S[reflect101(w + k, w)]; }
        memcpy(P + 3, S, (size_t)w);
        uint16_t* Hr = H + (size_t)y * w;
        for (int x = 0; x < w; ++x)
            Hr[x] = (uint16_t)(18u * (P[x] + P[x + 6]) + 34u * (P[x + 1] + P[x + 5]) + 48u * (P[x + 2] + P[x + 4]) + 56u * P[x + 3]);
    }
    for (int y = 0; y < h; ++y) {
        const uint16_t* R0 = H + (size_t)reflect101(y - 3, h) * w; const uint16_t* R1 = H + (size_t)reflect101(y - 2, h) * w;
        const uint16_t* R2 = H + (size_t)reflect101(y - 1, h) * w; const uint16_t* R3 = H + (size_t)y * w;
        const uint16_t* R4 = H + (size_t)reflect101(y + 1, h) * w; const uint16_t* R5 = H + (size_t)reflect101(y + 2, h) * w;
        const uint16_t* R6 = H + (size_t)reflect101(y + 3, h) * w;
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) {
            uint32_t acc = 18u * ((uint32_t)R0[x] + R6[x]) + 34u * ((uint32_t)R1[x] + R5[x]) + 48u * ((uint32_t)R2[x] + R4[x]) + 56u * (uint32_t)R3[x];
            D[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
    free(H); free(P);
}

void orbo_border_reflect101_u8(const uint8_t* src, int w, int h, size_t sstride,
                               uint8_t* dst, size_t dstride, int b)
{
    for (int y = -b; y < h + b; ++y) {
        const uint8_t* S = src + (size_t)reflect101(y, h) * sstride;
        uint8_t* D = dst + (size_t)(y + b) * dstride;
        for (int x = -b; x < w + b; ++x) D[x + b] = S[reflect101(x, w)];
    }
}

/* ------------------------------------------------------------------------------------------
 * FAST-9/16 with non-max suppression.  Follows OpenCV features2d fast.cpp / fast_score.cpp:
 * circle of 16 at radius 3; a pixel is a corner at threshold t iff 9 contiguous circle pixels
 * are all > v+t or all < v-t; its score is (largest t for which that still holds), which
 * equals V-1 with V = max over arcs of the arc's minimum signed difference.  NMS keeps a
 * corner iff its score is strictly greater than its 8 neighbours' scores (non-corners and the
 * 3-pixel rim count as 0).  Output is row-major.  Reference call sites:
 * src/ORBextractor.cc:811,816.
 * ---------------------------------------------------------------------------------------- */
static const int CIRC_DX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int CIRC_DY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

int orbo_fast_arc_value(const uint8_t* center, size_t stride)
{
    int d[16];
    int v = center[0];
    for (int k = 0; k < 16; ++k)
        d[k] = v - (int)center[(ptrdiff_t)CIRC_DY[k] * (ptrdiff_t)stride + CIRC_DX[k]];
    int best = -255;
    for (int k = 0; k < 16; ++k) {
        int mn = 255, mx = -255;               /* min of d (pixel darker arc) / max of d */
        for (int j = 0; j < 9; ++j) {
            int e = d[(k + j) & 15];
            if (e < mn) mn = e;
            if (e > mx) mx = e;
        }
        if (mn > best) best = mn;              /* all circle pixels darker than centre by >= mn */
        if (-mx > best) best = -mx;            /* all brighter by >= -mx */
    }
    return best;
}

int orbo_fast9_16(const uint8_t* img, int w, int h, size_t stride, int threshold, int nonmax,
                  int* xs, int* ys, int* scores, int cap)
{
    if (threshold < 0) threshold = 0;
    if (threshold > 255) threshold = 255;
    if (w < 7 || h < 7) return 0;
    uint8_t* S = (uint8_t*)calloc((size_t)w * (size_t)h, 1);     /* score map, 0 = not a corner */
    uint8_t* C = (uint8_t*)calloc((size_t)w * (size_t)h, 1);     /* corner flag */
    ptrdiff_t off[16];
    for (int k = 0; k < 16; ++k) off[k] = (ptrdiff_t)CIRC_DY[k] * (ptrdiff_t)stride + CIRC_DX[k];
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            /* cheap rejection: an arc of 9 contains one pixel of every opposite pair, so each pair
             * must offer a pixel darker than v-t (bit 0) or brighter than v+t (bit 1) */
            const uint8_t* c = img + (size_t)y * stride + x;
            const int lo = c[0] - threshold, hi = c[0] + threshold;
#define ORBO_CLS(k) ((c[off[k]] < lo ? 1 : 0) | (c[off[k]] > hi ? 2 : 0))
            int d = ORBO_CLS(0) | ORBO_CLS(8);
            if (!d) continue;
            d &= ORBO_CLS(4) | ORBO_CLS(12);
            if (!d) continue;
            d &= ORBO_CLS(2) | ORBO_CLS(10);
            d &= ORBO_CLS(6) | ORBO_CLS(14);
            if (!d) continue;
            d &= ORBO_CLS(1) | ORBO_CLS(9);
            d &= ORBO_CLS(3) | ORBO_CLS(11);
            d &= ORBO_CLS(5) | ORBO_CLS(13);
            d &= ORBO_CLS(7) | ORBO_CLS(15);
            if (!d) continue;
#undef ORBO_CLS
            int V = orbo_fast_arc_value(c, stride);
            if (V > threshold) {
                C[(size_t)y * w + x] = 1;
                S[(size_t)y * w + x] = (uint8_t)(V - 1);
            }
        }
    int n = 0;
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            if (!C[(size_t)y * w + x]) continue;
            int s = S[(size_t)y * w + x];
            int keep = 1;
            if (nonmax) {
                for (int dy = -1; dy <= 1 && keep; ++dy)
                    for (int dx = -1; dx <= 1; ++dx) {
                        if (!dx && !dy) continue;
                        if (s <= S[(size_t)(y + dy) * w + (x + dx)]) { keep = 0; break; }
                    }
            }
            if (keep) {
                if (n < cap) { xs[n] = x; ys[n] = y; scores[n] = nonmax ? s : 0; }
                ++n;
            }
        }
    free(S); free(C);
    return n;
}

/* ------------------------------------------------------------------------------------------
 * cv::fastAtan2 (core mathfuncs_core): degree-7 odd polynomial of min/max ratio, constants
 * pre-multiplied by 180/pi in float.  Reference call site: src/ORBextractor.cc:105.
 * ---------------------------------------------------------------------------------------- */
float orbo_fast_atan2(float y, float x)
{
    const float k180pi = (float)(180.0 / 3.1415926535897932384626433832795);
    const float P1 = 0.9997878412794807f * k180pi;
    const float P3 = -0.3258083974640975f * k180pi;
    const float P5 = 0.1555786518463281f * k180pi;
    const float P7 = -0.04432655554792128f * k180pi;
    const float eps = (float)2.2204460492503131e-16;
    float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + eps);
        c2 = c * c;
        a = (((P7 * c2 + P5) * c2 + P3) * c2 + P1) * c;
    } else {
        c = ax / (ay + eps);
        c2 = c * c;
        a = 90.f - (((P7 * c2 + P5) * c2 + P3) * c2 + P1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

/* ------------------------------------------------------------------------------------------
 * sincosf as glibc 2.39 evaluates it on an FMA-capable x86-64 (the ifunc picks the -mfma build
 * of sysdeps/x86_64/fpu/s_sincosf.c): reduce by pi/2 in double with one fused step, then the
 * sine and cosine minimax polynomials evaluated pairwise with fused multiply-adds.
 * Constants are glibc's __sincosf_table.  Reference call site: src/ORBextractor.cc:115
 * (cos(float)/sin(float) compile to one sincosf call).
 * ---------------------------------------------------------------------------------------- */
void orbo_sincosf(float y, float* sinp, float* cosp)
{
    static const double HPI_INV = 0x1.45f306dc9c883p+23;   /* 2/pi * 2^24 */
    static const double HPI = 0x1.921fb54442d18p+0;
    static const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5,
                        C3 = -0x1.6c087e89a359dp-10, C4 = 0x1.99343027bf8c3p-16;
    static const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7,
                        S3 = -0x1.994eb3774cf24p-13;
    static const double SIGN[4] = {1.0, -1.0, -1.0, 1.0};
    uint32_t bits; memcpy(&bits, &y, 4);
    uint32_t top12 = (bits >> 20) & 0x7ff;
    double x = (double)y, xs, flip = 1.0;
    int n = 0;
    if (top12 < 0x3f4) {                 /* |y| < pi/4 */
        if (top12 < 0x398) { *sinp = y; *cosp = 1.0f; return; }   /* |y| < 2^-12 */
        xs = x;
    } else {                             /* pi/4 <= |y| < 120 */
        double r = x * HPI_INV;
        n = ((int32_t)r + 0x800000) >> 24;
        x = fma(-(double)n, HPI, x);
        xs = x * SIGN[n & 3];
        if (n & 2) flip = -1.0;          /* second table: cosine coefficients negated */
    }
    double x2 = x * x;
    double sp = fma(x2, S3, S2);
    double cp = fma(x2, flip * C4, flip * C3);
    double x3 = x2 * xs, x4 = x2 * x2;
    double x5 = x2 * x3, x6 = x2 * x4;
    double c1 = fma(x2, flip * C1, flip * C0);
    double s = fma(x3, S1, xs);
    double c = fma(x4, flip * C2, c1);
    float fs = (float)fma(sp, x5, s);
    float fc = (float)fma(cp, x6, c);
    if (n & 1) { *sinp = fc; *cosp = fs; } else { *sinp = fs; *cosp = fc; }
}

/* Order-independent checksum of (x, sin x, cos x) over the float bit patterns first + i*step, i < count.
 * which = 0: this file's restatement; 1: the host's glibc sincosf.  Same mixing as the CUDA test hook. */
extern void sincosf(float, float*, float*);
uint64_t orbo_sincos_checksum(uint32_t first, uint32_t count, uint32_t step, int which)
{
    uint64_t sum = 0;
    for (uint32_t i = 0; i < count; ++i) {
        uint32_t bits = first + i * step, sb, cb;
        float x, s, c;
        memcpy(&x, &bits, 4);
        if (which) sincosf(x, &s, &c); else orbo_sincosf(x, &s, &c);
        memcpy(&sb, &s, 4); memcpy(&cb, &c, 4);
        uint64_t h = (uint64_t)bits * 0x9E3779B97F4A7C15ull ^ (((uint64_t)sb << 32) | cb);
        h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        sum += h;
    }
    return sum;
}


/* ---- cv::undistortPoints, OpenCV 4.13 (see orb_prims.h) ------------------------------------------------- */
void orbo_undistort_points(const float* xy, int n, const float* K4, const float* dist, int ndist, float* out)
{
    double k[14] = {0};
    const double fx = K4[0], fy = K4[1], cx = K4[2], cy = K4[3];
    const double ifx = 1. / fx, ify = 1. / fy;
    for (int i = 0; i < ndist && i < 14; ++i) k[i] = dist[i];
    for (int i = 0; i < n; ++i) {
        const double u = xy[2 * i], v = xy[2 * i + 1];
        double x = (u - cx) * ifx, y = (v - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; ++j) {                       /* TermCriteria(MAX_ITER, 5, 0.01): count only */
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = (u - cx) * ifx; y = (v - cy) * ify; break; }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        /* RR = P * I with P = K */
        const double xx = fx * x + 0. * y + cx;
        const double yy = 0. * x + fy * y + cy;
        const double ww = 1. / (0. * x + 0. * y + 1.);
        out[2 * i] = (float)(xx * ww);
        out[2 * i + 1] = (float)(yy * ww);
    }
}
