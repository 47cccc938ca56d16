/* oracle/orb_prims.h -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatements of the OpenCV / libm primitives that the reference's ORB front end calls
 * (the arithmetic lives in an un-vendored dependency: OpenCV, unpinned by the reference's
 * CMakeLists.txt:33-39; pinned here to OpenCV 4.13.0 = the cv2 wheel available in the build
 * container, against which every primitive below is checked byte-for-byte by
 * tests/golden/make_golden.py and tests/test_oracle_prims.py).
 *
 * Call sites in the reference (file:line under /root/reference):
 *   resize          src/ORBextractor.cc:1123      -> orbo_resize_linear_u8
 *   copyMakeBorder  src/ORBextractor.cc:1127,1132 -> orbo_border_reflect101_u8
 *   FAST            src/ORBextractor.cc:811,816   -> orbo_fast9_16
 *   GaussianBlur    src/ORBextractor.cc:1088      -> orbo_gauss7_u8
 *   fastAtan2       src/ORBextractor.cc:105       -> orbo_fast_atan2
 *   cos/sin(float)  src/ORBextractor.cc:115       -> orbo_sincosf (glibc 2.39 x86-64 FMA variant)
 *   cvRound         src/ORBextractor.cc:83,117,...-> orbo_round (round-half-even)
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * use anything under oracle/.  Compile with -ffp-contract=off (the Makefile does).
 */
#ifndef ORBO_PRIMS_H
#define ORBO_PRIMS_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* cvRound(double|float): round to nearest, ties to even. */
int orbo_round(double v);
int orbo_roundf(float v);

/* cv::resize(src, dst, dsize, 0, 0, INTER_LINEAR) for CV_8UC1 (11-bit fixed-point coefficients). */
void orbo_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride,
                           uint8_t* dst, int dw, int dh, size_t dstride);
/* the per-axis tables of the same resize: ofs[d] = first source index, a0/a1 = 11-bit weights */
void orbo_resize_axis_tables(int ssize, int dsize, int* ofs, short* a0, short* a1);

/* cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) for CV_8UC1 (OpenCV >= 4.x
 * bit-exact fixed-point path, taps [18,34,48,56,48,34,18]/256).  src may equal dst. */
void orbo_gauss7_u8(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst, size_t dstride);

/* cv::copyMakeBorder(..., b,b,b,b, BORDER_REFLECT_101): writes the (w+2b) x (h+2b) image. */
void orbo_border_reflect101_u8(const uint8_t* src, int w, int h, size_t sstride,
                               uint8_t* dst, size_t dstride, int b);

/* cv::FAST(img, kps, threshold, nonmax=true), TYPE_9_16, on a w x h sub-image.
 * Writes up to cap (x, y, score) triples in OpenCV's output order (row-major); returns the
 * number found (may exceed cap; only the first cap are written). */
int orbo_fast9_16(const uint8_t* img, int w, int h, size_t stride, int threshold, int nonmax,
                  int* xs, int* ys, int* scores, int cap);
/* V(p) = max over the 16 arcs of 9 contiguous circle pixels of min(|d|) with one sign;
 * p is a FAST corner at threshold t iff V > t, and OpenCV's score is V-1. */
int orbo_fast_arc_value(const uint8_t* center, size_t stride);

/* cv::fastAtan2(y, x): degrees in [0,360), polynomial, plain float mul/add (no FMA). */
float orbo_fast_atan2(float y, float x);

/* glibc 2.39 sincosf, x86-64 FMA ifunc variant (sysdeps/x86_64/fpu/s_sincosf.c built -mfma):
 * restated from the published algorithm (double-precision polynomial, interleaved sin/cos
 * evaluation with fused multiply-adds).  Valid for |x| < 120. */
void orbo_sincosf(float x, float* s, float* c);

/* checksum of (x, sin x, cos x) over float bit patterns; which = 0 restatement, 1 host glibc */
uint64_t orbo_sincos_checksum(uint32_t first, uint32_t count, uint32_t step, int which);

/* cv::undistortPoints(src, dst, K, D, noArray(), K) of OpenCV 4.13 (calib3d undistort: 5 fixed-point iterations in
 * double, then re-projection with P = K), as Frame::UndistortKeyPoints / ComputeImageBounds call it
 * (src/Frame.cc:404-463).  K4 = fx, fy, cx, cy and dist = k1, k2, p1, p2[, k3] are the float values of mK / mDistCoef.
 * Pinned: bit-equal to cv2 4.13.0 on tests/golden/undistort_golden.npz. */
void orbo_undistort_points(const float* xy, int n, const float* K4, const float* dist, int ndist, float* out);

#ifdef __cplusplus
}
#endif
#endif
