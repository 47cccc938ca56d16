/* oracle/orb_port.h -- TEST INFRASTRUCTURE, not product code.
 *
 * CPU restatement ("port") of the reference's ORB front end and Hamming search, written from
 * the behaviour of /root/reference/src/ORBextractor.cc and src/ORBmatcher.cc (file:line cited
 * per function in orb_port.cc).  Parity status: PINNED -- the port is checked (tests/) against
 * (a) cv2 4.13.0 golden vectors for the pixel primitives (tests/golden/) and (b) the unmodified
 * reference extractor compiled here (oracle/_ref/liborbref.so, ascending-arena tie-break pin).
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may use it.
 */
#ifndef ORB_PORT_H
#define ORB_PORT_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orbport orbport;
typedef struct { float x, y, size, angle, response; int32_t octave, class_id; } orbport_keypoint; /* cv::KeyPoint */

orbport* orbport_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th);
void orbport_destroy(orbport*);
/* scale4 = [scale | inv scale | sigma2 | inv sigma2] x nlevels; feats[nlevels]; umax[16] */
int orbport_tables(orbport*, float* scale4, int* feats, int* umax);

/* ORBextractor::operator().  Returns 0, -1 if cap is too small (n_out still set). */
int orbport_extract(orbport*, const uint8_t* img, int w, int h, size_t stride,
                    orbport_keypoint* kps, uint8_t* desc, int cap, int* n_out);

/* per-stage views of the last orbport_extract / orbport_pyramid call */
int orbport_pyramid(orbport*, const uint8_t* img, int w, int h, size_t stride);
int orbport_level_size(orbport*, int level, int* w, int* h);
int orbport_level_pixels(orbport*, int level, uint8_t* dst, size_t dstride);
int orbport_level_blurred(orbport*, int level, uint8_t* dst, size_t dstride);
/* FAST candidates / octree survivors of a level: coordinates relative to the (16,16) cell-grid
 * origin (src/ORBextractor.cc:824-825), in list order.  Return the count (writes <= cap). */
int orbport_level_candidates(orbport*, int level, int* x, int* y, int* score, int cap);
int orbport_level_selected(orbport*, int level, int* x, int* y, int* score, int cap);

/* DistributeOctTree on an arbitrary candidate list; W = maxX-minX, H = maxY-minY.
 * out_idx receives indices into the input list, in result order.  Returns count, or -3 if the
 * aspect ratio gives zero root nodes (undefined behaviour in the reference). */
int orbport_octree(const int* xs, const int* ys, const int* resp, int n, int W, int H, int N,
                   int* out_idx, int cap);

/* ORBmatcher::DescriptorDistance and the best / second-best scan */
int orbport_hamming256(const uint8_t* a, const uint8_t* b);
void orbport_match_bruteforce(const uint8_t* A, int nA, const uint8_t* B, int nB,
                              int32_t* best_idx, uint16_t* best, uint16_t* second);
/* accepted-match count of one keyframe pair: best <= th_low && float(best) < ratio*float(second) */
int orbport_match_count(const uint8_t* A, int nA, const uint8_t* B, int nB, int th_low, float ratio);
/* all-pairs counts over a keyframe table (BASELINE config 5), `threads` host threads; popcnt = 0: the reference's SWAR
 * DescriptorDistance (src/ORBmatcher.cc:1647-1663), 1: hardware popcount.  Returns wall seconds (the timed CPU baseline). */
double orbport_count_pairs_bench(const uint8_t* desc, int n_per_kf, const int32_t* pairs, int num_pairs, int th_low, float ratio,
                                 int32_t* counts, int threads, int popcnt);

/* Frame::ComputeStereoMatches (src/Frame.cc:466-640).  PINNED: the reference's own src/Frame.cc compiles
 * unchanged against oracle/cvshim2 into oracle/_ref/libframeref.so (oracle/ref_frame_wrap.cc) and
 * tests/test_ref_frame.py checks this restatement bit-for-bit against Frame::ComputeStereoMatches executed there.
 * lvlL/lvlR[l] = dense unblurred level l (w[l] x h[l], pitch w[l]) of the left / right pyramid; pixels outside a
 * level are its REFLECT_101 continuation (the 19-px frame of mvImagePyramid).  Outputs mvuRight / mvDepth [nL]. */
void orbport_stereo_matches(const orbport_keypoint* kL, const uint8_t* dL, int nL, const orbport_keypoint* kR,
                            const uint8_t* dR, int nR, const uint8_t* const* lvlL, const uint8_t* const* lvlR,
                            const int* w, const int* h, const float* scale, const float* inv_scale, float bf, float b,
                            float* u_right, float* depth);

#ifdef __cplusplus
}
#endif
#endif
