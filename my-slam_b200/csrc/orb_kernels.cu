// orb_kernels.cu -- hand-written CUDA kernels (sm_100a) of the ORB front end.
//
// Stage -> reference function it reproduces bit-for-bit (paths relative to the reference tree):
//   k_resize_level   ORBextractor::ComputePyramid      src/ORBextractor.cc:1109-1137 (cv::resize INTER_LINEAR 8U)
//   k_fast_cells     ComputeKeyPointsOctTree cell loop src/ORBextractor.cc:767-831   (cv::FAST 9/16 + NMS + fallback)
//   k_octree         DistributeOctTree / DivideNode    src/ORBextractor.cc:483-765
//   k_blur           GaussianBlur 7x7 sigma 2          src/ORBextractor.cc:1083-1088
//   k_describe       IC_Angle + computeOrbDescriptor   src/ORBextractor.cc:79-149, 474-481, 1036-1043, 1097-1103
// All of it is byte / integer work bound by on-chip bandwidth and issue rate, not a dense float
// contraction: no tensor cores.  Float arithmetic that feeds rounding decisions uses explicit
// __fmul_rn/__fadd_rn (no FMA contraction) and __float2int_rn (round-half-even == cvRound).
#include <cuda_runtime.h>
#include <cuda_pipeline.h>
#include <stdint.h>
#include "orb_common.cuh"
#include "orb_pattern.h"

// rBRIEF pattern in global memory: lane i of a warp owns bytes [32i, 32i+32) = its 16 sample points
__device__ __align__(16) int8_t g_pattern[1024];
// umax[|v|] of the 31-px disc packed as nibbles (src/ORBextractor.cc:456-471 with HALF_PATCH_SIZE 15):
// 15,15,15,15,14,14,14,13,13,12,11,10,9,8,6,3.  The disc is symmetric, so column u holds rows |v| <= umax[|u|].
#define ORB_UMAX_NIBBLES 0x3689abcddeeeffffull

// IC_Angle weights for the word path of k_describe: the 31 x 31 window around a keypoint is read as 31 rows x 9 aligned
// words starting at column cx - 15 - s, s = (cx - 15) & 3.  Entry [s][row * 9 + word] = (the four bytes' u as signed bytes
// where the pixel lies in the disc, else 0; 1 per byte in the disc, else 0): row v = row - 15 holds |u| <= umax[|v|]
// (src/ORBextractor.cc:86-103).
#ifndef ORB_IC_WORDS_PATH
#define ORB_IC_WORDS_PATH 1
#endif
#define IC_WORDS 9
#define IC_ITEMS 288          // 31 * 9 = 279 items, padded to 9 per lane
__device__ __align__(16) uint2 g_ic_weights[4 * IC_ITEMS];

int orb_upload_constants(const int* umax16)
{
    // the disc table is a constant of the algorithm (HALF_PATCH_SIZE 15); the kernels carry it packed
    for (int i = 0; i < 16; ++i)
        if (umax16[i] != (int)((ORB_UMAX_NIBBLES >> (4 * i)) & 15ull)) return -1;
    if (cudaMemcpyToSymbol(g_pattern, ORB_PATTERN_XY, 1024) != cudaSuccess) return -1;
    static uint2 tab[4 * IC_ITEMS];
    for (int sft = 0; sft < 4; ++sft)
        for (int i = 0; i < IC_ITEMS; ++i) {
            const int row = i / IC_WORDS, word = i % IC_WORDS, v = row - ORB_HALF_PATCH;
            unsigned wu = 0, wm = 0;
            for (int b = 0; b < 4; ++b) {
                const int u = 4 * word + b - sft - ORB_HALF_PATCH;
                if (row < ORB_PATCH && abs(u) <= ORB_HALF_PATCH && abs(u) <= umax16[abs(v)]) {
                    wu |= (unsigned)(u & 0xFF) << (8 * b);
                    wm |= 1u << (8 * b);
                }
            }
            tab[sft * IC_ITEMS + i] = make_uint2(wu, wm);
        }
    if (cudaMemcpyToSymbol(g_ic_weights, tab, sizeof(tab)) != cudaSuccess) return -1;
    return 0;
}

// =================================================================================================
// Level 0 whose rows are not 4-byte aligned (dense 1241-px KITTI frames ...) is copied once into a buffer
// with a 16-byte-multiple pitch, so that every other kernel stays on its aligned word path.
// =================================================================================================
__global__ void __launch_bounds__(256) k_repitch(const uint8_t* __restrict__ src, long long src_frame_stride, int src_pitch,
                                                 uint8_t* __restrict__ dst, int dst_pitch, int w, int h)
{
    const int wpr = dst_pitch >> 2;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const int y = t / wpr, x4 = (t - y * wpr) * 4;
    if (y >= h) return;
    const uint8_t* __restrict__ r = src + (long long)blockIdx.y * src_frame_stride + (long long)y * src_pitch;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) v |= (uint32_t)__ldg(r + min(x4 + k, w - 1)) << (8 * k);
    *reinterpret_cast<uint32_t*>(dst + ((long long)blockIdx.y * h + y) * dst_pitch + x4) = v;
}

// cv::cvtColor(..., CV_BGR2GRAY / RGB2GRAY / BGRA2GRAY / RGBA2GRAY) on 8-bit pixels as Tracking::GrabImage* applies it
// before the extractor (src/Tracking.cc:244-257), in OpenCV 4.13's fixed point (pinned against cv2, tests/golden):
// gray = (c0*k0 + c1*19235 + c2*k2 + 2^14) >> 15 with (k0, k2) = (3735, 9798) for B,G,R order and swapped for R,G,B.
__global__ void __launch_bounds__(256) k_cvt_gray(const uint8_t* __restrict__ src, long long src_frame_stride, int src_pitch, int bpp,
                                                  int k0, int k2, uint8_t* __restrict__ dst, int dst_pitch, int w, int h)
{
    const int wpr = dst_pitch >> 2;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const int y = t / wpr, x4 = (t - y * wpr) * 4;
    if (y >= h) return;
    const uint8_t* __restrict__ r = src + (long long)blockIdx.y * src_frame_stride + (long long)y * src_pitch;
    uint32_t v = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint8_t* px = r + (long long)min(x4 + k, w - 1) * bpp;
        const int g = ((int)__ldg(px) * k0 + (int)__ldg(px + 1) * 19235 + (int)__ldg(px + 2) * k2 + (1 << 14)) >> 15;
        v |= (uint32_t)g << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dst + ((long long)blockIdx.y * h + y) * dst_pitch + x4) = v;
}

int orb_launch_cvt_gray(const uint8_t* src, long long src_frame_stride, int src_pitch, int channels, int rgb_order, uint8_t* dst,
                        int dst_pitch, int w, int h, int nframes, cudaStream_t st)
{
    const int ntask = (dst_pitch >> 2) * h;
    k_cvt_gray<<<dim3((ntask + 255) / 256, nframes), 256, 0, st>>>(src, src_frame_stride, src_pitch, channels, rgb_order ? 9798 : 3735,
                                                                    rgb_order ? 3735 : 9798, dst, dst_pitch, w, h);
    return 1;
}

int orb_launch_repitch(const uint8_t* src, long long src_frame_stride, int src_pitch, uint8_t* dst, int dst_pitch, int w, int h,
                       int nframes, cudaStream_t st)
{
    const int ntask = (dst_pitch >> 2) * h;
    k_repitch<<<dim3((ntask + 255) / 256, nframes), 256, 0, st>>>(src, src_frame_stride, src_pitch, dst, dst_pitch, w, h);
    return 1;
}

// =================================================================================================
// Pyramid: level l from level l-1, OpenCV's fixed-point bilinear (11-bit weights; horizontal pass in
// int32, vertical pass ((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2.  The per-axis tables
// (source index, weights) are built on the host in double/float exactly as OpenCV does.
// One thread = 4 horizontally adjacent output pixels = one aligned 32-bit store.
// =================================================================================================
#ifndef RS_ROWS
#define RS_ROWS 8             // output rows per thread
#endif
#ifndef RS_THREADS
#define RS_THREADS 128
#endif

// horizontal pass of one source row for this thread's 4 output pixels: H[j] = S[s_j]*a0_j + S[s_j + 1]*a1_j.
// WORDS: the 4 pixel pairs lie inside three aligned words starting at word wb (byte offsets o_j = s_j - 4wb).
// horizontal pass of one source row for this thread's 4 output pixels: H[j] = S[s_j]*a0_j + S[s_j + 1]*a1_j.
// WORDS: the 4 pixel pairs lie inside three aligned words starting at word wb.  Pixel j's pair (bytes o_j, o_j + 1
// of the 12-byte window) is cut out by one PRMT whose selector sel[j] was fixed at thread start; lowpair[j] tells
// whether it lies in words (0,1) or (1,2).
// NC: the source level was written by an earlier launch and may be read through the non-coherent path (__ldg); the
// fused tail kernel reads levels it wrote itself a barrier ago and must use ordinary loads.
template <bool NC> __device__ __forceinline__ unsigned rs_ldw(const uint32_t* p) { return NC ? __ldg(p) : *p; }
template <bool NC> __device__ __forceinline__ unsigned rs_ldb(const uint8_t* p) { return NC ? (unsigned)__ldg(p) : (unsigned)*p; }

template <bool WORDS, bool NC>
__device__ __forceinline__ void resize_hrow(const uint8_t* __restrict__ row, int wb, int wlast, const int s[4], const unsigned sel[4],
                                            const bool lowpair[4], const unsigned a01[4], int sw, int H[4])
{
    if (WORDS) {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(row);
        const unsigned w0 = rs_ldw<NC>(p + wb), w1 = rs_ldw<NC>(p + min(wb + 1, wlast)), w2 = rs_ldw<NC>(p + min(wb + 2, wlast));
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const unsigned pair = __byte_perm(lowpair[j] ? w0 : w1, lowpair[j] ? w1 : w2, sel[j]);
            H[j] = (int)__dp2a_lo(a01[j], pair, 0u);           // a0 * byte0 + a1 * byte1
        }
    } else {
#pragma unroll
        for (int j = 0; j < 4; ++j)
            H[j] = (int)(rs_ldb<NC>(row + s[j]) * (a01[j] & 0xFFFFu) + rs_ldb<NC>(row + min(s[j] + 1, sw - 1)) * (a01[j] >> 16));
    }
}

#ifndef RS_SRC_ROWS
#define RS_SRC_ROWS 12        // source rows one strip of RS_ROWS output rows may touch on the shared-memory path (a multiple of 6)
#endif

// s_h: this thread's private shared-memory column, rows STRIDE entries apart
template <bool WORDS, bool NC, int STRIDE>
__device__ __forceinline__ void resize_strip(const OrbPlanDev& P, const OrbLevel& D, const uint8_t* src, int sp,
                                             int sw, int sh, uint8_t* dst, int wc, int y0, const int s[4],
                                             const unsigned sel[4], const bool lowpair[4], const unsigned a01[4], int wb,
                                             int wlast, uint2* s_h)
{
    // y-axis table entries of all RS_ROWS rows first (independent loads)
    int sy[RS_ROWS];
    unsigned b01[RS_ROWS];
#pragma unroll
    for (int k = 0; k < RS_ROWS; ++k) {
        const int dy = min(y0 + k, D.h - 1);
        sy[k] = P.rs_ofs[D.yofs_off + dy];
        b01[k] = *reinterpret_cast<const unsigned*>(P.rs_a + 2 * (D.yofs_off + dy));
    }
    const int rfirst = min(max(sy[0], 0), sh - 1), rlast = min(max(sy[RS_ROWS - 1] + 1, 0), sh - 1);
    if (rlast - rfirst < RS_SRC_ROWS) {
        // every source row of the strip exactly once: H rows go to this thread's private shared-memory column
        // (no reuse logic, no divergence, all row loads independent), then the vertical pass reads two of them
        uint2* hcol = s_h;                                     // H >> 4 <= 32640: two 16-bit values per word
        if (WORDS) {
            // six rows at a time: their 18 word loads are issued before the first use (row indices clamped to the last
            // needed row instead of a branch per row, which kept each row's loads behind the previous row's arithmetic)
            const int wb1 = min(wb + 1, wlast), wb2 = min(wb + 2, wlast);
#pragma unroll
            for (int h = 0; h < RS_SRC_ROWS; h += 6) {
                unsigned w0[6], w1[6], w2[6];
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const uint32_t* p = reinterpret_cast<const uint32_t*>(src + (size_t)min(rfirst + h + i, rlast) * sp);
                    w0[i] = rs_ldw<NC>(p + wb); w1[i] = rs_ldw<NC>(p + wb1); w2[i] = rs_ldw<NC>(p + wb2);
                }
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    int H[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const unsigned pair = __byte_perm(lowpair[j] ? w0[i] : w1[i], lowpair[j] ? w1[i] : w2[i], sel[j]);
                        H[j] = (int)__dp2a_lo(a01[j], pair, 0u);
                    }
                    hcol[(h + i) * STRIDE] = make_uint2(__byte_perm(H[0] >> 4, H[1] >> 4, 0x5410), __byte_perm(H[2] >> 4, H[3] >> 4, 0x5410));
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < RS_SRC_ROWS; ++i) {
                if (rfirst + i <= rlast) {
                    int H[4];
                    resize_hrow<WORDS, NC>(src + (size_t)(rfirst + i) * sp, wb, wlast, s, sel, lowpair, a01, sw, H);
                    hcol[i * STRIDE] = make_uint2(__byte_perm(H[0] >> 4, H[1] >> 4, 0x5410), __byte_perm(H[2] >> 4, H[3] >> 4, 0x5410));
                }
            }
        }
#pragma unroll
        for (int k = 0; k < RS_ROWS; ++k) {
            const int dy = y0 + k;
            if (dy < D.h) {
                const int b0 = (int)(b01[k] & 0xFFFFu), b1 = (int)(b01[k] >> 16);
                const uint2 A = hcol[(min(max(sy[k], 0), sh - 1) - rfirst) * STRIDE];
                const uint2 B = hcol[(min(max(sy[k] + 1, 0), sh - 1) - rfirst) * STRIDE];
                const int v0 = (((b0 * (int)(A.x & 0xFFFFu)) >> 16) + ((b1 * (int)(B.x & 0xFFFFu)) >> 16) + 2) >> 2;
                const int v1 = (((b0 * (int)(A.x >> 16)) >> 16) + ((b1 * (int)(B.x >> 16)) >> 16) + 2) >> 2;
                const int v2 = (((b0 * (int)(A.y & 0xFFFFu)) >> 16) + ((b1 * (int)(B.y & 0xFFFFu)) >> 16) + 2) >> 2;
                const int v3 = (((b0 * (int)(A.y >> 16)) >> 16) + ((b1 * (int)(B.y >> 16)) >> 16) + 2) >> 2;
                *reinterpret_cast<uint32_t*>(dst + (size_t)dy * D.pitch + 4 * wc) =
                    __byte_perm(__byte_perm(v0, v1, 0x0040), __byte_perm(v2, v3, 0x0040), 0x5410);
            }
        }
        return;
    }
    // steep scale factors (more than RS_SRC_ROWS source rows per strip): row by row with a two-row cache
    int Ha[4], Hb[4], ya = -1, yb = -1;
#pragma unroll
    for (int k = 0; k < RS_ROWS; ++k) {
        const int dy = y0 + k;
        if (dy >= D.h) break;
        const int b0 = (int)(b01[k] & 0xFFFFu), b1 = (int)(b01[k] >> 16);
        const int r0 = min(max(sy[k], 0), sh - 1), r1 = min(max(sy[k] + 1, 0), sh - 1);
        if (r0 == yb) {
#pragma unroll
            for (int j = 0; j < 4; ++j) Ha[j] = Hb[j];
            ya = yb;
        } else if (r0 != ya) {
            resize_hrow<WORDS, NC>(src + (size_t)r0 * sp, wb, wlast, s, sel, lowpair, a01, sw, Ha);
            ya = r0;
        }
        if (r1 == ya) {
#pragma unroll
            for (int j = 0; j < 4; ++j) Hb[j] = Ha[j];
            yb = ya;
        } else if (r1 != yb) {
            resize_hrow<WORDS, NC>(src + (size_t)r1 * sp, wb, wlast, s, sel, lowpair, a01, sw, Hb);
            yb = r1;
        }
        uint32_t packed = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int v = (((b0 * (Ha[j] >> 4)) >> 16) + ((b1 * (Hb[j] >> 4)) >> 16) + 2) >> 2;
            packed |= (uint32_t)(v & 0xFF) << (8 * j);
        }
        *reinterpret_cast<uint32_t*>(dst + (size_t)dy * D.pitch + 4 * wc) = packed;
    }
}

// One task = 4 horizontally adjacent output pixels (one aligned 32-bit store per row) x RS_ROWS rows of level l.  The
// x-axis table entries of the 4 pixels stay in registers; a source row costs three aligned word loads and
// four IDP.2A dot products, and is reused when consecutive output rows share it.
template <bool NC, int STRIDE>
__device__ __forceinline__ void resize_task(const OrbPlanDev& P, int l, int frame, int task, uint2* s_h)
{
    const OrbLevel& D = P.lv[l];
    const int nwc = (D.w + 3) >> 2;
    const int strip = task / nwc, wc = task - strip * nwc;
    const int y0 = strip * RS_ROWS;
    if (y0 >= D.h) return;
    const int sw = P.lv[l - 1].w, sh = P.lv[l - 1].h;
    const uint8_t* src = orb_level_ptr(P, l - 1, frame);
    const int sp = orb_level_pitch(P, l - 1);
    uint8_t* dst = P.pyr + (long long)frame * P.pyr_frame_bytes + D.pyr_off;

    int s[4];
    unsigned a01[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int dx = min(4 * wc + j, D.w - 1);               // tail lanes recompute the last pixel (pitch is padded)
        s[j] = P.rs_ofs[D.xofs_off + dx];
        a01[j] = *reinterpret_cast<const unsigned*>(P.rs_a + 2 * (D.xofs_off + dx));   // a0 | a1 << 16
    }
    const int smin = min(min(s[0], s[1]), min(s[2], s[3])), smax = max(max(s[0], s[1]), max(s[2], s[3]));
    const int wb = smin >> 2;
    unsigned sel[4];
    bool lowpair[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int o = s[j] - 4 * wb;                           // byte offset of the pair inside the 12-byte window
        lowpair[j] = o <= 6;
        const int q = lowpair[j] ? o : o - 4;
        sel[j] = (unsigned)q | ((unsigned)(q + 1) << 4);
    }
    // word path: rows 4-byte aligned and every (s_j, s_j + 1) pair inside words wb .. wb + 2 (clamped to the
    // last word that lies inside the pitch; a pair byte beyond the row only ever meets weight a1 = 0)
    const int wlast = (sp >> 2) - 1;
    const bool words = ((sp & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 3) == 0) && (smax + 1 - 4 * wb <= 11) &&
                       (wb + 2 <= wlast || 4 * wlast + 3 >= min(smax + 1, sw - 1));
    if (words) resize_strip<true, NC, STRIDE>(P, D, src, sp, sw, sh, dst, wc, y0, s, sel, lowpair, a01, wb, wlast, s_h);
    else resize_strip<false, NC, STRIDE>(P, D, src, sp, sw, sh, dst, wc, y0, s, sel, lowpair, a01, wb, wlast, s_h);
}

__global__ void __launch_bounds__(RS_THREADS, 8) k_resize_level(OrbPlanDev P, int l)
{
    __shared__ uint2 s_h[RS_SRC_ROWS * RS_THREADS];            // [source row][thread]: H >> 4 of the thread's 4 pixels, 16 bits each
    resize_task<true, RS_THREADS>(P, l, blockIdx.y, blockIdx.x * RS_THREADS + threadIdx.x, s_h + threadIdx.x);
}

// The small levels in ONE launch: level l reads level l - 1, so every level used to be its own launch, and from about
// 300 x 230 px down a launch is 7-9 us of launch floor for a grid of a few CTAs per frame.  Here a CLUSTER of RT_CLUSTER
// CTAs owns one frame and walks the levels first .. nlevels - 1 with a cluster barrier in between (release / acquire at
// cluster scope: the level just written by the sibling CTAs is visible); levels written inside this launch are read with
// L2-coherent loads (ld.global.cg), never through the non-coherent path.
#define RT_THREADS 512
#ifndef RT_CLUSTER
#define RT_CLUSTER 4
#endif
#ifndef RT_MINB
#define RT_MINB 2
#endif
template <> __device__ __forceinline__ unsigned rs_ldw<false>(const uint32_t* p) { return __ldcg(p); }
template <> __device__ __forceinline__ unsigned rs_ldb<false>(const uint8_t* p) { return (unsigned)__ldcg(p); }

__global__ void __cluster_dims__(RT_CLUSTER, 1, 1) __launch_bounds__(RT_THREADS, RT_MINB) k_resize_tail(OrbPlanDev P, int first)
{
    extern __shared__ __align__(16) uint8_t rt_smem[];
    uint2* s_h = reinterpret_cast<uint2*>(rt_smem) + threadIdx.x;
    const int frame = blockIdx.x / RT_CLUSTER;
    const int t0 = (blockIdx.x % RT_CLUSTER) * RT_THREADS + threadIdx.x;
    for (int l = first; l < P.nlevels; ++l) {
        const OrbLevel& D = P.lv[l];
        const int ntask = ((D.w + 3) >> 2) * ((D.h + RS_ROWS - 1) / RS_ROWS);
        for (int t = t0; t < ntask; t += RT_CLUSTER * RT_THREADS) resize_task<false, RT_THREADS>(P, l, frame, t, s_h);
        if (l + 1 < P.nlevels) {                               // level l complete in every CTA of the cluster before level l + 1 reads it
            asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        }
    }
}

// first level of the fused tail: the first level whose grid is below RT_MAX_TASKS tasks per frame, if at least two
// levels follow from there; nlevels = no tail
// MEASURED AND SWITCHED OFF (B200, C1, profiles/r2_resize_tail_variants.txt): one 1024-thread CTA per frame 0.78 -> 0.91 ms
// of pyramid time per 512 frames; clusters of 2 / 4 / 8 CTAs x 512 threads 0.75 / 0.82 / 1.05 ms, and with every variant the
// overlapped pipeline LOSES 3-8 % (182 k -> 168-176 k frames/s): the tail's few long-lived, latency-bound CTAs sit on SMs
// the other chunks' kernels would fill, while four tiny launches leave those SMs free.  Build with -DORB_RESIZE_TAIL=1 to
// measure again.
#ifndef ORB_RESIZE_TAIL
#define ORB_RESIZE_TAIL 0
#endif
#define RT_MAX_TASKS 2400
int orb_resize_tail_first(const OrbPlanDev& P)
{
#if !ORB_RESIZE_TAIL
    return P.nlevels;
#else
    for (int l = 1; l < P.nlevels; ++l) {
        const int ntask = ((P.lv[l].w + 3) / 4) * ((P.lv[l].h + RS_ROWS - 1) / RS_ROWS);
        if (ntask <= RT_MAX_TASKS) return (P.nlevels - l >= 2) ? l : P.nlevels;
    }
    return P.nlevels;
#endif
}

int orb_launch_pyramid(const OrbPlanDev& P, int nframes, cudaStream_t st)
{
    int launches = 0;
    const int tail = orb_resize_tail_first(P);
    for (int l = 1; l < tail; ++l) {
        const int ntask = ((P.lv[l].w + 3) / 4) * ((P.lv[l].h + RS_ROWS - 1) / RS_ROWS);
        k_resize_level<<<dim3((ntask + RS_THREADS - 1) / RS_THREADS, nframes), RS_THREADS, 0, st>>>(P, l);
        ++launches;
    }
    if (tail < P.nlevels) {
        const int smem = RS_SRC_ROWS * RT_THREADS * (int)sizeof(uint2);
        static OrbSmemOptIn optin;
        if (optin.ensure(k_resize_tail, smem) != cudaSuccess) return -1;
        k_resize_tail<<<nframes * RT_CLUSTER, RT_THREADS, smem, st>>>(P, tail);
        ++launches;
    }
    return launches;
}

// =================================================================================================
// FAST-9/16 per 30-px cell with per-cell non-max suppression and the iniTh -> minTh fallback.
// One warp owns one cell: the cell image (<= 59 x 59 bytes) is staged in shared memory, the arc
// value V = max over the 16 arcs of 9 contiguous circle pixels of min(+-(v - p)) is evaluated for
// every interior pixel with packed 16x2 three-input min/max (VIMNMX3, both polarities at once),
// score = V-1 where V > min(iniTh, minTh).  NMS uses the cell-local score map (neighbours outside
// the cell interior count as 0, as in cv::FAST on the cell sub-image).  Survivors are the strict
// local maxima with score >= iniTh, or, if the cell has none, those with score >= minTh.  They
// are written in row-major order (warp ballot + popc prefix) to the cell's slot area.
// =================================================================================================
// Shared-memory row pitch of the cell tile: 48 B when every cell image of the shape (plus up to 3 B of alignment shift) fits,
// else 64 B.  With 48 B a CTA needs 37 KB, four of them fit the 164 KB carve-out step and the SM keeps 64 KB of L1 for
// whatever runs beside FAST (k_describe takes 2.3 times as long with the 32 KB that the 196 KB step leaves).
inline int orb_fast_pitch(int max_cw) { return max_cw + 3 <= 48 ? 48 : 64; }
#ifndef FAST_WARPS
#define FAST_WARPS 8
#endif

template <int FAST_TP>
__device__ __forceinline__ int fast_arc_value(const uint8_t* c)
{
    // e_k = (v - p_k + 256) | (p_k - v + 256) << 16, built with one IMAD per circle pixel
    const unsigned v = c[0];
    const unsigned kv = v * (1u - 65536u) + 256u * 65537u;
#define E(dx, dy) ((unsigned)c[(dy) * FAST_TP + (dx)] * 65535u + kv)
    const unsigned e0 = E(0, 3), e1 = E(1, 3), e2 = E(2, 2), e3 = E(3, 1), e4 = E(3, 0), e5 = E(3, -1),
                   e6 = E(2, -2), e7 = E(1, -3), e8 = E(0, -3), e9 = E(-1, -3), e10 = E(-2, -2),
                   e11 = E(-3, -1), e12 = E(-3, 0), e13 = E(-3, 1), e14 = E(-2, 2), e15 = E(-1, 3);
#undef E
#define MN3(a, b, c) __vimin3_s16x2(a, b, c)
#define MX3(a, b, c) __vimax3_s16x2(a, b, c)
    // minima of 3 consecutive, then of 9 consecutive (3 + 3 + 3)
    const unsigned t0 = MN3(e0, e1, e2), t1 = MN3(e1, e2, e3), t2 = MN3(e2, e3, e4), t3 = MN3(e3, e4, e5),
                   t4 = MN3(e4, e5, e6), t5 = MN3(e5, e6, e7), t6 = MN3(e6, e7, e8), t7 = MN3(e7, e8, e9),
                   t8 = MN3(e8, e9, e10), t9 = MN3(e9, e10, e11), t10 = MN3(e10, e11, e12),
                   t11 = MN3(e11, e12, e13), t12 = MN3(e12, e13, e14), t13 = MN3(e13, e14, e15),
                   t14 = MN3(e14, e15, e0), t15 = MN3(e15, e0, e1);
    const unsigned n0 = MN3(t0, t3, t6), n1 = MN3(t1, t4, t7), n2 = MN3(t2, t5, t8), n3 = MN3(t3, t6, t9),
                   n4 = MN3(t4, t7, t10), n5 = MN3(t5, t8, t11), n6 = MN3(t6, t9, t12), n7 = MN3(t7, t10, t13),
                   n8 = MN3(t8, t11, t14), n9 = MN3(t9, t12, t15), n10 = MN3(t10, t13, t0),
                   n11 = MN3(t11, t14, t1), n12 = MN3(t12, t15, t2), n13 = MN3(t13, t0, t3),
                   n14 = MN3(t14, t1, t4), n15 = MN3(t15, t2, t5);
    const unsigned a = MX3(n0, n1, n2), b = MX3(n3, n4, n5), d = MX3(n6, n7, n8), f = MX3(n9, n10, n11),
                   g = MX3(n12, n13, n14);
    const unsigned r = MX3(MX3(a, b, d), MX3(f, g, n15), n15);
#undef MN3
#undef MX3
    const int lo = (int)(r & 0xFFFFu), hi = (int)(r >> 16);
    return max(lo, hi) - 256;
}

#ifndef FAST_STAGE8
#define FAST_STAGE8 1         // stage cells with 8-byte cp.async where the cell is narrow enough for a shift of up to 7 columns
#endif
#ifndef FAST_P1_UNROLL
#define FAST_P1_UNROLL 3      // row groups of the quick test per queue insertion (1 .. 3)
#endif
#define FAST_QCAP (FAST_P1_UNROLL == 1 ? 256 : 512)   // ring queue of quick-test survivors (per warp): power of two >= 31 + 128 * unroll
#ifndef FAST_CCAP
#define FAST_CCAP 256         // corner list (per warp); denser cells take the dense path
#endif
#define FAST_KCAP 96          // strict maxima ranked by pairwise comparison; more take the dense path

// flags (bit 7 of each byte) of  max(a, b) > th  for 4 packed bytes, th <= 126; K = (127 - th) * 0x01010101
__device__ __forceinline__ unsigned fast_gt4(unsigned a, unsigned b, unsigned K)
{
    return ((a & 0x7f7f7f7fu) + K) | a | ((b & 0x7f7f7f7fu) + K) | b;
}

__device__ __forceinline__ int fast_max8(const uint8_t* s, int p)
{
    return max(max(max((int)s[-p - 1], (int)s[-p]), max((int)s[-p + 1], (int)s[-1])),
               max(max((int)s[1], (int)s[p - 1]), max((int)s[p], (int)s[p + 1])));
}

// One warp = one cell, exactly as the reference runs it: cv::FAST(cell, iniThFAST, nonmax) and, when that
// returns nothing, cv::FAST(cell, minThFAST, nonmax) (src/ORBextractor.cc:811-818).  Per attempt:
//   1. quick rejection on 4 packed pixels per lane: every 9-arc of the ring contains pixel 0 or 8 and pixel
//      4 or 12, so a corner needs max(|p0-v|,|p8-v|) > th and max(|p4-v|,|p12-v|) > th (VABSDIFF4 + carry trick);
//   2. survivors go through a ring queue and are scored 32 at a time by the 16x2 min/max arc network;
//      score = V - 1 where V > th, into the cell-local score map; corners are also appended to a list;
//   3. strict 8-neighbour maxima of the listed corners, ranked by position (row-major) by pairwise comparison
//      and written in order.  Cells too dense for the lists fall back to an ordered scan of the score map.
template <int FAST_TP>
__global__ void __launch_bounds__(FAST_WARPS * 32) k_fast_cells(OrbPlanDev P)
{
    const int tile_rows = P.fast_rows, spw = P.fast_spw;
    extern __shared__ __align__(16) uint8_t fast_smem[];
    __shared__ uint16_t s_queue[FAST_WARPS][FAST_QCAP];
    __shared__ uint16_t s_corner[FAST_WARPS][FAST_CCAP];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int cell_id = blockIdx.x * FAST_WARPS + warp;
    const int frame = blockIdx.y;
    if (cell_id >= P.ncells) return;
    const OrbCell cell = P.cells[cell_id];
    const int l = cell.level;
    // per warp: the cell tile (pitch FAST_TP) and the score map of the cell interior with a 1-px rim (pitch spw): entry
    // (x + 1) + (y + 1) * spw for interior pixel (x, y).  The map is kept as small as the cell allows: k_fast_cells' shared
    // memory decides the L1 carve-out of every SM it runs on (4 CTAs x 40 KB fit the 164 KB step, 47 KB did not).
    const int warp_bytes = FAST_TP * tile_rows + ((spw * (tile_rows - 4) + 15) & ~15);
    uint8_t* tile = fast_smem + (size_t)warp * warp_bytes;
    uint8_t* smap = tile + FAST_TP * tile_rows;
    const int cw = cell.cw, ch = cell.ch, iw = cw - 6, ih = ch - 6;
    uint32_t* cnt_out = P.cellcnt + (long long)frame * P.ncells + cell_id;
    if (iw <= 0 || ih <= 0) { if (lane == 0) *cnt_out = 0; return; }

    // ---- stage the cell image (tile column = x - ini_x + sh) with asynchronous 4-byte copies
    const int pitch = orb_level_pitch(P, l);
    const uint8_t* __restrict__ img = orb_level_ptr(P, l, frame) + (long long)cell.ini_y * pitch + cell.ini_x;
    int sh = 0;
    if (FAST_STAGE8 && ((pitch & 7) == 0) && ((reinterpret_cast<uintptr_t>(img - cell.ini_x) & 7) == 0) && cw + 7 <= FAST_TP) {
        // 8-byte copies: eight lanes per row, four rows per step (half the LDGSTS of the 4-byte form below)
        sh = cell.ini_x & 7;
        const int nw = (cw + sh + 7) >> 3, c = lane & 7;
        if (c < nw) {
            const uint8_t* g = img - sh + (size_t)(lane >> 3) * pitch + 8 * c;
            uint8_t* t = tile + (lane >> 3) * FAST_TP + 8 * c;
            for (int r = lane >> 3; r < ch; r += 4, g += 4 * (size_t)pitch, t += 4 * FAST_TP) __pipeline_memcpy_async(t, g, 8);
        }
        __pipeline_commit();
        __pipeline_wait_prior(0);
    } else if (((pitch & 3) == 0) && ((reinterpret_cast<uintptr_t>(img - cell.ini_x) & 3) == 0)) {
        sh = cell.ini_x & 3;
        const int nw = (cw + sh + 3) >> 2, c = lane & 15;
        if (c < nw) {
            const uint8_t* g = img - sh + (size_t)(lane >> 4) * pitch + 4 * c;
            uint8_t* t = tile + (lane >> 4) * FAST_TP + 4 * c;
            ORB_CHECK(ch <= tile_rows && 4 * nw <= FAST_TP);
            for (int r = lane >> 4; r < ch; r += 2, g += 2 * (size_t)pitch, t += 2 * FAST_TP) __pipeline_memcpy_async(t, g, 4);
        }
        __pipeline_commit();
        __pipeline_wait_prior(0);
    } else {
        for (int r = 0; r < ch; ++r)
            for (int c = lane; c < cw; c += 32) tile[r * FAST_TP + c] = __ldg(img + (long long)r * pitch + c);
    }
    __syncwarp();

    const int cs = 3 + sh, ce = cs + iw;                       // interior tile columns [cs, ce), rows [3, 3 + ih)
    const int w0 = cs >> 2, nwi = ((ce - 1) >> 2) - w0 + 1;    // words of a row that hold interior pixels
    const float inv_iw = 1.0f / (float)iw;
    // pass-1 lane layout, fixed for the cell: rpi = 32 / nwi rows per iteration, lane -> (row lrow, word w0 + lcol)
    // 32 / nwi and lane / nwi by one approximate float reciprocal (nwi <= 16; the 0.01 keeps exact quotients above the integer)
    const float inv_nwi = __fdividef(1.0f, (float)nwi);
    const int rpi = __float2int_rz(32.01f * inv_nwi), lrow = __float2int_rz(((float)lane + 0.01f) * inv_nwi), lcol = lane - lrow * nwi;
    const bool lrow_ok = lrow < rpi;
    const int loff = (3 + lrow) * FAST_TP + 4 * (w0 + lcol);   // tile offset of the lane's word in the first iteration
    unsigned lmask = 0x80808080u;                              // flag bits of the word's bytes that are interior pixels
    if (4 * (w0 + lcol) < cs) lmask &= 0xFFFFFFFFu << (8 * (cs - 4 * (w0 + lcol)));
    if (4 * (w0 + lcol) + 4 > ce) lmask &= 0xFFFFFFFFu >> (8 * (4 * (w0 + lcol) + 4 - ce));
    uint16_t* q = s_queue[warp];
    uint16_t* cl = s_corner[warp];
    uint32_t* slot = P.slots + (long long)frame * P.slot_frame_words + P.lv[l].slot_off + cell.slot_off;
    int cnt = 0;

    int ncor = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        const int th = attempt ? P.min_th : P.ini_th;
#ifdef FAST_SKIP_REZERO
        if (attempt == 0 || ncor > 0)                                       // a first attempt without a single corner left the map all zero
#endif
        {
            uint4* z = reinterpret_cast<uint4*>(smap);                      // interior + NMS rim
            for (int i = lane; i < ((ih + 2) * spw + 15) >> 4; i += 32) z[i] = make_uint4(0, 0, 0, 0);
        }
        __syncwarp();
        int head = 0, tail = 0;
        ncor = 0;

        // scores 32 queued pixels with the arc network; corners go to the score map and the corner list
#define FAST_DRAIN(active)                                                               \
        {                                                                                \
            int o_ = 0, V_ = 0;                                                          \
            if (active) { o_ = q[(head + lane) & (FAST_QCAP - 1)]; V_ = fast_arc_value<FAST_TP>(tile + o_); } \
            const bool c_ = V_ > th;                                                     \
            const unsigned bc_ = __ballot_sync(0xffffffffu, c_);                         \
            if (c_) {                                                                    \
                ORB_CHECK(o_ >= 3 * FAST_TP + 3 && o_ < (tile_rows - 3) * FAST_TP);      \
                smap[(o_ / FAST_TP - 2) * spw + o_ % FAST_TP - cs + 1] = (uint8_t)(V_ - 1);    \
                const int p_ = ncor + __popc(bc_ & lt);                                  \
                if (p_ < FAST_CCAP) cl[p_] = (uint16_t)o_;                               \
            }                                                                            \
            ncor += __popc(bc_);                                                         \
        }

        if (th <= 126) {
            const unsigned K = (unsigned)(127 - th) * 0x01010101u;
            // FAST_P1_UNROLL row groups per iteration: the "any survivor?" vote, the prefix scan of the per-lane counts and the
            // loop overhead are paid once per FAST_P1_UNROLL * rpi rows (at most 31 + 128 * FAST_P1_UNROLL queue entries are live)
            for (int y0 = 0; y0 < ih; y0 += FAST_P1_UNROLL * rpi) {
                unsigned f[FAST_P1_UNROLL];
                const int off0 = loff + y0 * FAST_TP;
                int mine = 0;
                unsigned any = 0;
#pragma unroll
                for (int g = 0; g < FAST_P1_UNROLL; ++g) {
                    f[g] = 0;
                    if (lrow_ok && y0 + g * rpi + lrow < ih) {
                        const uint32_t* wp = reinterpret_cast<const uint32_t*>(tile + off0 + g * rpi * FAST_TP);
                        const unsigned wm = wp[0], wl = wp[-1], wr = wp[1], wn = wp[3 * FAST_TP / 4], ws = wp[-3 * FAST_TP / 4];
                        const unsigned we = __byte_perm(wm, wr, 0x6543), ww = __byte_perm(wl, wm, 0x4321);
                        f[g] = fast_gt4(__vabsdiffu4(wn, wm), __vabsdiffu4(ws, wm), K) & fast_gt4(__vabsdiffu4(we, wm), __vabsdiffu4(ww, wm), K) & lmask;
                    }
                    any |= f[g];
                    mine += __popc(f[g]);
                }
                if (__ballot_sync(0xffffffffu, any != 0) == 0) continue;
                // queue order is irrelevant (results are keyed by position): exclusive prefix of the per-lane counts
                int incl = mine;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int u = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += u;
                }
                int pos = tail + incl - mine;
#pragma unroll
                for (int g = 0; g < FAST_P1_UNROLL; ++g) {
#pragma unroll
                    for (int b = 0; b < 4; ++b)
                        if (f[g] & (0x80u << (8 * b))) { q[pos & (FAST_QCAP - 1)] = (uint16_t)(off0 + g * rpi * FAST_TP + b); ++pos; }
                }
                tail += __shfl_sync(0xffffffffu, incl, 31);
                ORB_CHECK(tail - head <= FAST_QCAP);
                __syncwarp();
                while (tail - head >= 32) { FAST_DRAIN(true); head += 32; }
                __syncwarp();
            }
        } else {                                               // thresholds above 126: score every pixel
            for (int i = 0; i < iw * ih; i += 32) {
                const int idx = i + lane, y = idx / iw, x = idx - y * iw;
                if (idx < iw * ih) q[lane] = (uint16_t)((y + 3) * FAST_TP + cs + x);
                __syncwarp();
                head = 0;
                FAST_DRAIN(idx < iw * ih);
                __syncwarp();
            }
            head = tail = 0;
        }
        FAST_DRAIN(lane < tail - head);
#undef FAST_DRAIN
        __syncwarp();

        // ---- strict 8-neighbour maxima, emitted in row-major order
        cnt = 0;
        bool dense = ncor > FAST_CCAP;
        if (!dense) {
            uint32_t* kl = reinterpret_cast<uint32_t*>(q);     // kept list (the queue is idle now): offset << 8 | score
            int K2 = 0;
            for (int i = 0; i < ncor; i += 32) {
                bool kp = false;
                unsigned key = 0;
                if (i + lane < ncor) {
                    const int o = cl[i + lane];
                    const uint8_t* sp = smap + (o / FAST_TP - 2) * spw + o % FAST_TP - cs + 1;
                    const int v = sp[0];
                    kp = v > fast_max8(sp, spw);
                    key = ((unsigned)o << 8) | (unsigned)v;
                }
                const unsigned bk = __ballot_sync(0xffffffffu, kp);
                if (kp) { const int p = K2 + __popc(bk & lt); if (p < FAST_KCAP) kl[p] = key; }
                K2 += __popc(bk);
            }
            __syncwarp();
            if (K2 <= FAST_KCAP) {
                for (int i = lane; i < K2; i += 32) {
                    const unsigned key = kl[i];
                    int rank = 0;
                    for (int j = 0; j < K2; ++j) rank += kl[j] < key;
                    const int o = (int)(key >> 8);
                    if (rank < cell.cap) slot[rank] = ORB_PACK(o % FAST_TP - cs + cell.off_x, o / FAST_TP - 3 + cell.off_y, key & 0xFF);
                }
                cnt = K2;
            } else dense = true;
            __syncwarp();
        }
        if (dense) {                                           // ordered scan of the whole score map
            const int tot = iw * ih;
            for (int base = 0; base < tot; base += 32) {
                const int idx = base + lane;
                int v = 0, x = 0, y = 0;
                bool kp = false;
                if (idx < tot) {
                    y = __float2int_rd(((float)idx + 0.5f) * inv_iw); x = idx - y * iw;
                    const uint8_t* sp = smap + (y + 1) * spw + x + 1;
                    v = sp[0];
                    kp = v > 0 && v > fast_max8(sp, spw);
                }
                const unsigned bk = __ballot_sync(0xffffffffu, kp);
                if (kp) { const int p = cnt + __popc(bk & lt); if (p < cell.cap) slot[p] = ORB_PACK(x + cell.off_x, y + cell.off_y, v); }
                cnt += __popc(bk);
            }
        }
        if (cnt > 0 || P.min_th == P.ini_th) break;
    }
    if (lane == 0) *cnt_out = (uint32_t)min(cnt, cell.cap);
}

int orb_fast_tile_pitch(int max_cw) { return orb_fast_pitch(max_cw); }

int orb_fast_smem_bytes(int max_cw, int max_ch)
{
    if (max_cw + 3 > 64) return -1;                       // + up to 3 bytes of alignment shift
    return FAST_WARPS * (orb_fast_pitch(max_cw) * max_ch + ((orb_fast_spw(max_cw) * (max_ch - 4) + 15) & ~15));
}

int orb_launch_fast(const OrbPlanDev& P, int nframes, int smem_bytes, cudaStream_t st)
{
    dim3 grid((P.ncells + FAST_WARPS - 1) / FAST_WARPS, nframes);
    if (P.fast_tp == 48) {
        static OrbSmemOptIn optin;
        if (optin.ensure(k_fast_cells<48>, smem_bytes) != cudaSuccess) return -1;
        k_fast_cells<48><<<grid, FAST_WARPS * 32, smem_bytes, st>>>(P);
    } else {
        static OrbSmemOptIn optin;
        if (optin.ensure(k_fast_cells<64>, smem_bytes) != cudaSuccess) return -1;
        k_fast_cells<64><<<grid, FAST_WARPS * 32, smem_bytes, st>>>(P);
    }
    return 1;
}

// =================================================================================================
// DistributeOctTree as array passes; one CTA per (level, frame).
//
// The reference keeps a std::list of nodes and splits them with push_front (src/ORBextractor.cc:
// 596-739).  Here a node IS its list position: every pass computes, for all nodes at once, the new
// list  reverse(children in creation order) ++ (unsplit nodes in old order)  with prefix scans, and
// remaps each key's node index.  Only nodes created by the previous pass can hold more than one key,
// so the expandable set is {i < M : cnt[i] > 1} (M = number of nodes the last pass created).
//   full sweep   (:596-667)  split every expandable node, in list order
//   budgeted pass(:675-739)  order expandable nodes by (key count desc, newest first) -- the
//                            reference's sort by (size, node address) read from the back, with the
//                            address tie pinned to creation order -- and split the prefix that
//                            keeps list size < N before each split (the `break` at :732-733)
// Selection (:743-762): per node the key with the highest response, first in input order on ties.
// Keys and node tables live in shared memory when they fit, else in a global scratch area.
// =================================================================================================
#ifndef OCT_THREADS
#define OCT_THREADS 256
#endif
#define OCT_NODE_BYTES 64     // per-node bytes in the node pool (see carve below)
#ifndef OCT_AGG_NODES
#define OCT_AGG_NODES 48      // sweeps over at most this many fresh nodes count children with warp-aggregated atomics
#endif

struct OctPools {
    uint32_t* kxy; uint16_t* knode; uint8_t* kq;                 // per key
    short4* bnd[2]; int* cnt[2]; int* cc; int* A; int* B; int* ord; int* X; int* np;   // per node
};

__device__ __forceinline__ void oct_carve(OctPools& o, uint8_t* keypool, int kcap, uint8_t* nodepool, int ncap)
{
    o.kxy = reinterpret_cast<uint32_t*>(keypool);
    o.knode = reinterpret_cast<uint16_t*>(keypool + 4 * (size_t)kcap);
    o.kq = keypool + 6 * (size_t)kcap;
    uint8_t* p = nodepool;
    o.bnd[0] = reinterpret_cast<short4*>(p); p += 8 * (size_t)ncap;
    o.bnd[1] = reinterpret_cast<short4*>(p); p += 8 * (size_t)ncap;
    o.cnt[0] = reinterpret_cast<int*>(p); p += 4 * (size_t)ncap;
    o.cnt[1] = reinterpret_cast<int*>(p); p += 4 * (size_t)ncap;
    o.cc = reinterpret_cast<int*>(p); p += 16 * (size_t)ncap;
    o.A = reinterpret_cast<int*>(p); p += 4 * (size_t)ncap;
    o.B = reinterpret_cast<int*>(p); p += 4 * (size_t)ncap;
    o.ord = reinterpret_cast<int*>(p); p += 4 * (size_t)ncap;
    o.X = reinterpret_cast<int*>(p); p += 4 * (size_t)ncap + 16;
    o.np = reinterpret_cast<int*>(p);
}

// in-place exclusive prefix sum of a[0..n) by the whole CTA; returns the total to every thread.
__device__ int block_excl_scan(int* a, int n, int* s_part /* 33 ints of shared memory */)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ipt = (n + OCT_THREADS - 1) / OCT_THREADS;
    const int lo = min(tid * ipt, n), hi = min(lo + ipt, n);
    __syncthreads();
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += a[i];
    int inc = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) s_part[warp] = inc;
    __syncthreads();
    // every warp scans the OCT_THREADS / 32 warp totals itself (one barrier less than a scan by warp 0 alone)
    const int w = lane < OCT_THREADS / 32 ? s_part[lane] : 0;
    int winc = w;
#pragma unroll
    for (int d = 1; d < OCT_THREADS / 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, winc, d);
        if (lane >= d) winc += t;
    }
    const int total = __shfl_sync(0xffffffffu, winc, OCT_THREADS / 32 - 1);
    int run = __shfl_sync(0xffffffffu, winc - w, warp) + inc - sum;
    for (int i = lo; i < hi; ++i) { const int v = a[i]; a[i] = run; run += v; }
    __syncthreads();
    return total;
}

// counter[idx] += 1 for every active lane, with ONE shared-memory atomic per distinct idx in the warp: while the tree has
// few nodes thousands of keys hit a handful of counters and per-lane atomics on one address serialise 32-fold
__device__ __forceinline__ void oct_count(int* counter, int idx, bool agg)
{
    if (agg) {
        const unsigned peers = __match_any_sync(__activemask(), idx);
        if ((peers & ((1u << (threadIdx.x & 31)) - 1u)) == 0) atomicAdd(&counter[idx], __popc(peers));
    } else atomicAdd(&counter[idx], 1);
}

// Runs the octree on keys o.kxy[0..nK); writes the survivors (packed) to out[0..count) in list order.
__device__ int octree_run(OctPools& o, int nK, int node_cap, int W, int H, int N, int n_ini, float hx,
                          uint32_t* out, int out_cap, int* s_part, int* s_misc)
{
    const int tid = threadIdx.x;
    (void)W;
    // ---- roots (:548-587)
    for (int i = tid; i < n_ini; i += OCT_THREADS) {
        o.bnd[0][i] = make_short4((short)(int)__fmul_rn(hx, (float)i), (short)(int)__fmul_rn(hx, (float)(i + 1)), 0, (short)H);
        o.cnt[0][i] = 0;
    }
    __syncthreads();
    for (int k = tid; k < nK; k += OCT_THREADS) {
        int r = (int)__fdiv_rn((float)ORB_PX(o.kxy[k]), hx);
        r = min(r, n_ini - 1);
        ORB_CHECK(r >= 0 && n_ini <= node_cap);
        o.knode[k] = (uint16_t)r;
        oct_count(o.cnt[0], r, true);
    }
    __syncthreads();
    for (int i = tid; i < n_ini; i += OCT_THREADS) o.A[i] = o.cnt[0][i] > 0;
    int n = block_excl_scan(o.A, n_ini, s_part);
    for (int i = tid; i < n_ini; i += OCT_THREADS)
        if (o.cnt[0][i] > 0) { o.bnd[1][o.A[i]] = o.bnd[0][i]; o.cnt[1][o.A[i]] = o.cnt[0][i]; }
    for (int k = tid; k < nK; k += OCT_THREADS) o.knode[k] = (uint16_t)o.A[o.knode[k]];
    __syncthreads();

    int cur = 1, M = n;
    bool phaseB = false;
    while (true) {
        short4* bnd = o.bnd[cur]; int* cnt = o.cnt[cur];
        short4* bnd2 = o.bnd[cur ^ 1]; int* cnt2 = o.cnt[cur ^ 1];
        // ---- child key counts of every expandable node (DivideNode, :483-539)
        for (int i = tid; i < 4 * M; i += OCT_THREADS) o.cc[i] = 0;
        if (tid == 0) { s_misc[0] = 0; s_misc[1] = 0; }
        __syncthreads();
        for (int k = tid; k < nK; k += OCT_THREADS) {
            const int nd = o.knode[k];
            if (nd < M && cnt[nd] > 1) {
                const short4 b = bnd[nd];
                const int mx = b.x + ((b.y - b.x + 1) >> 1), my = b.z + ((b.w - b.z + 1) >> 1);
                const uint32_t p = o.kxy[k];
                const int q = (ORB_PX(p) < mx ? 0 : 1) + (ORB_PY(p) < my ? 0 : 2);
                o.kq[k] = (uint8_t)q;
                oct_count(o.cc, 4 * nd + q, M <= OCT_AGG_NODES);
            }
        }
        __syncthreads();
        int nE, nCommit, Mnew, newSize;
        if (!phaseB && n <= 16384) {
            // full sweep: every expandable node is split, in list order, so the rank of a node, the list growth before it and
            // the number of committed nodes before ANY node come out of ONE scan over (expandable << 16 | nonempty children - 1)
            // instead of three scans (4 barriers each)
            for (int i = tid; i < n; i += OCT_THREADS) {
                const bool e = (i < M && cnt[i] > 1);
                const int g = e ? (o.cc[4 * i] > 0) + (o.cc[4 * i + 1] > 0) + (o.cc[4 * i + 2] > 0) + (o.cc[4 * i + 3] > 0) - 1 : 0;
                o.A[i] = ((int)e << 16) | g;
            }
            const int total = block_excl_scan(o.A, n, s_part);
            nE = total >> 16;
            for (int i = tid; i < n; i += OCT_THREADS) {
                const int v = o.A[i], rank = v >> 16;
                const bool e = (i < M && cnt[i] > 1);
                o.B[i] = e ? rank : -1;
                if (e) { o.ord[rank] = i; o.X[rank] = v & 0xFFFF; }
                o.A[i] = rank;                                // committed nodes before i
            }
            if (tid == 0) { o.X[nE] = total & 0xFFFF; s_misc[2] = 0; }
            __syncthreads();
            nCommit = nE;
            Mnew = o.X[nCommit] + nCommit;
            newSize = n - nCommit + Mnew;
            if (newSize > node_cap || newSize > out_cap) return -1;  // cannot happen (see node_cap derivation)
        } else {
            // ---- processing order of the expandable nodes
            if (!phaseB) {
                for (int i = tid; i < n; i += OCT_THREADS) o.A[i] = (i < M && cnt[i] > 1);
                nE = block_excl_scan(o.A, n, s_part);
                for (int i = tid; i < n; i += OCT_THREADS) {
                    const bool e = (i < M && cnt[i] > 1);
                    o.B[i] = e ? o.A[i] : -1;
                    if (e) o.ord[o.A[i]] = i;
                }
            } else {
                int local = 0;
                for (int i = tid; i < n; i += OCT_THREADS) {
                    int rank = -1;
                    if (i < M && cnt[i] > 1) {
                        const int ci = cnt[i];
                        rank = 0;
                        for (int j = 0; j < M; ++j) {
                            const int cj = cnt[j];
                            rank += (cj > 1) && (cj > ci || (cj == ci && j < i));
                        }
                        o.ord[rank] = i;
                        ++local;
                    }
                    o.B[i] = rank;
                }
                if (local) atomicAdd(&s_misc[0], local);
                __syncthreads();
                nE = s_misc[0];
            }
            __syncthreads();
            // ---- X[r] = list growth caused by the parents processed before r (nonempty children - 1)
            for (int r = tid; r < nE; r += OCT_THREADS) {
                const int p = o.ord[r];
                o.X[r] = (o.cc[4 * p] > 0) + (o.cc[4 * p + 1] > 0) + (o.cc[4 * p + 2] > 0) + (o.cc[4 * p + 3] > 0) - 1;
            }
            const int growth_all = block_excl_scan(o.X, nE, s_part);
            if (tid == 0) o.X[nE] = growth_all;
            __syncthreads();
            // committed prefix: everything in a full sweep; in a budgeted pass, parents reached while size < N
            nCommit = nE;
            if (phaseB) {
                int local = 0;
                for (int r = tid; r < nE; r += OCT_THREADS) local += (n + o.X[r] < N);
                if (local) atomicAdd(&s_misc[1], local);
                __syncthreads();
                nCommit = s_misc[1];
            }
            Mnew = o.X[nCommit] + nCommit;
            newSize = n - nCommit + Mnew;
            if (newSize > node_cap || newSize > out_cap) return -1;      // cannot happen (see node_cap derivation)
            // ---- unsplit nodes keep their order behind the new ones
            for (int i = tid; i < n; i += OCT_THREADS) o.A[i] = (o.B[i] >= 0 && o.B[i] < nCommit);
            block_excl_scan(o.A, n, s_part);
            if (tid == 0) s_misc[2] = 0;
            __syncthreads();
        }
        int to_expand = 0;
        for (int i = tid; i < n; i += OCT_THREADS) {
            const int r = o.B[i];
            if (r >= 0 && r < nCommit) {
                int c = o.X[r] + r;                       // creation index of this parent's first child
                const short4 b = bnd[i];
                const int mx = b.x + ((b.y - b.x + 1) >> 1), my = b.z + ((b.w - b.z + 1) >> 1);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int kc = o.cc[4 * i + q];
                    if (kc > 0) {
                        const int pos = Mnew - 1 - c; ++c;
                        ORB_CHECK(pos >= 0 && pos < newSize && newSize <= node_cap);
                        bnd2[pos] = make_short4((q & 1) ? (short)mx : b.x, (q & 1) ? b.y : (short)mx,
                                                (q & 2) ? (short)my : b.z, (q & 2) ? b.w : (short)my);
                        cnt2[pos] = kc;
                        o.cc[4 * i + q] = pos;
                        to_expand += (kc > 1);
                    }
                }
            } else {
                const int pos = Mnew + i - o.A[i];
                ORB_CHECK(pos >= Mnew && pos < newSize);
                o.np[i] = pos;
                bnd2[pos] = bnd[i];
                cnt2[pos] = cnt[i];
            }
        }
        if (to_expand) atomicAdd(&s_misc[2], to_expand);
        __syncthreads();
        const int prev_n = n; (void)prev_n;
        for (int k = tid; k < nK; k += OCT_THREADS) {
            const int nd = o.knode[k];
            ORB_CHECK(nd < prev_n);
            const int r = o.B[nd];
            o.knode[k] = (uint16_t)((r >= 0 && r < nCommit) ? o.cc[4 * nd + o.kq[k]] : o.np[nd]);
        }
        const int nToExpand = s_misc[2];
        __syncthreads();
        const int prev = n;
        n = newSize; M = Mnew; cur ^= 1;
        if (n >= N || n == prev) break;                   // :671-674, :736-737
        if (!phaseB && n + 3 * nToExpand > N) phaseB = true;   // :675
    }
    // ---- best key per node (:743-762)
    unsigned* best = reinterpret_cast<unsigned*>(o.A);
    for (int i = tid; i < n; i += OCT_THREADS) best[i] = 0;
    __syncthreads();
    for (int k = tid; k < nK; k += OCT_THREADS)
        atomicMax(&best[o.knode[k]], ((unsigned)ORB_PS(o.kxy[k]) << 24) | (0xFFFFFFu - (unsigned)k));
    __syncthreads();
    for (int i = tid; i < n; i += OCT_THREADS) out[i] = o.kxy[0xFFFFFFu - (best[i] & 0xFFFFFFu)];
    return n;
}

__device__ __forceinline__ size_t oct_key_pool_bytes(int kcap) { return ((size_t)7 * kcap + 15) & ~(size_t)15; }

__global__ void __launch_bounds__(OCT_THREADS) k_octree(OrbPlanDev P, int key_smem_cap, int node_smem_cap)
{
    extern __shared__ __align__(16) uint8_t oct_smem[];
    __shared__ int s_part[33];
    __shared__ int s_misc[4];
    const int l = blockIdx.x, frame = blockIdx.y, tid = threadIdx.x;
    const OrbLevel& L = P.lv[l];
    int* selcnt = P.selcnt + (long long)frame * ORB_MAX_LEVELS + l;
    int* candcnt = P.candcnt + (long long)frame * ORB_MAX_LEVELS + l;
    if (L.ncells == 0 || L.n_ini < 1) { if (tid == 0) { *selcnt = 0; *candcnt = 0; } return; }

    // candidate list order = cell row-major, then in-cell row-major (:791-828): exclusive scan of the
    // per-cell survivor counts gives every cell's offset (kept in place for the debug accessors)
    int* cbase = reinterpret_cast<int*>(P.cellcnt + (long long)frame * P.ncells + L.cell0);
    const int nK = block_excl_scan(cbase, L.ncells, s_part);
    if (tid == 0) *candcnt = nK;

    const bool keys_in_smem = nK <= key_smem_cap, nodes_in_smem = L.node_cap <= node_smem_cap;
    uint8_t* scratch = P.scratch + (long long)frame * P.scratch_frame_bytes;
    OctPools o;
    oct_carve(o, keys_in_smem ? oct_smem : scratch + L.key_scratch_off, keys_in_smem ? key_smem_cap : L.slot_total,
              nodes_in_smem ? oct_smem + oct_key_pool_bytes(key_smem_cap) : scratch + L.node_scratch_off, L.node_cap);

    const uint32_t* slots = P.slots + (long long)frame * P.slot_frame_words + L.slot_off;
    const OrbCell* cells = P.cells + L.cell0;
    // thread = one candidate: its cell is the last one whose offset is <= k (binary search over the scanned counts, which
    // sit in L1).  A warp-per-cell walk made every warp chase cell table -> slot through ~35 dependent round trips.
    for (int k = tid; k < nK; k += OCT_THREADS) {
        int lo = 0, hi = L.ncells;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (cbase[mid] <= k) lo = mid + 1; else hi = mid;
        }
        const int c = lo - 1;
        ORB_CHECK(c >= 0 && c < L.ncells && k - cbase[c] < cells[c].cap && (keys_in_smem ? k < key_smem_cap : k < L.slot_total));
        o.kxy[k] = slots[cells[c].slot_off + (k - cbase[c])];
    }
    __syncthreads();

    uint32_t* out = P.sel + (long long)frame * P.sel_total + L.sel_off;
    const int n = octree_run(o, nK, L.node_cap, L.W, L.H, L.nfeat, L.n_ini, L.hx, out, L.sel_cap, s_part, s_misc);
    if (tid == 0) *selcnt = n;
}

int orb_launch_octree(const OrbPlanDev& P, int nframes, int key_smem_cap, int node_smem_cap, cudaStream_t st)
{
    const int smem = (int)((((size_t)7 * key_smem_cap + 15) & ~(size_t)15) + (size_t)OCT_NODE_BYTES * node_smem_cap + 64);
    static OrbSmemOptIn optin;
    if (optin.ensure(k_octree, smem) != cudaSuccess) return -1;
    k_octree<<<dim3(P.nlevels, nframes), OCT_THREADS, smem, st>>>(P, key_smem_cap, node_smem_cap);
    return 1;
}

__global__ void __launch_bounds__(OCT_THREADS) k_octree_single(const uint32_t* keys, int nK, int W, int H, int N, int n_ini,
                                                               float hx, int node_cap, uint8_t* scratch, uint32_t* out,
                                                               int* count, int key_smem_cap, int node_smem_cap)
{
    extern __shared__ __align__(16) uint8_t oct_smem[];
    __shared__ int s_part[33];
    __shared__ int s_misc[4];
    const bool keys_in_smem = nK <= key_smem_cap, nodes_in_smem = node_cap <= node_smem_cap;
    OctPools o;
    const size_t key_bytes = ((size_t)7 * nK + 15) & ~(size_t)15;
    oct_carve(o, keys_in_smem ? oct_smem : scratch, keys_in_smem ? key_smem_cap : nK,
              nodes_in_smem ? oct_smem + oct_key_pool_bytes(key_smem_cap) : scratch + key_bytes, node_cap);
    for (int k = threadIdx.x; k < nK; k += OCT_THREADS) o.kxy[k] = keys[k];
    __syncthreads();
    const int n = octree_run(o, nK, node_cap, W, H, N, n_ini, hx, out, node_cap, s_part, s_misc);
    if (threadIdx.x == 0) *count = n;
}

int orb_launch_octree_single(const uint32_t* d_keys, int n, int W, int H, int N, int n_ini, float hx, int node_cap,
                             uint8_t* d_scratch, uint32_t* d_out, int* d_count, int key_smem_cap, int node_smem_cap,
                             cudaStream_t st)
{
    const int smem = (int)((((size_t)7 * key_smem_cap + 15) & ~(size_t)15) + (size_t)OCT_NODE_BYTES * node_smem_cap + 64);
    static OrbSmemOptIn optin;
    if (optin.ensure(k_octree_single, smem) != cudaSuccess) return -1;
    k_octree_single<<<1, OCT_THREADS, smem, st>>>(d_keys, n, W, H, N, n_ini, hx, node_cap, d_scratch, d_out, d_count,
                                                  key_smem_cap, node_smem_cap);
    return 1;
}

// =================================================================================================
// GaussianBlur 7x7, sigma 2, REFLECT_101, in OpenCV's unsigned fixed point: 8.8 taps
// [18,34,48,56,48,34,18], horizontal sums in 16 bits, vertical sums in 32 bits, (v + 2^15) >> 16.
// One thread = one 4-pixel word column x ORB_BLUR_ROWS output rows: the horizontal sums of the last seven input rows
// live in registers (no shared memory), three aligned word loads per input row issued one row ahead; the row-end columns
// (REFLECT_101) are separate work items on a byte path.  The work list (pl.d_tiles) is built with the plan.
// =================================================================================================
#define BLUR_R ORB_BLUR_ROWS   // output rows per thread (4 x the 7-row register ring)
#define BLUR_THREADS ORB_BLUR_THREADS

__device__ __forceinline__ int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

// Input row number with REFLECT_101.  From 8 rows up one reflection per side is exact for every row a strip USES
// (-3 <= i <= hgt + 2); rows past the strip's last output are loaded but never used and are only clamped into the image.
__device__ __forceinline__ int blur_row(int i, int hgt)
{
    if (hgt < 8) return reflect101(i, hgt);
    const int a = i < 0 ? -i : i;
    const int b = a >= hgt ? 2 * (hgt - 1) - a : a;
    return min(max(b, 0), hgt - 1);
}

// The 12-byte window (wl, wm, wr) of one input row around word column wc: bytes I(4wc - 4 .. 4wc + 7).
// WORDS: three aligned 32-bit loads (the whole window lies inside the row); otherwise byte loads with
// REFLECT_101 (row ends, or rows that are not 4-byte aligned).
// Right-edge columns on aligned rows (mode 2): the window's bytes beyond the row end are REFLECT_101 copies of bytes that
// lie in the four aligned words wc - 2 .. wc + 1, and the four source bytes of each output word always fall into two
// ADJACENT words of those, so an output word is one PRMT of a word pair; which pair (0..2) and the selector are fixed per
// thread (blur_edge_sel) -- four word loads and two PRMT per row instead of ten byte loads and their reassembly.
struct BlurEdgeSel { unsigned sel_m, sel_r; int pair_m, pair_r, w3; };

__device__ __forceinline__ BlurEdgeSel blur_edge_sel(int wc, int w, int pitch)
{
    BlurEdgeSel e;
    unsigned sel[2]; int pr[2];
#pragma unroll
    for (int X = 1; X <= 2; ++X) {
        int o[4], mn = 16;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int i = 4 * wc - 4 + 4 * X + b;
            const int sidx = i >= w ? 2 * (w - 1) - i : i;              // one reflection: w >= 16 and i < w + 8
            o[b] = sidx - 4 * (wc - 2);
            mn = min(mn, o[b]);
        }
        const int pi = min(mn >> 2, 2);
        unsigned sl = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) sl |= (unsigned)(o[b] - 4 * pi) << (4 * b);
        sel[X - 1] = sl; pr[X - 1] = pi;
    }
    e.sel_m = sel[0]; e.sel_r = sel[1]; e.pair_m = pr[0]; e.pair_r = pr[1];
    e.w3 = min(wc + 1, (pitch >> 2) - 1) - wc;                        // word wc + 1 if it lies inside the pitch, else wc again
    return e;
}

template <int MODE>
__device__ __forceinline__ uint3 blur_load(const uint8_t* __restrict__ rowp, int wc, int w, const BlurEdgeSel& es)
{
    unsigned wl, wm, wr;
    const bool WORDS = MODE == 1;
    if (MODE == 2) {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(rowp) + wc;
        const unsigned q0 = __ldg(p - 2), q1 = __ldg(p - 1), q2 = __ldg(p), q3 = __ldg(p + es.w3);
        wl = q1;
        wm = __byte_perm(es.pair_m == 0 ? q0 : (es.pair_m == 1 ? q1 : q2), es.pair_m == 0 ? q1 : (es.pair_m == 1 ? q2 : q3), es.sel_m);
        wr = __byte_perm(es.pair_r == 0 ? q0 : (es.pair_r == 1 ? q1 : q2), es.pair_r == 0 ? q1 : (es.pair_r == 1 ? q2 : q3), es.sel_r);
    } else if (WORDS) {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(rowp) + wc;
        wm = __ldg(p);
        wr = __ldg(p + 1);
        wl = wc ? __ldg(p - 1) : __byte_perm(wm, 0, 0x1233);   // REFLECT_101: I(-1..-3) = I(1..3)
    } else {
        wl = wm = wr = 0;
        const int x0 = 4 * wc - 4, xr = 2 * (w - 1);
        const bool once = w >= 16;                             // one reflection is enough: -w < i < 2w - 1
        // five column numbers, then their five loads together, twice (a branch between two loads serialises them; ten at
        // once cost the word path two spilled registers)
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            int xi[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const int i = x0 + 1 + 5 * g + k;
                xi[k] = once ? (i < 0 ? -i : (i >= w ? xr - i : i)) : reflect101(i, w);
            }
            unsigned v[5];
#pragma unroll
            for (int k = 0; k < 5; ++k) v[k] = __ldg(rowp + xi[k]);
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const int kk = 1 + 5 * g + k;
                if (kk < 4) wl |= v[k] << (8 * kk); else if (kk < 8) wm |= v[k] << (8 * (kk - 4)); else wr |= v[k] << (8 * (kk - 8));
            }
        }
    }
    return make_uint3(wl, wm, wr);
}

// horizontal pass for the 4 pixels of the word column: h[p] = sum_i k_i * I(4wc + p + i - 3).
// window bytes b[k] = I(4wc - 3 + k): a_p = b[p..p+3], c_p = b[p+4..p+7]; taps (18,34,48,56 | 48,34,18,0) by IDP.4A
__device__ __forceinline__ void blur_hsum(const uint3 r, unsigned h[4])
{
    const unsigned K0 = 18u | (34u << 8) | (48u << 16) | (56u << 24), K1 = 48u | (34u << 8) | (18u << 16);
    h[0] = __dp4a(__funnelshift_r(r.x, r.y, 8), K0, __dp4a(__funnelshift_r(r.y, r.z, 8), K1, 0u));
    h[1] = __dp4a(__funnelshift_r(r.x, r.y, 16), K0, __dp4a(__funnelshift_r(r.y, r.z, 16), K1, 0u));
    h[2] = __dp4a(__funnelshift_r(r.x, r.y, 24), K0, __dp4a(__funnelshift_r(r.y, r.z, 24), K1, 0u));
    h[3] = __dp4a(r.y, K0, __dp4a(r.z, K1, 0u));
}

// One thread = one 4-pixel word column x BLUR_R output rows.  The horizontal sums of the last 7 input rows
// live in registers (28 values); every new input row costs three aligned 32-bit loads (L1 serves the
// overlap with the neighbouring threads), issued BLUR_AHEAD rows before their use; every output row is one
// aligned 32-bit store.  No shared memory.
#ifndef BLUR_AHEAD
#define BLUR_AHEAD 1
#endif
#ifndef BLUR_VDP2A
#define BLUR_VDP2A 1          // vertical pass with IDP.2A on row pairs (0: adds + IMADs on single rows)
#endif
#ifndef BLUR_EDGE_WORDS
#define BLUR_EDGE_WORDS 1     // right-edge columns of aligned rows on the word + PRMT path (0: byte path)
#endif
template <int MODE, bool INNER>
__device__ __forceinline__ void blur_strip(const uint8_t* __restrict__ src, int sp, uint8_t* __restrict__ dst, int dp,
                                           int w, int hgt, int wc, int y0)
{
    BlurEdgeSel es = {0, 0, 0, 0, 0};
    if (MODE == 2) es = blur_edge_sel(wc, w, sp);
    // input row y0 - 3 + k; rows are REFLECT_101-mapped only in the first / last strip; INNER: known at compile time not
    // to need it (the word path, where the row arithmetic is a visible share of the instructions)
    const bool inner = INNER || (y0 >= 3 && y0 + BLUR_R + 3 + BLUR_AHEAD <= hgt);
#if !BLUR_VDP2A
    unsigned ring[7][4];
#endif
    // the row numbers of the 6 + BLUR_AHEAD start-up rows first (reflect101 is a loop: computed between the loads it put a
    // branch after every row and serialised their round trips), then all their loads, then the sums
    int yy0[6 + BLUR_AHEAD];
#pragma unroll
    for (int k = 0; k < 6 + BLUR_AHEAD; ++k) yy0[k] = inner ? y0 - 3 + k : blur_row(y0 - 3 + k, hgt);
    uint3 first[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) first[k] = blur_load<MODE>(src + (size_t)yy0[k] * sp, wc, w, es);
    uint3 ahead[BLUR_AHEAD];                                   // rows y0 + 3 .. y0 + 3 + BLUR_AHEAD - 1, in flight
#pragma unroll
    for (int a = 0; a < BLUR_AHEAD; ++a) ahead[a] = blur_load<MODE>(src + (size_t)yy0[6 + a] * sp, wc, w, es);
#if BLUR_VDP2A
    // vertical pass on PAIRS of adjacent rows: pr[m % 7][p] = h_m | h_{m+1} << 16 (h < 2^16), so that
    // v_j = (18, 34).(h_j, h_j+1) + (48, 56).(h_j+2, h_j+3) + (48, 34).(h_j+4, h_j+5) + 18 h_j+6 is three IDP.2A and one IMAD
    // per pixel, plus one PRMT per pixel and row to form the newest pair, instead of three adds and four IMADs
    const unsigned K01 = 18u | (34u << 8), K23 = 48u | (56u << 8), K45 = 48u | (34u << 8);
    unsigned pr[7][4], hprev[4];
    {
        unsigned h0[4], h1[4];
        blur_hsum(first[0], h0);
#pragma unroll
        for (int k = 1; k < 6; ++k) {
            blur_hsum(first[k], h1);
#pragma unroll
            for (int p = 0; p < 4; ++p) { pr[k - 1][p] = __byte_perm(h0[p], h1[p], 0x5410); h0[p] = h1[p]; }
        }
#pragma unroll
        for (int p = 0; p < 4; ++p) hprev[p] = h0[p];
    }
    for (int yb = 0; yb < BLUR_R; yb += 7) {
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int y = y0 + yb + j;
            const uint3 cur = ahead[0];
#pragma unroll
            for (int a = 0; a + 1 < BLUR_AHEAD; ++a) ahead[a] = ahead[a + 1];
            {
                const int yn = y + 3 + BLUR_AHEAD;             // rows past the strip are loaded but never used
                const int yy = inner ? yn : blur_row(yn, hgt);
                ahead[BLUR_AHEAD - 1] = blur_load<MODE>(src + (size_t)yy * sp, wc, w, es);
            }
            unsigned hn[4], v[4];
            blur_hsum(cur, hn);
#pragma unroll
            for (int p = 0; p < 4; ++p) {
                pr[(j + 5) % 7][p] = __byte_perm(hprev[p], hn[p], 0x5410);
                hprev[p] = hn[p];
                v[p] = __dp2a_lo(pr[j % 7][p], K01, __dp2a_lo(pr[(j + 2) % 7][p], K23, __dp2a_lo(pr[(j + 4) % 7][p], K45, 18u * hn[p] + 32768u)));
            }
            const uint32_t out = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
            if (y < hgt) *reinterpret_cast<uint32_t*>(dst + (size_t)y * dp + 4 * wc) = out;
        }
    }
}
#else
#pragma unroll
    for (int k = 0; k < 6; ++k) blur_hsum(first[k], ring[k]);
    for (int yb = 0; yb < BLUR_R; yb += 7) {
#pragma unroll
        for (int j = 0; j < 7; ++j) {
            const int y = y0 + yb + j;
            const uint3 cur = ahead[0];
#pragma unroll
            for (int a = 0; a + 1 < BLUR_AHEAD; ++a) ahead[a] = ahead[a + 1];
            {
                const int yn = y + 3 + BLUR_AHEAD;             // rows past the strip are loaded but never used
                const int yy = inner ? yn : blur_row(yn, hgt);
                ahead[BLUR_AHEAD - 1] = blur_load<MODE>(src + (size_t)yy * sp, wc, w, es);
            }
            blur_hsum(cur, ring[(j + 6) % 7]);
            unsigned v[4];
#pragma unroll
            for (int p = 0; p < 4; ++p)
                v[p] = 18u * (ring[j % 7][p] + ring[(j + 6) % 7][p]) + 34u * (ring[(j + 1) % 7][p] + ring[(j + 5) % 7][p]) +
                       48u * (ring[(j + 2) % 7][p] + ring[(j + 4) % 7][p]) + 56u * ring[(j + 3) % 7][p] + 32768u;
            // byte 2 of each sum = (v + 2^15) >> 16 (v < 2^24)
            const uint32_t out = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
            if (y < hgt) *reinterpret_cast<uint32_t*>(dst + (size_t)y * dp + 4 * wc) = out;
        }
    }
}
#endif

// the rare paths as real calls: inlined beside the word path they cost it spilled registers under the kernel's 56-register cap
__device__ __noinline__ void blur_strip_edge(const uint8_t* src, int sp, uint8_t* dst, int dp, int w, int hgt, int wc, int y0)
{
    blur_strip<2, false>(src, sp, dst, dp, w, hgt, wc, y0);
}
__device__ __noinline__ void blur_strip_bytes(const uint8_t* src, int sp, uint8_t* dst, int dp, int w, int hgt, int wc, int y0)
{
    blur_strip<0, false>(src, sp, dst, dp, w, hgt, wc, y0);
}

// Work list entry: (level, first task, kind).  kind 0 = interior word columns [0, nfast) on the aligned
// word path; kind 1 = the remaining columns (row ends; or every column when the level's rows are not
// 4-byte aligned) on the byte path.  Splitting them keeps every warp on one path.
#ifndef BLUR_MINB
#define BLUR_MINB 9
#endif
__global__ void __launch_bounds__(BLUR_THREADS, BLUR_MINB) k_blur(OrbPlanDev P, const int4* __restrict__ tiles)
{
    const int4 t = tiles[blockIdx.x];
    const int l = t.x, frame = blockIdx.y;
    const OrbLevel& L = P.lv[l];
    const int w = L.w, hgt = L.h, nwc = (w + 3) >> 2;
    const uint8_t* __restrict__ src = orb_level_ptr(P, l, frame);
    const int sp = orb_level_pitch(P, l);
    uint8_t* __restrict__ dst = P.blur + (long long)frame * P.blur_frame_bytes + L.blur_off;
    const bool aligned = ((sp & 3) == 0) && ((reinterpret_cast<uintptr_t>(src) & 3) == 0);
    const int nfast = (aligned && w >= 8) ? (w - 4) >> 2 : 0;  // word columns with 4wc + 8 <= w
    const int task = t.y + threadIdx.x;
    if (t.z == 0) {
        if (nfast == 0) return;
        const int strip = task / nfast, wc = task - strip * nfast;
        if (strip * BLUR_R >= hgt) return;
        const int y0 = strip * BLUR_R;
        blur_strip<1, false>(src, sp, dst, L.pitch, w, hgt, wc, y0);
    } else {
        const int nslow = nwc - nfast;
        const int strip = task / nslow, wc = nfast + task - strip * nslow;
        if (strip * BLUR_R >= hgt) return;
        if (BLUR_EDGE_WORDS && aligned && w >= 16 && wc >= 2) blur_strip_edge(src, sp, dst, L.pitch, w, hgt, wc, strip * BLUR_R);
        else blur_strip_bytes(src, sp, dst, L.pitch, w, hgt, wc, strip * BLUR_R);
    }
}

int orb_launch_blur(const OrbPlanDev& P, int nframes, const int4* tiles, int ntiles, cudaStream_t st)
{
    k_blur<<<dim3(ntiles, nframes), BLUR_THREADS, 0, st>>>(P, tiles);
    return 1;
}

// =================================================================================================
// Orientation + descriptor + output record; one warp per octree survivor.
//   IC_Angle (:79-106): integer moments over the 749-pixel disc of the UNBLURRED level, lane = column;
//   cv::fastAtan2: degree-7 polynomial in plain float mul/add (no FMA), pinned to OpenCV 4.13;
//   steered BRIEF (:110-149) on the BLURRED level: lane i produces descriptor byte i;
//   cos/sin follow glibc 2.39's sincosf (x86-64 FMA variant) operation by operation in double,
//   so the rounded sample offsets agree with the host for every float angle.
// =================================================================================================
__device__ __forceinline__ float orb_fast_atan2(float y, float x)
{
    const float P1 = 57.283626556396484f, P3 = -18.66744613647461f, P5 = 8.914000511169434f, P7 = -2.539724588394165f;
    const float eps = 2.2204460492503131e-16f;
    const float ax = fabsf(x), ay = fabsf(y);
    float a;
    if (ax >= ay) {
        const float c = __fdiv_rn(ay, __fadd_rn(ax, eps)), c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(P7, c2), P5), c2), P3), c2), P1), c);
    } else {
        const float c = __fdiv_rn(ax, __fadd_rn(ay, eps)), c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(P7, c2), P5), c2), P3), c2), P1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

__device__ __forceinline__ void orb_sincosf(float y, float* sinp, float* cosp)
{
    const double HPI_INV = 0x1.45f306dc9c883p+23, HPI = 0x1.921fb54442d18p+0;
    const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10,
                 C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    const unsigned top12 = (__float_as_uint(y) >> 20) & 0x7ffu;
    double x = (double)y, xs, flip = 1.0;
    int n = 0;
    if (top12 < 0x3f4u) {
        if (top12 < 0x398u) { *sinp = y; *cosp = 1.0f; return; }
        xs = x;
    } else {
        const double r = __dmul_rn(x, HPI_INV);
        n = (__double2int_rz(r) + 0x800000) >> 24;
        x = __fma_rn(-(double)n, HPI, x);
        xs = ((n & 3) == 1 || (n & 3) == 2) ? -x : x;
        if (n & 2) flip = -1.0;
    }
    const double x2 = __dmul_rn(x, x);
    const double sp = __fma_rn(x2, S3, S2);
    const double cp = __fma_rn(x2, flip * C4, flip * C3);
    const double x3 = __dmul_rn(x2, xs), x4 = __dmul_rn(x2, x2);
    const double x5 = __dmul_rn(x2, x3), x6 = __dmul_rn(x2, x4);
    const double c1 = __fma_rn(x2, flip * C1, flip * C0);
    const double s = __fma_rn(x3, S1, xs);
    const double c = __fma_rn(x4, flip * C2, c1);
    const float fs = __double2float_rn(__fma_rn(sp, x5, s));
    const float fc = __double2float_rn(__fma_rn(cp, x6, c));
    if (n & 1) { *sinp = fc; *cosp = fs; } else { *sinp = fs; *cosp = fc; }
}

#ifndef DESC_PER_WARP
#define DESC_PER_WARP 2
#endif
#ifndef DESC_MINB
#define DESC_MINB 6
#endif
struct OrbKeypointRec { float x, y, size, angle, response; int octave, class_id; };

__global__ void __launch_bounds__(256, DESC_MINB) k_describe(OrbPlanDev P)
{
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, frame = blockIdx.y;
    const int* selcnt = P.selcnt + (long long)frame * ORB_MAX_LEVELS;
    // per-level survivor counts, one per lane; exclusive prefix = output offset of each level
    const int cl = lane < P.nlevels ? selcnt[lane] : 0;
    int incl = max(cl, 0);
#pragma unroll
    for (int d = 1; d < ORB_MAX_LEVELS; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    if (blockIdx.x == 0 && warp == 0) {
        const bool bad = __any_sync(0xffffffffu, cl < 0);
        const int tot = __shfl_sync(0xffffffffu, incl, ORB_MAX_LEVELS - 1);
        if (lane == 0) P.nout[frame] = (bad || tot > P.kp_capacity) ? -1 : tot;
    }
    // DESC_PER_WARP slots per warp (stride = the grid): fewer, longer-lived CTAs, the prologue above paid once per warp
    for (int g = blockIdx.x * 8 + warp; g < P.sel_total; g += gridDim.x * 8) {
    // level of slot g: the last level whose sel_off <= g
    const unsigned lv_mask = __ballot_sync(0xffffffffu, lane < P.nlevels && P.lv[min(lane, ORB_MAX_LEVELS - 1)].sel_off <= g &&
                                                        P.lv[min(lane, ORB_MAX_LEVELS - 1)].sel_cap > 0);
    const int l = lv_mask ? 31 - __clz(lv_mask) : 0;
    const OrbLevel& L = P.lv[l];
    const int idx = g - L.sel_off;
    const int cnt_l = __shfl_sync(0xffffffffu, cl, l);
    const int before = __shfl_sync(0xffffffffu, incl - max(cl, 0), l);
    if (idx >= cnt_l) continue;
    const int o = before + idx;
    if (o >= P.kp_capacity) continue;
    ORB_CHECK(o >= 0 && l < P.nlevels && idx >= 0);
    const uint32_t key = P.sel[(long long)frame * P.sel_total + g];
    const int cx = ORB_PX(key) + ORB_BORDER, cy = ORB_PY(key) + ORB_BORDER;

    // this lane's 16 pattern points (issued early; used after the angle is known)
    const uint4 pat0 = __ldg(reinterpret_cast<const uint4*>(g_pattern) + 2 * lane);
    const uint4 pat1 = __ldg(reinterpret_cast<const uint4*>(g_pattern) + 2 * lane + 1);

    // ---- IC_Angle on the unblurred level: lane = column u of the disc, rows v = -vm .. vm
    const uint8_t* __restrict__ img = orb_level_ptr(P, l, frame);
    const int pitch = orb_level_pitch(P, l);
    int m10 = 0, m01 = 0;
#if ORB_IC_WORDS_PATH
    {
        // lane -> items lane, lane + 32, ...: item = (row, word) of the 31 x 9-word window; consecutive lanes read consecutive
        // words of a row (one instruction spans about 3.5 rows).  Nine independent word loads + nine weight loads per lane,
        // then two IDP.4A each: sum of u * I (u8 x s8) and the row's plain sum (times v) -- the integer moments of :86-103.
        const int sft = (cx - ORB_HALF_PATCH) & 3;
        const uint8_t* w0 = img + (size_t)(cy - ORB_HALF_PATCH) * pitch + (cx - ORB_HALF_PATCH - sft);
        const uint2* wt = g_ic_weights + sft * IC_ITEMS + lane;
        unsigned pix[IC_WORDS];
        uint2 wgt[IC_WORDS];
        int rows[IC_WORDS];
#pragma unroll
        for (int k = 0; k < IC_WORDS; ++k) {
            const int i = lane + 32 * k;
            const int row = min((i * 57) >> 9, ORB_PATCH - 1);          // i / 9 for i < 288; the padding items reread row 30 with weight 0
            const int word = i - IC_WORDS * ((i * 57) >> 9);
            rows[k] = row;
            wgt[k] = __ldg(wt + 32 * k);
            pix[k] = __ldg(reinterpret_cast<const unsigned*>(w0 + (size_t)row * pitch) + (i < ORB_PATCH * IC_WORDS ? word : 0));
        }
#pragma unroll
        for (int k = 0; k < IC_WORDS; ++k) {
            int su;
            asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(su) : "r"(pix[k]), "r"(wgt[k].x), "r"(m10));
            m10 = su;
            m01 += (rows[k] - ORB_HALF_PATCH) * (int)__dp4a(pix[k], wgt[k].y, 0u);
        }
    }
#else
    const int u = lane - ORB_HALF_PATCH;
    if (lane < ORB_PATCH) {
        const int vm = (int)((ORB_UMAX_NIBBLES >> (4 * abs(u))) & 15ull);
        ORB_CHECK(cx >= ORB_HALF_PATCH && cy >= ORB_HALF_PATCH && cx + ORB_HALF_PATCH < L.w && cy + ORB_HALF_PATCH < L.h);
        const uint8_t* c = img + (size_t)cy * pitch + cx + u;
        int sum = __ldg(c);
#pragma unroll
        for (int v = 1; v <= ORB_HALF_PATCH; ++v) {
            if (v <= vm) {
                const int lo = __ldg(c + v * pitch), up = __ldg(c - v * pitch);
                sum += lo + up;
                m01 += v * (lo - up);
            }
        }
        m10 = u * sum;
    }
#endif
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, d);
        m01 += __shfl_xor_sync(0xffffffffu, m01, d);
    }
    const float angle = orb_fast_atan2((float)m01, (float)m10);

    // ---- steered BRIEF on the blurred level
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.0);
    float a, b;
    orb_sincosf(__fmul_rn(angle, factorPI), &b, &a);
    const uint8_t* __restrict__ bc = P.blur + (long long)frame * P.blur_frame_bytes + L.blur_off + (long long)cy * L.pitch + cx;
    unsigned byte = 0;
    const uint32_t pw[8] = {pat0.x, pat0.y, pat0.z, pat0.w, pat1.x, pat1.y, pat1.z, pat1.w};
    int t0[8], t1[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t w = pw[k];
        const float x0 = (float)(int8_t)(w & 0xFF), y0 = (float)(int8_t)((w >> 8) & 0xFF);
        const float x1 = (float)(int8_t)((w >> 16) & 0xFF), y1 = (float)(int8_t)(w >> 24);
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
        const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
        const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
        ORB_CHECK(cy + r0 >= 0 && cy + r0 < L.h && cx + c0 >= 0 && cx + c0 < L.w && cy + r1 >= 0 && cy + r1 < L.h && cx + c1 >= 0 && cx + c1 < L.w);
        t0[k] = __ldg(bc + r0 * L.pitch + c0); t1[k] = __ldg(bc + r1 * L.pitch + c1);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) byte |= (unsigned)(t0[k] < t1[k]) << k;
    P.desc[((long long)frame * P.kp_capacity + o) * 32 + lane] = (uint8_t)byte;

    if (lane == 0) {
        OrbKeypointRec kp;
        kp.x = l ? __fmul_rn((float)cx, L.scale) : (float)cx;
        kp.y = l ? __fmul_rn((float)cy, L.scale) : (float)cy;
        kp.size = L.kp_size; kp.angle = angle; kp.response = (float)ORB_PS(key); kp.octave = l; kp.class_id = -1;
        reinterpret_cast<OrbKeypointRec*>(P.kps)[(long long)frame * P.kp_capacity + o] = kp;
    }
    }
}

int orb_launch_describe(const OrbPlanDev& P, int nframes, cudaStream_t st)
{
    const int ctas = P.sel_total > 0 ? (P.sel_total + 8 * DESC_PER_WARP - 1) / (8 * DESC_PER_WARP) : 1;   // >= 1 block: block 0 writes nout
    k_describe<<<dim3(ctas, nframes), 256, 0, st>>>(P);
    return 1;
}

// =================================================================================================
// Test hooks for the float steps that decide descriptor bits (orbgpu_debug_* in include/orbgpu.h).
// =================================================================================================
__global__ void k_debug_atan2(const float* y, const float* x, int n, float* out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = orb_fast_atan2(y[i], x[i]);
}

__global__ void k_debug_sincosf(const float* rad, int n, float* s, float* c)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) orb_sincosf(rad[i], &s[i], &c[i]);
}

__device__ __host__ inline unsigned long long orb_mix64(unsigned bits, unsigned sbits, unsigned cbits)
{
    unsigned long long h = (unsigned long long)bits * 0x9E3779B97F4A7C15ull ^ (((unsigned long long)sbits << 32) | cbits);
    h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
    return h;
}

__global__ void k_debug_sincos_checksum(unsigned first_bits, unsigned count, unsigned step, unsigned long long* sum)
{
    unsigned long long acc = 0;
    for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; i < count;
         i += (unsigned long long)gridDim.x * blockDim.x) {
        const unsigned bits = first_bits + (unsigned)i * step;
        float s, c;
        orb_sincosf(__uint_as_float(bits), &s, &c);
        acc += orb_mix64(bits, __float_as_uint(s), __float_as_uint(c));
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
    if ((threadIdx.x & 31) == 0) atomicAdd(sum, acc);
}

int orb_debug_atan2(const float* d_y, const float* d_x, int n, float* d_out, cudaStream_t st)
{
    k_debug_atan2<<<(n + 255) / 256, 256, 0, st>>>(d_y, d_x, n, d_out);
    return 1;
}
int orb_debug_sincosf(const float* d_rad, int n, float* d_s, float* d_c, cudaStream_t st)
{
    k_debug_sincosf<<<(n + 255) / 256, 256, 0, st>>>(d_rad, n, d_s, d_c);
    return 1;
}
int orb_debug_sincos_checksum(unsigned first_bits, unsigned count, unsigned step, unsigned long long* d_sum, cudaStream_t st)
{
    k_debug_sincos_checksum<<<148 * 8, 256, 0, st>>>(first_bits, count, step, d_sum);
    return 1;
}
