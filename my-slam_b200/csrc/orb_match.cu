// orb_match.cu -- Hamming matching (ORBmatcher::DescriptorDistance + the best / second-best scan).
//
// Reference: src/ORBmatcher.cc:1647-1663 (256-bit XOR + popcount) and the scan body shared by all
// Search* routines (canonical form :201-226): candidates are visited in index order with strict
// '<' updates, so ties keep the lowest index.  The kernels are integer-ALU (popc) bound, not a
// float contraction: one thread owns one query descriptor in 8 registers, candidate descriptors are
// staged in shared memory as uint4 pairs and broadcast to the warp.  No tensor cores.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <algorithm>
#include <new>
#include <vector>

#include "../../include/orbgpu.h"
#include "orb_common.cuh"

#define MATCH_THREADS 128
#define MATCH_TILE 128        // candidates staged per shared-memory tile (4 KB)

void orbgpu_internal_set_error(const char* s);   // orbgpu_extractor.cu (thread-local text behind orbgpu_last_cuda_error)
#define CUM(call)                                                                                     \
    do {                                                                                              \
        cudaError_t _e = (call);                                                                      \
        if (_e != cudaSuccess) {                                                                      \
            char _b[256];                                                                             \
            snprintf(_b, sizeof(_b), "%s: %s", #call, cudaGetErrorString(_e));                        \
            orbgpu_internal_set_error(_b);                                                            \
            return ORBGPU_ERR_CUDA;                                                                   \
        }                                                                                             \
    } while (0)

// 256-bit Hamming distance.  POPC issues at a quarter of the ALU rate, so the 8 XOR words are first folded
// by three carry-save adders (sum = a^b^c, carry = maj(a,b,c), one LOP3 each): 8 words -> 2 sum words + 3 carry
// words, d = popc(s1) + popc(s2) + 2 (popc(c0) + popc(c1) + popc(c2)): 5 POPC instead of 8.
__device__ __forceinline__ unsigned csa_sum(unsigned a, unsigned b, unsigned c) { return a ^ b ^ c; }
__device__ __forceinline__ unsigned csa_carry(unsigned a, unsigned b, unsigned c) { return (a & b) | (c & (a | b)); }

__device__ __forceinline__ int hamming256(const uint4& q0, const uint4& q1, const uint4& c0, const uint4& c1)
{
    const unsigned x0 = q0.x ^ c0.x, x1 = q0.y ^ c0.y, x2 = q0.z ^ c0.z, x3 = q0.w ^ c0.w;
    const unsigned x4 = q1.x ^ c1.x, x5 = q1.y ^ c1.y, x6 = q1.z ^ c1.z, x7 = q1.w ^ c1.w;
    const unsigned s0 = csa_sum(x0, x1, x2), k0 = csa_carry(x0, x1, x2);
    const unsigned s1 = csa_sum(x3, x4, x5), k1 = csa_carry(x3, x4, x5);
    const unsigned s2 = csa_sum(x6, x7, s0), k2 = csa_carry(x6, x7, s0);
    return __popc(s1) + __popc(s2) + 2 * (__popc(k0) + __popc(k1) + __popc(k2));
}

// The scan kernels' distance: 4 POPC and 14 LOP3 per 256 bits.  ncu shows the scan bound by the ALU pipe (LOP3, 64
// lanes / clk / SM) with POPC (XU pipe, 16 lanes / clk / SM) close behind, so both counts matter:
//  * the three weight-2 carry words of hamming256 go through one more carry-save adder (2 LOP3) and come out as one
//    weight-2 and one weight-4 word: 4 POPC instead of 5;
//  * descriptors are held TRANSFORMED, t(c) = (c.x, c.x ^ c.y, c.x ^ c.y ^ c.z, c.w) per 128-bit half (the query in
//    registers, the candidates in the shared-memory tile, transformed once when the tile is staged).  XOR commutes with
//    the transform, so one LOP3 each gives x0 = q0 ^ c0, y = x0 ^ x1 and s = x0 ^ x1 ^ x2, and the carry of the group is
//    the 3-input function  maj(x0, x1, x2) = y ? ~s : x0  (x0 != x1 makes the third input x2 = s ^ 1 decide, else x0):
//    4 LOP3 per three-word group instead of 3 XOR + 2.
__device__ __forceinline__ uint4 hamming_transform(const uint4& c) { return make_uint4(c.x, c.x ^ c.y, c.x ^ c.y ^ c.z, c.w); }

__device__ __forceinline__ int hamming256_t(const uint4& q0, const uint4& q1, const uint4& c0, const uint4& c1)
{
    const unsigned xa = q0.x ^ c0.x, ya = q0.y ^ c0.y, sa = q0.z ^ c0.z, ka = (ya & ~sa) | (~ya & xa);
    const unsigned xb = q1.x ^ c1.x, yb = q1.y ^ c1.y, sb = q1.z ^ c1.z, kb = (yb & ~sb) | (~yb & xb);
    const unsigned x3 = q0.w ^ c0.w, x7 = q1.w ^ c1.w;
    const unsigned s2 = csa_sum(x3, x7, sa), kc = csa_carry(x3, x7, sa);
    const unsigned t0 = csa_sum(ka, kb, kc), f0 = csa_carry(ka, kb, kc);
    // the weighted sum as three IMADs (FMA pipe, nearly idle here) instead of an IADD3 on the ALU pipe that LOP3 / VIMNMX fill
    int d = __popc(sb);
    asm("mad.lo.s32 %0, %1, 2, %0;" : "+r"(d) : "r"(__popc(t0)));
    asm("mad.lo.s32 %0, %1, 4, %0;" : "+r"(d) : "r"(__popc(f0)));
    asm("mad.lo.s32 %0, %1, 1, %0;" : "+r"(d) : "r"(__popc(s2)));
    return d;
}

// (distance, index) packed as d << 22 | j: min over the packed value is the lowest distance and, among equal
// distances, the lowest index -- the strict '<' update of src/ORBmatcher.cc:201-226.  Needs j < 2^22.
#define MATCH_IDX_BITS 22
#define MATCH_INIT ((256u << MATCH_IDX_BITS) | ((1u << MATCH_IDX_BITS) - 1u))

// Stage one tile of candidates in shared memory, transformed (hamming_transform).  The CTA cooperates.
__device__ __forceinline__ void stage_tile(const uint4* __restrict__ B, int t0, int tn, uint4* s_tile)
{
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * tn; i += MATCH_THREADS) s_tile[i] = hamming_transform(__ldg(B + 2 * (size_t)t0 + i));
    __syncthreads();
}

// Best / second-best scan of all nB candidates (index order) for this thread's (transformed) query, two candidates per
// step: inside a tile the key is 16 bits (distance << 6 | pair number, at most 64 pairs), the even candidate in the low
// half and the odd one in the high half of one register, so one packed 16x2 min / max serves two candidates
// (3 VIMNMX per two distances instead of six).  After each tile the four 16-bit winners are widened to the 32-bit
// (distance, global index) key and merged: all keys are distinct, so the two smallest keys overall are among the two
// smallest of each half.
__device__ __forceinline__ void scan_candidates(const uint4* __restrict__ B, int nB, bool active, const uint4& q0,
                                                const uint4& q1, uint4* s_tile, int& best, int& second, int& best_idx)
{
    unsigned pb = MATCH_INIT, ps = MATCH_INIT;                 // best and second-best packed values
    for (int t0 = 0; t0 < nB; t0 += MATCH_TILE) {
        const int tn = min(MATCH_TILE, nB - t0);
        stage_tile(B, t0, tn, s_tile);
        if (active) {
            unsigned b2 = 0xFFFFFFFFu, s2 = 0xFFFFFFFFu;
            const int np = tn >> 1;
#pragma unroll 4
            for (int jj = 0; jj < np; ++jj) {
                const int da = hamming256_t(q0, q1, s_tile[4 * jj], s_tile[4 * jj + 1]);
                const int db = hamming256_t(q0, q1, s_tile[4 * jj + 2], s_tile[4 * jj + 3]);
                const unsigned cur = (unsigned)(db * 64 + jj) * 65536u + (unsigned)(da * 64 + jj);
                const unsigned hi = __vmaxu2(b2, cur);
                b2 = __vminu2(b2, cur);
                s2 = __vminu2(s2, hi);
            }
#pragma unroll
            for (int h = 0; h < 4; ++h) {
                const unsigned v = ((h & 2) ? s2 : b2) >> ((h & 1) * 16) & 0xFFFFu;           // 0xFFFF = empty: widens to a key above MATCH_INIT
                const unsigned cur = ((v >> 6) << MATCH_IDX_BITS) | (unsigned)(t0 + 2 * (int)(v & 63u) + (h & 1));
                const unsigned hi = max(pb, cur);
                pb = min(pb, cur);
                ps = min(ps, hi);
            }
            if (tn & 1) {
                const int j = tn - 1;
                const unsigned cur = ((unsigned)hamming256_t(q0, q1, s_tile[2 * j], s_tile[2 * j + 1]) << MATCH_IDX_BITS) | (unsigned)(t0 + j);
                const unsigned hi = max(pb, cur);
                pb = min(pb, cur);
                ps = min(ps, hi);
            }
        }
    }
    best = (int)(pb >> MATCH_IDX_BITS);
    second = (int)(ps >> MATCH_IDX_BITS);
    best_idx = pb == MATCH_INIT ? -1 : (int)(pb & ((1u << MATCH_IDX_BITS) - 1u));
    if (nB == 0) best_idx = -1;
}

// The all-pairs count needs the best and the second-best DISTANCE only (accept rule of src/ORBmatcher.cc:228-232), not the
// index: the key is the bare distance, two candidates per register, and the running best / second of the even and the
// odd candidates are merged once at the end (multiset semantics: a distance that occurs twice is its own second-best,
// exactly as the strict '<' scan leaves it).
__device__ __forceinline__ void scan_distances(const uint4* __restrict__ B, int nB, bool active, const uint4& q0,
                                               const uint4& q1, uint4* s_tile, int& best, int& second)
{
    unsigned b2 = 0x01000100u, s2 = 0x01000100u;               // 256 | 256 << 16
    for (int t0 = 0; t0 < nB; t0 += MATCH_TILE) {
        const int tn = min(MATCH_TILE, nB - t0);
        stage_tile(B, t0, tn, s_tile);
        if (active) {
            const int np = tn >> 1;
#pragma unroll 4
            for (int jj = 0; jj < np; ++jj) {
                const int da = hamming256_t(q0, q1, s_tile[4 * jj], s_tile[4 * jj + 1]);
                const int db = hamming256_t(q0, q1, s_tile[4 * jj + 2], s_tile[4 * jj + 3]);
                const unsigned cur = (unsigned)db * 65536u + (unsigned)da;
                const unsigned hi = __vmaxu2(b2, cur);
                b2 = __vminu2(b2, cur);
                s2 = __vminu2(s2, hi);
            }
            if (tn & 1) {                                       // odd tail: the last candidate alone, its partner is "no candidate" (256)
                const int da = hamming256_t(q0, q1, s_tile[2 * (tn - 1)], s_tile[2 * (tn - 1) + 1]);
                const unsigned cur = 256u * 65536u + (unsigned)da;
                const unsigned hi = __vmaxu2(b2, cur);
                b2 = __vminu2(b2, cur);
                s2 = __vminu2(s2, hi);
            }
        }
    }
    const int be = (int)(b2 & 0xFFFFu), bo = (int)(b2 >> 16), se = (int)(s2 & 0xFFFFu), so = (int)(s2 >> 16);
    best = min(be, bo);
    second = min(max(be, bo), min(se, so));
}

__global__ void __launch_bounds__(MATCH_THREADS) k_match_bruteforce(const uint4* __restrict__ A, int nA,
                                                                    const uint4* __restrict__ B, int nB,
                                                                    int32_t* best_idx, uint16_t* best, uint16_t* second)
{
    __shared__ uint4 s_tile[2 * MATCH_TILE];
    const int q = blockIdx.x * MATCH_THREADS + threadIdx.x;
    const bool active = q < nA;
    uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
    if (active) { q0 = hamming_transform(__ldg(A + 2 * (size_t)q)); q1 = hamming_transform(__ldg(A + 2 * (size_t)q + 1)); }
    int b1 = 256, b2 = 256, bi = -1;
    scan_candidates(B, nB, active, q0, q1, s_tile, b1, b2, bi);
    if (active) { best_idx[q] = bi; best[q] = (uint16_t)b1; second[q] = (uint16_t)b2; }
}

// one CTA = MATCH_THREADS queries of keyframe a against all descriptors of keyframe b
__global__ void __launch_bounds__(MATCH_THREADS) k_match_count_pairs(const uint4* __restrict__ desc, int n_per_kf,
                                                                     const int2* __restrict__ pairs, int th_low, float ratio,
                                                                     int32_t* counts)
{
    __shared__ uint4 s_tile[2 * MATCH_TILE];
    const int2 pr = pairs[blockIdx.y];
    const uint4* A = desc + 2 * (size_t)pr.x * n_per_kf;
    const uint4* B = desc + 2 * (size_t)pr.y * n_per_kf;
    const int q = blockIdx.x * MATCH_THREADS + threadIdx.x;
    const bool active = q < n_per_kf;
    uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
    if (active) { q0 = hamming_transform(__ldg(A + 2 * (size_t)q)); q1 = hamming_transform(__ldg(A + 2 * (size_t)q + 1)); }
    int b1 = 256, b2 = 256;
    scan_distances(B, n_per_kf, active, q0, q1, s_tile, b1, b2);
    const bool ok = active && b1 <= th_low && (float)b1 < __fmul_rn(ratio, (float)b2);
    const int c = __syncthreads_count(ok);
    if (threadIdx.x == 0 && c) atomicAdd(&counts[blockIdx.y], c);
}

// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-303), one warp per map point.  Lane r owns rows r, r + 32, ...:
// the row's N distances go into the lane's own 257-bin histogram (distances are 0..256; layout [bin][lane] so the lanes of
// a warp never collide), the median is the first bin whose running count exceeds k = int(0.5 (N - 1)).  The best row is
// the minimum of (median << 16 | row): strict '<' in row order = the first row among equal medians (:296-300).
#define DD_WARPS 4
#define DD_BINS 257

__global__ void __launch_bounds__(DD_WARPS * 32) k_distinctive(const uint4* __restrict__ desc, const int* __restrict__ group_start,
                                                                int n_groups, int* __restrict__ best_idx)
{
    extern __shared__ unsigned short dd_hist[];                // [DD_WARPS][DD_BINS][32]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = blockIdx.x * DD_WARPS + warp;
    if (g >= n_groups) return;
    const int s0 = group_start[g], N = group_start[g + 1] - s0;
    if (N <= 0) { if (lane == 0) best_idx[g] = -1; return; }
    unsigned short* h = dd_hist + (size_t)warp * DD_BINS * 32 + lane;
    const int k = (N - 1) >> 1;                                // vDists[0.5 * (N - 1)]
    unsigned best = 0xFFFFFFFFu;
    for (int r = lane; r < N; r += 32) {
        for (int b = 0; b < DD_BINS; ++b) h[b * 32] = 0;
        const uint4 q0 = __ldg(desc + 2 * (size_t)(s0 + r)), q1 = __ldg(desc + 2 * (size_t)(s0 + r) + 1);
        for (int j = 0; j < N; ++j) {
            const uint4 c0 = __ldg(desc + 2 * (size_t)(s0 + j)), c1 = __ldg(desc + 2 * (size_t)(s0 + j) + 1);
            h[hamming256(q0, q1, c0, c1) * 32]++;              // j == r gives Distances[i][i] = 0
        }
        int run = 0, med = 0;
        for (int b = 0; b < DD_BINS; ++b) { run += h[b * 32]; if (run > k) { med = b; break; } }
        best = min(best, ((unsigned)med << 16) | (unsigned)r);
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0) best_idx[g] = (int)(best & 0xFFFFu);
}

// Instruction-rate micro-benchmarks behind the roofline of the matching kernels ("a popc
// micro-benchmark peak").  Eight independent dependency chains per thread; which = 0: one POPC per chain step (plus the
// XOR and the add that keep the value alive, both on the faster ALU pipe), which = 1: LOP3 only.  out keeps the result
// live.  The kernel reports nothing itself: the host divides the op count by the CUDA-event time.
__global__ void __launch_bounds__(256) k_pipe_peak(unsigned* out, int iters, int which)
{
    unsigned a[8];
    const unsigned k = (blockIdx.x * 256u + threadIdx.x) * 0x9E3779B9u + 0x7F4A7C15u;
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = k ^ (0x01010101u * (unsigned)(j + 1));
    if (which == 0) {
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = (unsigned)__popc(a[j] ^ k) + a[j];
        }
    } else {
        unsigned m = k | 1u;
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int j = 0; j < 8; ++j) a[j] = (a[j] & m) ^ a[(j + 1) & 7];                // one LOP3 (a 3-input function)
            m += 2u;                                                                       // keeps the loop from collapsing
        }
    }
    unsigned r = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) r ^= a[j];
    if (r == 0x12345678u) out[0] = r;                                                      // practically never true
}

struct orbgpu_matcher {
    int device = 0;
    cudaStream_t st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    uint4* d_desc = nullptr; bool own_desc = false;
    int num_kf = 0, n_per_kf = 0;
    int2* d_pairs = nullptr; int32_t* d_counts = nullptr; int pair_cap = 0;
    double last_ms = 0;
    // all-pairs sharded over ranks (orbgpu_matcher_all_pairs_sharded): this rank's pair list, its counts padded to the
    // per-rank length, and the all_gather receive buffer; rebuilt when (keyframes, rank, world) change
    int sh_kf = 0, sh_rank = -1, sh_world = 0, sh_block = 0, sh_mine = 0, sh_per_rank = 0;
    int2* d_sh_pairs = nullptr; int32_t* d_sh_local = nullptr; int32_t* d_sh_all = nullptr;
    cudaEvent_t e2 = nullptr;
    double sh_kernel_ms = 0, sh_gather_ms = 0;
};

// NCCL is bound at run time (dlopen of the libnccl.so.2 the process already holds, or the system one): liborbgpu.so has
// no link-time dependency on it, and hosts that never shard never load it.
namespace {
struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
std::mutex g_nccl_mu;
NcclApi g_nccl;
bool g_nccl_tried = false;

const NcclApi* nccl_api()
{
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (!g_nccl_tried) {
        g_nccl_tried = true;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (h) {
            NcclApi a;
            a.lib = h;
            *(void**)&a.GetUniqueId = dlsym(h, "ncclGetUniqueId");
            *(void**)&a.CommInitRank = dlsym(h, "ncclCommInitRank");
            *(void**)&a.CommInitAll = dlsym(h, "ncclCommInitAll");
            *(void**)&a.CommDestroy = dlsym(h, "ncclCommDestroy");
            *(void**)&a.AllGather = dlsym(h, "ncclAllGather");
            *(void**)&a.GroupStart = dlsym(h, "ncclGroupStart");
            *(void**)&a.GroupEnd = dlsym(h, "ncclGroupEnd");
            *(void**)&a.GetErrorString = dlsym(h, "ncclGetErrorString");
            if (a.GetUniqueId && a.CommInitRank && a.CommInitAll && a.CommDestroy && a.AllGather && a.GroupStart && a.GroupEnd && a.GetErrorString)
                g_nccl = a;
        }
    }
    return g_nccl.lib ? &g_nccl : nullptr;
}
}  // namespace

#define NC(call)                                                                                      \
    do {                                                                                              \
        const ncclResult_t _r = (call);                                                               \
        if (_r != ncclSuccess) {                                                                      \
            char _b[256];                                                                             \
            snprintf(_b, sizeof(_b), "%s: %s", #call, nccl_api() ? nccl_api()->GetErrorString(_r) : "?"); \
            orbgpu_internal_set_error(_b);                                                            \
            return ORBGPU_ERR_CUDA;                                                                   \
        }                                                                                             \
    } while (0)

struct orbgpu_comm {
    int device = 0, rank = 0, world = 1;
    ncclComm_t comm = nullptr;
};

// all_gather output (rank-major, per_rank padded slots each) -> pair order.  Slot i of rank r holds pair
// ((i / block) * world + r) * block + i % block (blocks of `block` pairs dealt round-robin).
__global__ void __launch_bounds__(256) k_scatter_pair_counts(const int32_t* __restrict__ all, int per_rank, int world, int block,
                                                             int num_pairs, int32_t* __restrict__ out)
{
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= per_rank * world) return;
    const int r = t / per_rank, i = t - r * per_rank;
    const long long p = ((long long)(i / block) * world + r) * block + i % block;
    if (p < num_pairs) out[p] = all[t];
}

// MapPoint::UpdateNormalAndDepth (src/MapPoint.cc:330-371), one thread per map point: the mean viewing direction is a float
// accumulation in observation order, so a point's walk stays serial while the points run in parallel.  OpenCV's arithmetic:
// cv::norm in double, "normal + normali / norm" = cv::scaleAdd = fmaf(v, (float)(1.0 / norm), acc), "normal / n" = v * (float)(1.0 / n).
__global__ void __launch_bounds__(128) k_update_normals(const float* __restrict__ pos, const int* __restrict__ obs_start,
                                                        const float* __restrict__ centers, const float* __restrict__ ref_center,
                                                        const float* __restrict__ ref_scale, const float* __restrict__ last_scale,
                                                        int n_points, float* __restrict__ normal, float* __restrict__ max_dist,
                                                        float* __restrict__ min_dist)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_points) return;
    const int b = obs_start[g], e = obs_start[g + 1];
    if (e <= b) return;                                                           // observations.empty(): nothing changes (:345-346)
    const float px = pos[3 * g], py = pos[3 * g + 1], pz = pos[3 * g + 2];
    float ax = 0.f, ay = 0.f, az = 0.f;
    for (int j = b; j < e; ++j) {
        const float vx = __fsub_rn(px, __ldg(centers + 3 * (size_t)j)), vy = __fsub_rn(py, __ldg(centers + 3 * (size_t)j + 1)),
                    vz = __fsub_rn(pz, __ldg(centers + 3 * (size_t)j + 2));
        const double nrm = sqrt(__dadd_rn(__dadd_rn(__dmul_rn((double)vx, (double)vx), __dmul_rn((double)vy, (double)vy)),
                                          __dmul_rn((double)vz, (double)vz)));
        const float a = (float)(1.0 / nrm);
        ax = __fmaf_rn(vx, a, ax); ay = __fmaf_rn(vy, a, ay); az = __fmaf_rn(vz, a, az);
    }
    const float cx = __fsub_rn(px, ref_center[3 * g]), cy = __fsub_rn(py, ref_center[3 * g + 1]), cz = __fsub_rn(pz, ref_center[3 * g + 2]);
    const float dist = (float)sqrt(__dadd_rn(__dadd_rn(__dmul_rn((double)cx, (double)cx), __dmul_rn((double)cy, (double)cy)),
                                             __dmul_rn((double)cz, (double)cz)));
    const float mx = __fmul_rn(dist, ref_scale[g]);
    max_dist[g] = mx;
    min_dist[g] = __fdiv_rn(mx, last_scale[g]);
    const float inv_n = (float)(1.0 / (double)(e - b));
    normal[3 * g] = __fmul_rn(ax, inv_n); normal[3 * g + 1] = __fmul_rn(ay, inv_n); normal[3 * g + 2] = __fmul_rn(az, inv_n);
}

// Per-thread scratch of the stateless entry points (device block + pinned mirror + stream), grown on demand and kept: no
// cudaMalloc / cudaFree and no pageable copies on the call path.
namespace {
struct MatchScratch {
    int device = -1;
    size_t cap = 0;
    uint8_t* d = nullptr;
    uint8_t* h = nullptr;
    cudaStream_t st = nullptr;
    int ensure(int dev, size_t bytes)
    {
        if (dev != device) {
            if (device >= 0) { cudaSetDevice(device); cudaFree(d); cudaFreeHost(h); if (st) cudaStreamDestroy(st); }
            d = nullptr; h = nullptr; st = nullptr; cap = 0; device = -1;
            CUM(cudaSetDevice(dev));
            CUM(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            device = dev;
        }
        CUM(cudaSetDevice(dev));
        if (bytes <= cap) return ORBGPU_OK;
        cudaFree(d); cudaFreeHost(h); d = nullptr; h = nullptr; cap = 0;
        const size_t want = (bytes + bytes / 2 + 65535) & ~(size_t)65535;
        CUM(cudaMalloc(&d, want));
        CUM(cudaHostAlloc(&h, want, cudaHostAllocDefault));
        cap = want;
        return ORBGPU_OK;
    }
};
thread_local MatchScratch g_scratch;
inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }
}  // namespace

// Scratch of orbgpu_match_bruteforce_noalloc: a fixed device array that every GPU loading the library holds until
// exit (both descriptor sets and the three result arrays of the largest accepted call, each region 256-byte aligned).
// In host code `g_noalloc_scratch` is only a stand-in for the symbol: the device address comes from
// cudaGetSymbolAddress, once per device with that device current (noalloc_scratch_ptr).
#define NOALLOC_MAX ORBGPU_NOALLOC_MAX_DESCRIPTORS
#define NOALLOC_BYTES (2 * 32 * NOALLOC_MAX + 4 * NOALLOC_MAX + 2 * 2 * NOALLOC_MAX)
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

namespace {
#define NOALLOC_MAX_DEVICES 64
std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[NOALLOC_MAX_DEVICES];     // device address of g_noalloc_scratch, per device ordinal

// `device` must be current.  Returns ORBGPU_OK and the device's own copy of the array.
int noalloc_scratch_ptr(int device, unsigned char** out)
{
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* p = nullptr;
        CUM(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[device];
    return ORBGPU_OK;
}
}  // namespace

extern "C" {

int orbgpu_hamming256(const void* a, const void* b)
{
    uint64_t x[4], y[4];
    memcpy(x, a, 32); memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) +
           __builtin_popcountll(x[2] ^ y[2]) + __builtin_popcountll(x[3] ^ y[3]);
}

int orbgpu_match_bruteforce(int device, const uint8_t* A, int nA, const uint8_t* B, int nB,
                            int32_t* best_idx, uint16_t* best, uint16_t* second)
{
    if (!A || !B || !best_idx || !best || !second || nA < 1 || nB < 0) return ORBGPU_ERR_ARG;
    if (nB >= (1 << 22) - 1) return ORBGPU_ERR_UNSUPPORTED;   // candidate index is packed into 22 bits
    const size_t o_a = 0, o_b = up256(32 * (size_t)nA), o_i = up256(o_b + 32 * (size_t)nB), o_1 = up256(o_i + 4 * (size_t)nA),
                 o_2 = up256(o_1 + 2 * (size_t)nA), total = up256(o_2 + 2 * (size_t)nA);
    MatchScratch& S = g_scratch;
    const int rc = S.ensure(device, total);
    if (rc != ORBGPU_OK) return rc;
    memcpy(S.h + o_a, A, 32 * (size_t)nA);
    if (nB > 0) memcpy(S.h + o_b, B, 32 * (size_t)nB);
    CUM(cudaMemcpyAsync(S.d, S.h, o_b + 32 * (size_t)nB, cudaMemcpyHostToDevice, S.st));
    k_match_bruteforce<<<(nA + MATCH_THREADS - 1) / MATCH_THREADS, MATCH_THREADS, 0, S.st>>>(
        reinterpret_cast<const uint4*>(S.d + o_a), nA, reinterpret_cast<const uint4*>(S.d + o_b), nB,
        reinterpret_cast<int32_t*>(S.d + o_i), reinterpret_cast<uint16_t*>(S.d + o_1), reinterpret_cast<uint16_t*>(S.d + o_2));
    CUM(cudaGetLastError());
    CUM(cudaMemcpyAsync(S.h + o_i, S.d + o_i, total - o_i, cudaMemcpyDeviceToHost, S.st));
    CUM(cudaStreamSynchronize(S.st));
    memcpy(best_idx, S.h + o_i, 4 * (size_t)nA);
    memcpy(best, S.h + o_1, 2 * (size_t)nA);
    memcpy(second, S.h + o_2, 2 * (size_t)nA);
    return ORBGPU_OK;
}

// The allocation-free variant (include/orbgpu.h): same kernel, same results, the staging area is g_noalloc_scratch
// instead of the grown-on-demand MatchScratch; plain cudaMemcpy on the legacy default stream, so no stream, event,
// pinned or device allocation is made here.  Nothing in the array has to start at zero: the two descriptor regions
// are overwritten by this call's uploads before the kernel reads them, and the kernel writes all nA entries of the
// three result regions before they are copied back; bytes beyond nA / nB keep the previous call's contents and are
// never read.
int orbgpu_match_bruteforce_noalloc(int device, const uint8_t* A, int nA, const uint8_t* B, int nB,
                                    int32_t* best_idx, uint16_t* best, uint16_t* second)
{
    if (!A || !B || !best_idx || !best || !second || nA < 1 || nB < 0) return ORBGPU_ERR_ARG;
    if (nA > NOALLOC_MAX || nB > NOALLOC_MAX) {
        char b[160];
        snprintf(b, sizeof(b), "orbgpu_match_bruteforce_noalloc: nA = %d / nB = %d exceed the fixed scratch (at most %d descriptors per set)",
                 nA, nB, NOALLOC_MAX);
        orbgpu_internal_set_error(b);
        return ORBGPU_ERR_CAPACITY;
    }
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return ORBGPU_ERR_ARG;
    CUM(cudaSetDevice(device));
    unsigned char* d = nullptr;
    const int rc = noalloc_scratch_ptr(device, &d);
    if (rc != ORBGPU_OK) return rc;
    unsigned char* d_a = d;
    unsigned char* d_b = d_a + 32 * (size_t)NOALLOC_MAX;
    unsigned char* d_i = d_b + 32 * (size_t)NOALLOC_MAX;
    unsigned char* d_1 = d_i + 4 * (size_t)NOALLOC_MAX;
    unsigned char* d_2 = d_1 + 2 * (size_t)NOALLOC_MAX;
    CUM(cudaMemcpy(d_a, A, 32 * (size_t)nA, cudaMemcpyHostToDevice));
    if (nB > 0) CUM(cudaMemcpy(d_b, B, 32 * (size_t)nB, cudaMemcpyHostToDevice));
    k_match_bruteforce<<<(nA + MATCH_THREADS - 1) / MATCH_THREADS, MATCH_THREADS>>>(
        reinterpret_cast<const uint4*>(d_a), nA, reinterpret_cast<const uint4*>(d_b), nB,
        reinterpret_cast<int32_t*>(d_i), reinterpret_cast<uint16_t*>(d_1), reinterpret_cast<uint16_t*>(d_2));
    CUM(cudaGetLastError());
    CUM(cudaMemcpy(best_idx, d_i, 4 * (size_t)nA, cudaMemcpyDeviceToHost));
    CUM(cudaMemcpy(best, d_1, 2 * (size_t)nA, cudaMemcpyDeviceToHost));
    CUM(cudaMemcpy(second, d_2, 2 * (size_t)nA, cudaMemcpyDeviceToHost));
    return ORBGPU_OK;
}

int orbgpu_distinctive_descriptors(int device, const uint8_t* desc, const int32_t* group_start, int n_groups, int32_t* best_idx)
{
    if (!group_start || !best_idx || n_groups < 0) return ORBGPU_ERR_ARG;
    if (n_groups == 0) return ORBGPU_OK;
    const int total = group_start[n_groups];
    if (group_start[0] != 0 || total < 0 || (total > 0 && !desc)) return ORBGPU_ERR_ARG;
    for (int g = 0; g < n_groups; ++g) {
        if (group_start[g + 1] < group_start[g]) return ORBGPU_ERR_ARG;
        if (group_start[g + 1] - group_start[g] > 0xFFFF) return ORBGPU_ERR_UNSUPPORTED;     // row index rides in 16 bits
    }
    CUM(cudaSetDevice(device));
    const size_t smem = (size_t)DD_WARPS * DD_BINS * 32 * sizeof(unsigned short);
    static OrbSmemOptIn optin;
    CUM(optin.ensure(k_distinctive, (int)smem));
    const size_t o_gs = up256((size_t)32 * (size_t)total), o_out = up256(o_gs + (size_t)4 * (size_t)(n_groups + 1)),
                 bytes = up256(o_out + (size_t)4 * (size_t)n_groups);
    MatchScratch& S = g_scratch;
    const int rc = S.ensure(device, bytes);
    if (rc != ORBGPU_OK) return rc;
    if (total > 0) memcpy(S.h, desc, (size_t)32 * total);
    memcpy(S.h + o_gs, group_start, (size_t)4 * (n_groups + 1));
    CUM(cudaMemcpyAsync(S.d, S.h, o_gs + (size_t)4 * (n_groups + 1), cudaMemcpyHostToDevice, S.st));
    k_distinctive<<<(n_groups + DD_WARPS - 1) / DD_WARPS, DD_WARPS * 32, smem, S.st>>>(reinterpret_cast<const uint4*>(S.d),
                                                                                       reinterpret_cast<const int*>(S.d + o_gs), n_groups,
                                                                                       reinterpret_cast<int*>(S.d + o_out));
    CUM(cudaGetLastError());
    CUM(cudaMemcpyAsync(S.h + o_out, S.d + o_out, (size_t)4 * n_groups, cudaMemcpyDeviceToHost, S.st));
    CUM(cudaStreamSynchronize(S.st));
    memcpy(best_idx, S.h + o_out, (size_t)4 * n_groups);
    return ORBGPU_OK;
}

int orbgpu_update_normals_and_depths(int device, const float* pos, const int32_t* obs_start, const float* centers,
                                     const float* ref_center, const float* ref_scale, const float* last_scale, int n_points,
                                     float* normal, float* max_dist, float* min_dist)
{
    if (n_points < 0) return ORBGPU_ERR_ARG;
    if (n_points == 0) return ORBGPU_OK;
    if (!pos || !obs_start || !ref_center || !ref_scale || !last_scale || !normal || !max_dist || !min_dist) return ORBGPU_ERR_ARG;
    const int total = obs_start[n_points];
    if (obs_start[0] != 0 || total < 0 || (total > 0 && !centers)) return ORBGPU_ERR_ARG;
    for (int g = 0; g < n_points; ++g)
        if (obs_start[g + 1] < obs_start[g]) return ORBGPU_ERR_ARG;
    const size_t np = (size_t)n_points;
    // inputs: pos | ref_center | ref_scale | last_scale | obs_start | centers; in/out: normal | max | min
    const size_t o_pos = 0, o_rc = up256(12 * np), o_rs = up256(o_rc + 12 * np), o_ls = up256(o_rs + 4 * np), o_os = up256(o_ls + 4 * np),
                 o_c = up256(o_os + 4 * (np + 1)), o_n = up256(o_c + 12 * (size_t)total), o_mx = up256(o_n + 12 * np),
                 o_mn = up256(o_mx + 4 * np), bytes = up256(o_mn + 4 * np);
    MatchScratch& S = g_scratch;
    const int rc = S.ensure(device, bytes);
    if (rc != ORBGPU_OK) return rc;
    memcpy(S.h + o_pos, pos, 12 * np); memcpy(S.h + o_rc, ref_center, 12 * np);
    memcpy(S.h + o_rs, ref_scale, 4 * np); memcpy(S.h + o_ls, last_scale, 4 * np);
    memcpy(S.h + o_os, obs_start, 4 * (np + 1));
    if (total > 0) memcpy(S.h + o_c, centers, 12 * (size_t)total);
    memcpy(S.h + o_n, normal, 12 * np); memcpy(S.h + o_mx, max_dist, 4 * np); memcpy(S.h + o_mn, min_dist, 4 * np);
    CUM(cudaMemcpyAsync(S.d, S.h, bytes, cudaMemcpyHostToDevice, S.st));
    k_update_normals<<<(n_points + 127) / 128, 128, 0, S.st>>>(
        reinterpret_cast<const float*>(S.d + o_pos), reinterpret_cast<const int*>(S.d + o_os), reinterpret_cast<const float*>(S.d + o_c),
        reinterpret_cast<const float*>(S.d + o_rc), reinterpret_cast<const float*>(S.d + o_rs), reinterpret_cast<const float*>(S.d + o_ls),
        n_points, reinterpret_cast<float*>(S.d + o_n), reinterpret_cast<float*>(S.d + o_mx), reinterpret_cast<float*>(S.d + o_mn));
    CUM(cudaGetLastError());
    CUM(cudaMemcpyAsync(S.h + o_n, S.d + o_n, bytes - o_n, cudaMemcpyDeviceToHost, S.st));
    CUM(cudaStreamSynchronize(S.st));
    memcpy(normal, S.h + o_n, 12 * np); memcpy(max_dist, S.h + o_mx, 4 * np); memcpy(min_dist, S.h + o_mn, 4 * np);
    return ORBGPU_OK;
}

int orbgpu_debug_pipe_peak(int device, int which, double* ops_per_s)
{
    if (!ops_per_s || which < 0 || which > 1) return ORBGPU_ERR_ARG;
    MatchScratch& S = g_scratch;
    const int rc = S.ensure(device, 256);
    if (rc != ORBGPU_OK) return rc;
    int sms = 0;
    CUM(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    cudaEvent_t e0, e1;
    CUM(cudaEventCreate(&e0)); CUM(cudaEventCreate(&e1));
    const int iters = 8192, ctas = sms * 8;
    double best = 0.0;
    for (int rep = 0; rep < 4; ++rep) {                                                    // first pass warms the clocks
        cudaEventRecord(e0, S.st);
        k_pipe_peak<<<ctas, 256, 0, S.st>>>(reinterpret_cast<unsigned*>(S.d), iters, which);
        cudaEventRecord(e1, S.st);
        cudaError_t ce = cudaStreamSynchronize(S.st);
        float ms = 0.f;
        if (ce == cudaSuccess) ce = cudaEventElapsedTime(&ms, e0, e1);
        if (ce != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); CUM(ce); }
        const double rate = (double)ctas * 256.0 * iters * 8.0 / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *ops_per_s = best;
    return ORBGPU_OK;
}

int orbgpu_matcher_create(orbgpu_matcher** out, int device)
{
    if (!out) return ORBGPU_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    CUM(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return ORBGPU_ERR_ARG;
    CUM(cudaSetDevice(device));
    orbgpu_matcher* m = new (std::nothrow) orbgpu_matcher();
    if (!m) return ORBGPU_ERR_NOMEM;
    m->device = device;
    CUM(cudaStreamCreateWithFlags(&m->st, cudaStreamNonBlocking));
    CUM(cudaEventCreate(&m->e0)); CUM(cudaEventCreate(&m->e1)); CUM(cudaEventCreate(&m->e2));
    *out = m;
    return ORBGPU_OK;
}

void orbgpu_matcher_destroy(orbgpu_matcher* m)
{
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->st) cudaStreamSynchronize(m->st);
    if (m->own_desc) cudaFree(m->d_desc);
    cudaFree(m->d_pairs); cudaFree(m->d_counts);
    cudaFree(m->d_sh_pairs); cudaFree(m->d_sh_local); cudaFree(m->d_sh_all);
    if (m->e2) cudaEventDestroy(m->e2);
    if (m->e0) cudaEventDestroy(m->e0);
    if (m->e1) cudaEventDestroy(m->e1);
    if (m->st) cudaStreamDestroy(m->st);
    delete m;
}

int orbgpu_matcher_set_keyframes(orbgpu_matcher* m, const uint8_t* desc, int num_keyframes, int n_per_keyframe)
{
    if (!m || !desc || num_keyframes < 1 || n_per_keyframe < 1) return ORBGPU_ERR_ARG;
    if (n_per_keyframe >= (1 << 22) - 1) return ORBGPU_ERR_UNSUPPORTED;
    CUM(cudaSetDevice(m->device));
    if (m->own_desc) cudaFree(m->d_desc);
    m->d_desc = nullptr; m->own_desc = false;
    const size_t bytes = 32 * (size_t)num_keyframes * n_per_keyframe;
    CUM(cudaMalloc(&m->d_desc, bytes));
    m->own_desc = true;
    CUM(cudaMemcpy(m->d_desc, desc, bytes, cudaMemcpyHostToDevice));
    m->num_kf = num_keyframes; m->n_per_kf = n_per_keyframe;
    return ORBGPU_OK;
}

int orbgpu_matcher_set_keyframes_device(orbgpu_matcher* m, const uint8_t* d_desc, int num_keyframes, int n_per_keyframe)
{
    if (!m || !d_desc || num_keyframes < 1 || n_per_keyframe < 1 || ((uintptr_t)d_desc & 15)) return ORBGPU_ERR_ARG;
    if (n_per_keyframe >= (1 << 22) - 1) return ORBGPU_ERR_UNSUPPORTED;
    CUM(cudaSetDevice(m->device));
    if (m->own_desc) cudaFree(m->d_desc);
    m->d_desc = reinterpret_cast<uint4*>(const_cast<uint8_t*>(d_desc));
    m->own_desc = false;
    m->num_kf = num_keyframes; m->n_per_kf = n_per_keyframe;
    return ORBGPU_OK;
}

int orbgpu_matcher_count_pairs_device(orbgpu_matcher* m, const int32_t* d_pairs, int num_pairs, int th_low, float ratio,
                                      int32_t* d_counts)
{
    if (!m || !m->d_desc || !d_pairs || !d_counts || num_pairs < 1) return ORBGPU_ERR_ARG;
    CUM(cudaSetDevice(m->device));
    CUM(cudaMemsetAsync(d_counts, 0, sizeof(int32_t) * (size_t)num_pairs, m->st));
    CUM(cudaEventRecord(m->e0, m->st));
    const int qblocks = (m->n_per_kf + MATCH_THREADS - 1) / MATCH_THREADS;
    for (int p0 = 0; p0 < num_pairs; p0 += 65535) {          // gridDim.y limit
        const int np = num_pairs - p0 < 65535 ? num_pairs - p0 : 65535;
        k_match_count_pairs<<<dim3(qblocks, np), MATCH_THREADS, 0, m->st>>>(m->d_desc, m->n_per_kf,
                                                                           reinterpret_cast<const int2*>(d_pairs) + p0,
                                                                           th_low, ratio, d_counts + p0);
    }
    CUM(cudaGetLastError());
    CUM(cudaEventRecord(m->e1, m->st));
    CUM(cudaStreamSynchronize(m->st));
    float ms = 0.f;
    CUM(cudaEventElapsedTime(&ms, m->e0, m->e1));
    m->last_ms = ms;
    return ORBGPU_OK;
}

int orbgpu_matcher_count_pairs(orbgpu_matcher* m, const int32_t* kf_a, const int32_t* kf_b, int num_pairs, int th_low,
                               float ratio, int32_t* counts)
{
    if (!m || !m->d_desc || !kf_a || !kf_b || !counts || num_pairs < 1) return ORBGPU_ERR_ARG;
    CUM(cudaSetDevice(m->device));
    for (int p = 0; p < num_pairs; ++p)
        if (kf_a[p] < 0 || kf_a[p] >= m->num_kf || kf_b[p] < 0 || kf_b[p] >= m->num_kf) return ORBGPU_ERR_ARG;
    if (num_pairs > m->pair_cap) {
        cudaFree(m->d_pairs); cudaFree(m->d_counts); m->d_pairs = nullptr; m->d_counts = nullptr; m->pair_cap = 0;
        CUM(cudaMalloc(&m->d_pairs, sizeof(int2) * (size_t)num_pairs));
        CUM(cudaMalloc(&m->d_counts, sizeof(int32_t) * (size_t)num_pairs));
        m->pair_cap = num_pairs;
    }
    int2* h = new (std::nothrow) int2[num_pairs];
    if (!h) return ORBGPU_ERR_NOMEM;
    for (int p = 0; p < num_pairs; ++p) h[p] = make_int2(kf_a[p], kf_b[p]);
    cudaError_t ce = cudaMemcpy(m->d_pairs, h, sizeof(int2) * (size_t)num_pairs, cudaMemcpyHostToDevice);
    delete[] h;
    CUM(ce);
    int rc = orbgpu_matcher_count_pairs_device(m, reinterpret_cast<const int32_t*>(m->d_pairs), num_pairs, th_low, ratio, m->d_counts);
    if (rc != ORBGPU_OK) return rc;
    CUM(cudaMemcpy(counts, m->d_counts, sizeof(int32_t) * (size_t)num_pairs, cudaMemcpyDeviceToHost));
    return ORBGPU_OK;
}

// ---- all-pairs matching sharded over ranks, result gather by ncclAllGather on the kernel's own stream -------------------

int orbgpu_comm_unique_id(void* id128)
{
    if (!id128) return ORBGPU_ERR_ARG;
    const NcclApi* N = nccl_api();
    if (!N) { orbgpu_internal_set_error("libnccl.so.2 not found"); return ORBGPU_ERR_UNSUPPORTED; }
    ncclUniqueId id;
    NC(N->GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return ORBGPU_OK;
}

int orbgpu_comm_init_rank(orbgpu_comm** out, int device, int rank, int world, const void* id128)
{
    if (!out || !id128 || world < 1 || rank < 0 || rank >= world) return ORBGPU_ERR_ARG;
    *out = nullptr;
    const NcclApi* N = nccl_api();
    if (!N) { orbgpu_internal_set_error("libnccl.so.2 not found"); return ORBGPU_ERR_UNSUPPORTED; }
    CUM(cudaSetDevice(device));
    orbgpu_comm* c = new (std::nothrow) orbgpu_comm();
    if (!c) return ORBGPU_ERR_NOMEM;
    c->device = device; c->rank = rank; c->world = world;
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    const ncclResult_t r = N->CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) { delete c; NC(r); }
    *out = c;
    return ORBGPU_OK;
}

int orbgpu_comm_init_all(orbgpu_comm** out, const int* devices, int n)
{
    if (!out || !devices || n < 1 || n > 64) return ORBGPU_ERR_ARG;
    for (int i = 0; i < n; ++i) out[i] = nullptr;
    const NcclApi* N = nccl_api();
    if (!N) { orbgpu_internal_set_error("libnccl.so.2 not found"); return ORBGPU_ERR_UNSUPPORTED; }
    ncclComm_t comms[64];
    NC(N->CommInitAll(comms, n, devices));
    for (int i = 0; i < n; ++i) {
        orbgpu_comm* c = new (std::nothrow) orbgpu_comm();
        if (!c) return ORBGPU_ERR_NOMEM;
        c->device = devices[i]; c->rank = i; c->world = n; c->comm = comms[i];
        out[i] = c;
    }
    return ORBGPU_OK;
}

void orbgpu_comm_destroy(orbgpu_comm* c)
{
    if (!c) return;
    const NcclApi* N = nccl_api();
    if (N && c->comm) { cudaSetDevice(c->device); N->CommDestroy(c->comm); }
    delete c;
}

static int adaptive_pair_block(long long num_pairs, int world)
{
    long long b = num_pairs / (16LL * (world > 0 ? world : 1));          // every rank owns at least 16 blocks (my-slam_b200/shard.py)
    if (b > 4096) b = 4096;
    if (b < 1) b = 1;
    return (int)b;
}

// this rank's share of the upper-triangular pair list (a < b, row-major), blocks of `block` pairs dealt round-robin
static int sharded_prepare(orbgpu_matcher* m, int rank, int world)
{
    if (m->sh_kf == m->num_kf && m->sh_rank == rank && m->sh_world == world && m->d_sh_pairs) return ORBGPU_OK;
    const long long K = m->num_kf, P = K * (K - 1) / 2;
    if (P < 1 || P > 0x7FFFFFFF) return ORBGPU_ERR_UNSUPPORTED;
    const int block = adaptive_pair_block(P, world);
    const long long nblocks = (P + block - 1) / block;
    const long long per_rank_blocks = (nblocks + world - 1) / world;
    const long long per_rank = per_rank_blocks * block;
    std::vector<int2> mine;
    mine.reserve((size_t)per_rank);
    // walk the pair list once, keeping the pairs whose block belongs to this rank
    long long p = 0;
    for (int a = 0; a < K - 1; ++a) {
        for (int b = a + 1; b < K; ++b, ++p)
            if ((p / block) % world == rank) mine.push_back(make_int2(a, b));
    }
    cudaFree(m->d_sh_pairs); cudaFree(m->d_sh_local); cudaFree(m->d_sh_all);
    m->d_sh_pairs = nullptr; m->d_sh_local = nullptr; m->d_sh_all = nullptr;
    CUM(cudaMalloc(&m->d_sh_pairs, sizeof(int2) * std::max<size_t>(mine.size(), 1)));
    CUM(cudaMalloc(&m->d_sh_local, sizeof(int32_t) * (size_t)per_rank));
    CUM(cudaMalloc(&m->d_sh_all, sizeof(int32_t) * (size_t)per_rank * world));
    if (!mine.empty()) CUM(cudaMemcpy(m->d_sh_pairs, mine.data(), sizeof(int2) * mine.size(), cudaMemcpyHostToDevice));
    m->sh_kf = m->num_kf; m->sh_rank = rank; m->sh_world = world; m->sh_block = block;
    m->sh_mine = (int)mine.size(); m->sh_per_rank = (int)per_rank;
    return ORBGPU_OK;
}

// enqueue: counts of this rank's pairs -> (gather is enqueued by the caller) -> scatter into pair order
static int sharded_enqueue_kernel(orbgpu_matcher* m, int th_low, float ratio)
{
    CUM(cudaMemsetAsync(m->d_sh_local, 0, sizeof(int32_t) * (size_t)m->sh_per_rank, m->st));
    CUM(cudaEventRecord(m->e0, m->st));
    const int qblocks = (m->n_per_kf + MATCH_THREADS - 1) / MATCH_THREADS;
    for (int p0 = 0; p0 < m->sh_mine; p0 += 65535) {
        const int np = m->sh_mine - p0 < 65535 ? m->sh_mine - p0 : 65535;
        k_match_count_pairs<<<dim3(qblocks, np), MATCH_THREADS, 0, m->st>>>(m->d_desc, m->n_per_kf, m->d_sh_pairs + p0, th_low, ratio,
                                                                           m->d_sh_local + p0);
    }
    CUM(cudaGetLastError());
    CUM(cudaEventRecord(m->e1, m->st));
    return ORBGPU_OK;
}
static int sharded_enqueue_scatter(orbgpu_matcher* m, int world, int32_t* d_counts)
{
    const long long K = m->num_kf, P = K * (K - 1) / 2;
    const int slots = m->sh_per_rank * world;
    k_scatter_pair_counts<<<(slots + 255) / 256, 256, 0, m->st>>>(world > 1 ? m->d_sh_all : m->d_sh_local, m->sh_per_rank, world, m->sh_block,
                                                                  (int)P, d_counts);
    CUM(cudaGetLastError());
    CUM(cudaEventRecord(m->e2, m->st));
    return ORBGPU_OK;
}
static int sharded_finish(orbgpu_matcher* m, double* ms3)
{
    CUM(cudaStreamSynchronize(m->st));
    float k = 0.f, g = 0.f;
    CUM(cudaEventElapsedTime(&k, m->e0, m->e1));
    CUM(cudaEventElapsedTime(&g, m->e1, m->e2));
    m->sh_kernel_ms = k; m->sh_gather_ms = g; m->last_ms = k;
    if (ms3) { ms3[0] = k; ms3[1] = g; ms3[2] = (double)k + g; }
    return ORBGPU_OK;
}

int orbgpu_matcher_all_pairs_sharded(orbgpu_matcher* m, orbgpu_comm* c, int th_low, float ratio, int32_t* d_counts, double* ms3)
{
    if (!m || !m->d_desc || !d_counts || m->num_kf < 2) return ORBGPU_ERR_ARG;
    const int rank = c ? c->rank : 0, world = c ? c->world : 1;
    if (c && c->device != m->device) return ORBGPU_ERR_ARG;
    CUM(cudaSetDevice(m->device));
    int rc;
    if ((rc = sharded_prepare(m, rank, world)) != ORBGPU_OK) return rc;
    if ((rc = sharded_enqueue_kernel(m, th_low, ratio)) != ORBGPU_OK) return rc;
    if (world > 1) {                                           // on the kernel's stream: no host synchronisation between the last tile and the gather
        const NcclApi* N = nccl_api();
        NC(N->AllGather(m->d_sh_local, m->d_sh_all, (size_t)m->sh_per_rank, ncclInt32, c->comm, m->st));
    }
    if ((rc = sharded_enqueue_scatter(m, world, d_counts)) != ORBGPU_OK) return rc;
    return sharded_finish(m, ms3);
}

int orbgpu_matcher_all_pairs_multi(orbgpu_matcher** ms, orbgpu_comm** cs, int n, int th_low, float ratio, int32_t** d_counts, double* ms3)
{
    if (!ms || !cs || !d_counts || n < 1) return ORBGPU_ERR_ARG;
    for (int r = 0; r < n; ++r)
        if (!ms[r] || !cs[r] || !d_counts[r] || !ms[r]->d_desc || cs[r]->world != n || cs[r]->rank != r || cs[r]->device != ms[r]->device ||
            ms[r]->num_kf != ms[0]->num_kf || ms[r]->n_per_kf != ms[0]->n_per_kf || ms[r]->num_kf < 2)
            return ORBGPU_ERR_ARG;
    int rc;
    for (int r = 0; r < n; ++r) {                              // every GPU starts its own pairs
        CUM(cudaSetDevice(ms[r]->device));
        if ((rc = sharded_prepare(ms[r], r, n)) != ORBGPU_OK) return rc;
        if ((rc = sharded_enqueue_kernel(ms[r], th_low, ratio)) != ORBGPU_OK) return rc;
    }
    if (n > 1) {                                               // one grouped all_gather, each rank's part on that rank's stream
        const NcclApi* N = nccl_api();
        NC(N->GroupStart());
        for (int r = 0; r < n; ++r) {
            const ncclResult_t g = N->AllGather(ms[r]->d_sh_local, ms[r]->d_sh_all, (size_t)ms[r]->sh_per_rank, ncclInt32, cs[r]->comm, ms[r]->st);
            if (g != ncclSuccess) { N->GroupEnd(); NC(g); }
        }
        NC(N->GroupEnd());
    }
    for (int r = 0; r < n; ++r) {
        CUM(cudaSetDevice(ms[r]->device));
        if ((rc = sharded_enqueue_scatter(ms[r], n, d_counts[r])) != ORBGPU_OK) return rc;
    }
    double worst[3] = {0, 0, 0};
    for (int r = 0; r < n; ++r) {
        CUM(cudaSetDevice(ms[r]->device));
        double t[3];
        if ((rc = sharded_finish(ms[r], t)) != ORBGPU_OK) return rc;
        for (int k = 0; k < 3; ++k) worst[k] = t[k] > worst[k] ? t[k] : worst[k];
    }
    if (ms3) { ms3[0] = worst[0]; ms3[1] = worst[1]; ms3[2] = worst[2]; }
    return ORBGPU_OK;
}

double orbgpu_matcher_last_kernel_ms(orbgpu_matcher* m) { return m ? m->last_ms : 0.0; }

}  // extern "C"
