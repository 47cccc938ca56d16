// orb_search.cu -- the grid-gated searches between two frames:
//   Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea     src/Frame.cc:230-245, 327-391
//   ORBmatcher::SearchForInitialization                             src/ORBmatcher.cc:405-520
//   ORBmatcher::SearchByProjection(Frame&, const Frame&, th, mono)  src/ORBmatcher.cc:1328-1470
//   ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th)  src/ORBmatcher.cc:44-125 (local-map search)
//   ORBmatcher::SearchForTriangulation + CheckDistEpipolarLine      src/ORBmatcher.cc:657-820, 140-157
//   Frame::isInFrustum + MapPoint::PredictScale                     src/Frame.cc:269-325, src/MapPoint.cc:402-417
//   the search of ORBmatcher::Fuse(KeyFrame*, vector<MapPoint*>, th) src/ORBmatcher.cc:825-957, src/KeyFrame.cc:569-613
//   ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)  src/ORBmatcher.cc:1472-1597 (relocalisation)
//
// Device layout.  The 64 x 48 grid of the searched frame is a counting sort of its keypoint indices by cell id
// ix * 48 + iy with the index order kept inside a cell (= the push_back order of AssignFeaturesToGrid); the cells
// (ix, iyMin..iyMax) of one window column are then ONE contiguous range, so a window is nMaxCellX - nMinCellX + 1
// ranges visited in the reference's order.  Three steps per call:
//   k_grid_build        one CTA: cell ids, per-cell counts, exclusive scan, in-cell rank by index (also as the last block of
//                       the kernels that compute the queries: it does not depend on them)
//   k_area_candidates   one warp per query: walks the window's column ranges laid end to end, keeps order with ballot prefix
//                       sums, evaluates the Hamming distance of every candidate and stores (index, distance) lists
//   k_*_resolve         one CTA stages lists + state in shared memory and applies the reference's order-dependent rules
//                       (mvpMapPoints occupancy: k_proj_resolve / k_local_resolve iterate the rule to its fixed point, all
//                       queries at once; vMatchedDistance / vnMatches21 stealing: k_init_resolve walks the queries in
//                       index order, 1..32 per pass) -- bit-exact with the sequential loops either way
// The rotation histogram + ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642) finish on the host, O(matches).
#include <cuda_runtime.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <time.h>
#include <string.h>
#include <map>
#include <memory>
#include <mutex>
#include <new>
#include <algorithm>
#include <vector>
#include "../../include/orbgpu.h"
#include "orb_common.cuh"

void orbgpu_internal_set_error(const char* s);

#define GRID_COLS 64                                           // include/Frame.h:37-38
#define GRID_ROWS 48
#define GRID_CELLS (GRID_COLS * GRID_ROWS)
#define GRID_MAX_KEYS 16384
#define CAND_DIST_BITS 9

struct GridDev {
    const float* keys;            // 28-byte records (x, y, size, angle, response, octave, class_id)
    const uint8_t* desc;
    int n;
    float minX, minY, invW, invH;
    float winX, winY;             // origin the search window is measured from: minX, minY for a Frame, their int truncation for a KeyFrame
    int* cell_start;              // GRID_CELLS + 1
    int* sorted;                  // n
};

struct AreaQuery {                // one GetFeaturesInArea call + the descriptor compared against its result
    float x, y, r;
    int min_level, max_level;
    int active;
    float ur;                     // SearchByProjection: u - mbf * invzc for the mvuRight check (:1408-1414)
    int check_ur;
};

__device__ __forceinline__ int key_octave(const float* keys, int i) { return __float_as_int(keys[7 * (size_t)i + 5]); }

// Frame::AssignFeaturesToGrid (src/Frame.cc:230-245) as a counting sort by cell: one CTA; dynamic shared memory holds the
// cell of every key and the keys in cell order (2 x 2 bytes x n).  Order inside a cell = index order (the push_back order of
// the reference): keys are dropped into their cell's range in arrival order and every key then takes its place by its rank
// among the keys of its own range -- shared-memory reads only, the same code for a cell of two keys and a cell of two hundred.
#define GRID_BUILD_SMEM (4 * GRID_MAX_KEYS)
__device__ __forceinline__ void grid_build_cta(const GridDev& G)      // one CTA of 1024 threads, 4 * G.n bytes of dynamic shared memory
{
    extern __shared__ __align__(16) short grid_smem[];
    __shared__ int s_cnt[GRID_CELLS];                          // counts, then the running end of every cell's range
    __shared__ int s_start[GRID_CELLS];
    __shared__ int s_warp[32];
    short* s_cell = grid_smem;                                 // [n] cell of key i, -1 = outside the grid
    short* s_sorted = grid_smem + G.n;                         // [n] keys in cell order
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int c = t; c < GRID_CELLS; c += 1024) s_cnt[c] = 0;
    __syncthreads();
    for (int i = t; i < G.n; i += 1024) {
        // PosInGrid (src/Frame.cc:383-391): round() of the float product, keypoints outside the grid are left out
        const float fx = roundf(__fmul_rn(__fsub_rn(G.keys[7 * (size_t)i], G.minX), G.invW));
        const float fy = roundf(__fmul_rn(__fsub_rn(G.keys[7 * (size_t)i + 1], G.minY), G.invH));
        int c = -1;
        if (fx >= 0.f && fx < (float)GRID_COLS && fy >= 0.f && fy < (float)GRID_ROWS) c = (int)fx * GRID_ROWS + (int)fy;
        s_cell[i] = (short)c;
        if (c >= 0) atomicAdd(&s_cnt[c], 1);
    }
    __syncthreads();
    // exclusive scan of the 3072 counts: 3 per thread, a shuffle scan per warp and one over the warp totals
    const int c0 = s_cnt[3 * t], c1 = s_cnt[3 * t + 1], c2 = s_cnt[3 * t + 2];
    int incl = c0 + c1 + c2;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += u; }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = s_warp[lane];
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += u; }
        s_warp[lane] = w;
    }
    __syncthreads();
    const int base = (warp ? s_warp[warp - 1] : 0) + incl - (c0 + c1 + c2);
    const int total = s_warp[31];
    s_cnt[3 * t] = base; s_cnt[3 * t + 1] = base + c0; s_cnt[3 * t + 2] = base + c0 + c1;
    s_start[3 * t] = base; s_start[3 * t + 1] = base + c0; s_start[3 * t + 2] = base + c0 + c1;
    G.cell_start[3 * t] = base; G.cell_start[3 * t + 1] = base + c0; G.cell_start[3 * t + 2] = base + c0 + c1;
    if (t == 1023) G.cell_start[GRID_CELLS] = total;
    __syncthreads();
    for (int i = t; i < G.n; i += 1024) {
        const int c = s_cell[i];
        if (c >= 0) s_sorted[atomicAdd(&s_cnt[c], 1)] = (short)i;     // s_cnt[c] runs from the cell's start to its end
    }
    __syncthreads();
    for (int k = t; k < total; k += 1024) {
        const int v = s_sorted[k], c = s_cell[v];
        const int start = s_start[c], end = s_cnt[c];
        int rank = 0;
        for (int j = start; j < end; ++j) rank += s_sorted[j] < v;
        G.sorted[start + rank] = v;
    }
}

__global__ void __launch_bounds__(1024) k_grid_build(GridDev G) { grid_build_cta(G); }

// The kernels that turn map points into window queries do not depend on the grid of the searched frame: launched with
// 1024-thread blocks and one block more than the queries need, that last block builds the grid (one launch and one kernel
// boundary less per call).
#define GRID_RIDER_BLOCK(G)                                              \
    if (blockIdx.x == gridDim.x - 1) { grid_build_cta(G); return; }

__device__ __forceinline__ int hamming_regs(const uint4& a0, const uint4& a1, const uint4* __restrict__ b)
{
    const uint4 b0 = __ldg(b), b1 = __ldg(b + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// One warp per query.  cand[q * cap + k] = (i2 << 9) | dist for the k-th index GetFeaturesInArea returns
// (src/Frame.cc:327-380), after the order-independent mvuRight check of SearchByProjection.
__global__ void __launch_bounds__(256) k_area_candidates(GridDev G, const AreaQuery* __restrict__ Q, const uint8_t* __restrict__ qdesc,
                                                          int nq, const float* __restrict__ u_right, uint32_t* __restrict__ cand, int cap,
                                                          int* __restrict__ cand_n)
{
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (q >= nq) return;
    const AreaQuery A = Q[q];
    int total = 0;
    if (A.active) {
        const float x = A.x, y = A.y, r = A.r;
        const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, G.winX), r), G.invW)));
        const int nMaxCellX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, G.winX), r), G.invW)));
        const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, G.winY), r), G.invH)));
        const int nMaxCellY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, G.winY), r), G.invH)));
        if (nMinCellX < GRID_COLS && nMaxCellX >= 0 && nMinCellY < GRID_ROWS && nMaxCellY >= 0) {
            const bool bCheckLevels = (A.min_level > 0) || (A.max_level >= 0);
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(qdesc) + 2 * (size_t)q);
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(qdesc) + 2 * (size_t)q + 1);
            uint32_t* out = cand + (size_t)q * cap;
            // the window's columns (a column of cells is one contiguous range of the sorted keys), 32 at a time: lane = column
            // loads its range, a shuffle scan lays the ranges end to end, and the lanes then walk the concatenation -- one chain
            // of dependent loads (range -> key index -> key, descriptor) for the whole window instead of one per column
            for (int cb = nMinCellX; cb <= nMaxCellX && nMinCellY <= nMaxCellY; cb += 32) {
                const int ncols = min(32, nMaxCellX - cb + 1);
                int p0 = 0, len = 0;
                if (lane < ncols) {
                    p0 = G.cell_start[(cb + lane) * GRID_ROWS + nMinCellY];
                    len = G.cell_start[(cb + lane) * GRID_ROWS + nMaxCellY + 1] - p0;
                }
                int incl = len;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += u; }
                const int all = __shfl_sync(0xffffffffu, incl, 31);
                for (int base = 0; base < all; base += 32) {
                    const int idx = base + lane;
                    int col = 0;
                    for (int c = 0; c + 1 < ncols; ++c) col += idx >= __shfl_sync(0xffffffffu, incl, c);
                    const int cp0 = __shfl_sync(0xffffffffu, p0, col), cin = __shfl_sync(0xffffffffu, incl, col), cl = __shfl_sync(0xffffffffu, len, col);
                    const int p = cp0 + idx - (cin - cl);
                    bool ok = idx < all;
                    int i2 = 0;
                    if (ok) {
                        i2 = G.sorted[p];
                        const float kx = G.keys[7 * (size_t)i2], ky = G.keys[7 * (size_t)i2 + 1];
                        if (bCheckLevels) {
                            const int oct = key_octave(G.keys, i2);
                            if (oct < A.min_level) ok = false;
                            if (A.max_level >= 0 && oct > A.max_level) ok = false;
                        }
                        const float distx = __fsub_rn(kx, x), disty = __fsub_rn(ky, y);
                        if (!(fabsf(distx) < r && fabsf(disty) < r)) ok = false;
                    }
                    if (ok && A.check_ur) {                       // src/ORBmatcher.cc:1408-1414
                        const float ur2 = u_right[i2];
                        if (ur2 > 0.f && fabsf(__fsub_rn(A.ur, ur2)) > r) ok = false;
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, ok);
                    if (ok) {
                        const int d = hamming_regs(a0, a1, reinterpret_cast<const uint4*>(G.desc) + 2 * (size_t)i2);
                        out[total + __popc(m & ((1u << lane) - 1u))] = ((uint32_t)i2 << CAND_DIST_BITS) | (uint32_t)d;
                    }
                    total += __popc(m);
                }
            }
        }
    }
    if (lane == 0) cand_n[q] = total;
}

// ---- in-order resolution -------------------------------------------------------------------------------------------
// One CTA.  All 1024 threads first stage the walk's state in shared memory -- per-query list offsets and counts, the
// order-dependent tables, and (when they fit) the compacted candidate lists themselves -- then warp 0 walks the
// queries in index order; a query costs a few shared-memory round trips and two warp reductions instead of a chain
// of dependent global loads.
#define RESOLVE_THREADS 1024
#define RESOLVE_SMEM (200 * 1024)

struct ResolveStage {
    int* off;                     // [nq] start of query q's list, relative to `base`
    unsigned short* cnt;          // [nq]
    const uint32_t* base;         // shared-memory copy of the lists, or the global buffer when they do not fit
};

// carve-out after `used` bytes of the dynamic shared memory; every thread returns the same stage
__device__ ResolveStage stage_lists(unsigned char* smem, size_t used, const uint32_t* __restrict__ cand, const int* __restrict__ cand_n,
                                    int cap, int nq)
{
    __shared__ int s_warp[RESOLVE_THREADS / 32];
    __shared__ int s_total;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    ResolveStage S;
    S.off = reinterpret_cast<int*>(smem + used); used += sizeof(int) * (size_t)nq;
    S.cnt = reinterpret_cast<unsigned short*>(smem + used); used += sizeof(unsigned short) * (size_t)nq;
    used = (used + 15) & ~(size_t)15;
    uint32_t* s_cand = reinterpret_cast<uint32_t*>(smem + used);
    const int room = (int)((RESOLVE_SMEM - used) / sizeof(uint32_t));
    // exclusive prefix of the list lengths: `per` consecutive queries per thread, then one shuffle scan per warp and one
    // over the warp totals (three barriers)
    const int per = (nq + RESOLVE_THREADS - 1) / RESOLVE_THREADS;
    const int q0 = min(t * per, nq), q1 = min(q0 + per, nq);
    int sum = 0;
    for (int q = q0; q < q1; ++q) sum += cand_n[q];
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += u; }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = s_warp[lane];                                  // RESOLVE_THREADS / 32 == 32 warp totals
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int u = __shfl_up_sync(0xffffffffu, w, d); if (lane >= d) w += u; }
        s_warp[lane] = w;
        if (lane == 31) s_total = w;
    }
    __syncthreads();
    const bool fits = s_total <= room;
    int run = (warp ? s_warp[warp - 1] : 0) + incl - sum;
    for (int q = q0; q < q1; ++q) {
        const int c = cand_n[q];
        S.off[q] = fits ? run : q * cap;
        S.cnt[q] = (unsigned short)c;
        run += c;
    }
    __syncthreads();
    if (fits) {
        // one thread per query: its loads are independent of each other, so a list costs one memory round trip, and all
        // lists are in flight at once (a warp per list walked 1/32 of the queries one dependent round trip after the other)
        for (int q = t; q < nq; q += RESOLVE_THREADS) {
            const int c = S.cnt[q];
            uint32_t* dst = s_cand + S.off[q];
            const uint32_t* src = cand + (size_t)q * cap;
#pragma unroll 4
            for (int p = 0; p < c; ++p) dst[p] = src[p];
        }
    }
    S.base = fits ? s_cand : cand;
    __syncthreads();
    return S;
}

// ---- the speculative in-order walk ---------------------------------------------------------------------------------
// The reference visits the queries strictly in index order because an accepted match changes what later queries may
// take (vMatchedDistance / mvpMapPoints occupancy).  Warp 0 keeps that order exactly but works on 32 consecutive queries
// at a time, one per lane: every lane scans its own candidate list against the committed state and decides; the
// features the deciding lanes would take are posted in `taker` (lowest lane per feature, found with match.any, so a byte
// per feature and no atomics).  Taking a feature only ever removes it from later queries' eligible sets, and a query's
// decision depends on its eligible set only through its best and second-best candidate, so a lane may have decided on
// stale state -- and is invalid -- only when an EARLIER lane posted its best or its second-best feature.  The lanes
// before the first invalid one are committed together (their scans saw exactly the state the sequential loop would have
// shown them), and the window restarts at the first invalid query.  Lane 0 is never invalid, so every pass commits at
// least one query; scenes without contention commit 32 per pass.
#define TAKER_NONE 64

// posts `lane` in taker[take] for the lowest lane among those that want the same feature
__device__ __forceinline__ void post_taker(unsigned char* taker, bool wants, int take, int lane)
{
    const unsigned same = __match_any_sync(0xffffffffu, wants ? take : -1 - lane);
    if (wants && (same & ((1u << lane) - 1u)) == 0) taker[take] = (unsigned char)lane;
    __syncwarp();
}

// SearchForInitialization, the loop over i1 (src/ORBmatcher.cc:418-484) on the candidate lists.
// accepted[i1] = bestIdx2 at the moment i1 was accepted (feeds rotHist), else -1.
__global__ void __launch_bounds__(RESOLVE_THREADS) k_init_resolve(const uint32_t* __restrict__ cand, const int* __restrict__ cand_n, int cap,
                                                                  int n1, int n2, float nnratio, int* __restrict__ matches12,
                                                                  int* __restrict__ accepted, int* __restrict__ nmatches_out)
{
    extern __shared__ __align__(16) unsigned char smem[];
    const int t = threadIdx.x;
    unsigned short* matched_dist = reinterpret_cast<unsigned short*>(smem);                            // vMatchedDistance, 0xFFFF = INT_MAX
    short* matches21 = reinterpret_cast<short*>(smem + 2 * (size_t)n2);                                // vnMatches21
    unsigned char* taker = smem + 4 * (size_t)n2;
    for (int i = t; i < n2; i += RESOLVE_THREADS) { taker[i] = TAKER_NONE; matched_dist[i] = 0xFFFFu; matches21[i] = -1; }
    for (int i = t; i < n1; i += RESOLVE_THREADS) { matches12[i] = -1; accepted[i] = -1; }
    const ResolveStage S = stage_lists(smem, (5 * (size_t)n2 + 15) & ~(size_t)15, cand, cand_n, cap, n1);
    if (t >= 32) return;
    const int lane = t;
    int nmatches = 0;
    for (int qb = 0; qb < n1;) {
        const int i1 = qb + lane;
        const int cnt = i1 < n1 ? S.cnt[i1] : 0;
        const uint32_t* c = S.base + (i1 < n1 ? S.off[i1] : 0);
        unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;           // (dist << 16) | position: earliest candidate wins ties
        for (int p = 0; p < cnt; ++p) {
            const uint32_t e = c[p];
            const unsigned dist = e & ((1u << CAND_DIST_BITS) - 1u);
            if ((unsigned)matched_dist[e >> CAND_DIST_BITS] <= dist) continue;      // :444-445
            const unsigned key = (dist << 16) | (unsigned)p;
            if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
        }
        int take = -1, bestDist = 0;
        if (k1 != 0xFFFFFFFFu) {
            bestDist = (int)(k1 >> 16);
            const int bestDist2 = k2 == 0xFFFFFFFFu ? INT_MAX : (int)(k2 >> 16);
            if (bestDist <= 50 && (float)bestDist < __fmul_rn((float)bestDist2, nnratio))             // TH_LOW, :459-461
                take = (int)(c[k1 & 0xFFFFu] >> CAND_DIST_BITS);
        }
        const unsigned takers = __ballot_sync(0xffffffffu, take >= 0);
        int ncommit = 32;
        if (takers) {                                          // later lanes may have decided on stale state
            post_taker(taker, take >= 0, take, lane);
            bool invalid = false;
            if (k1 != 0xFFFFFFFFu) invalid = taker[c[k1 & 0xFFFFu] >> CAND_DIST_BITS] < lane;
            if (k2 != 0xFFFFFFFFu) invalid = invalid || taker[c[k2 & 0xFFFFu] >> CAND_DIST_BITS] < lane;
            const unsigned inv = __ballot_sync(0xffffffffu, invalid);
            if (inv) ncommit = __ffs(inv) - 1;
            if (take >= 0) taker[take] = TAKER_NONE;
        }
        const bool commit = take >= 0 && lane < ncommit;
        bool fresh = false;
        if (commit) {                                          // committed lanes hold distinct features
            const int prev = matches21[take];
            if (prev >= 0) matches12[prev] = -1;               // :463-468: the earlier owner loses the feature
            else fresh = true;
            matches12[i1] = take;
            matches21[take] = (short)i1;
            matched_dist[take] = (unsigned short)bestDist;
            accepted[i1] = take;
        }
        nmatches += __popc(__ballot_sync(0xffffffffu, fresh));
        __syncwarp();
        qb += ncommit;
    }
    if (lane == 0) *nmatches_out = nmatches;
}

// SearchByProjection(Frame, Frame): projection of the last frame's map points (src/ORBmatcher.cc:1361-1391)
struct ProjCam { float R[9], t[3], fx, fy, cx, cy, bf, minX, maxX, minY, maxY, th; int forward, backward; };

__global__ void __launch_bounds__(1024) k_proj_queries(ProjCam C, const float* __restrict__ world_pos, const uint8_t* __restrict__ valid,
                                                       const int* __restrict__ last_octave, const float* __restrict__ scale, int n,
                                                       AreaQuery* __restrict__ Q, GridDev G)
{
    GRID_RIDER_BLOCK(G)
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    AreaQuery q;
    q.x = q.y = q.r = q.ur = 0.f; q.min_level = q.max_level = -1; q.active = 0; q.check_ur = 1;
    if (valid[i]) {
        const float X = world_pos[3 * (size_t)i], Y = world_pos[3 * (size_t)i + 1], Z = world_pos[3 * (size_t)i + 2];
        float c[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {                          // cv::gemm 3x3 float case: float dot, then double alpha/beta step
            const float t = __fadd_rn(__fadd_rn(__fmul_rn(C.R[3 * r], X), __fmul_rn(C.R[3 * r + 1], Y)), __fmul_rn(C.R[3 * r + 2], Z));
            c[r] = (float)__dadd_rn(__dmul_rn((double)t, 1.0), __dmul_rn((double)C.t[r], 1.0));
        }
        const float invzc = (float)__ddiv_rn(1.0, (double)c[2]);
        if (!(invzc < 0.f)) {
            const float u = __fadd_rn(__fmul_rn(__fmul_rn(C.fx, c[0]), invzc), C.cx);
            const float v = __fadd_rn(__fmul_rn(__fmul_rn(C.fy, c[1]), invzc), C.cy);
            if (!(u < C.minX || u > C.maxX) && !(v < C.minY || v > C.maxY)) {
                const int oct = last_octave[i];
                q.x = u; q.y = v; q.r = __fmul_rn(C.th, scale[oct]);
                if (C.forward) { q.min_level = oct; q.max_level = -1; }
                else if (C.backward) { q.min_level = 0; q.max_level = oct; }
                else { q.min_level = oct - 1; q.max_level = oct + 1; }
                q.ur = __fsub_rn(u, __fmul_rn(C.bf, invzc));
                q.active = 1;
            }
        }
    }
    Q[i] = q;
}

// ---- the in-order rule as a fixed point ----------------------------------------------------------------------------
// SearchByProjection visits its queries in index order because an accepted match occupies its feature for every later
// query (mvpMapPoints[i2] with Observations() > 0).  Instead of walking the queries, the two kernels below iterate: every
// query at once (one thread per query, the whole CTA) scans its candidate list for the best (and second best) feature
// that is neither occupied from the start nor currently taken by an EARLIER blocking query, and decides; taken[f] = the
// lowest index of a blocking query that takes f is rebuilt with shared-memory atomicMin; repeat until no decision
// changes.  After round k the first k queries hold their final decision (query 0 is never blocked, and a query whose
// predecessors are final sees exactly the sequential state), so the loop stops after at most n + 1 rounds in a state where
// every decision is consistent with all earlier ones: the sequential result.  Decisions only change when a predecessor
// took one of the query's best features, so real scenes need a handful of rounds, each a parallel scan -- against one
// dependent pass per 1..32 queries for the speculative walk this replaces (k_init_resolve above still walks: there a
// feature's state is a distance that falls with every new owner, not a flag).

// the loop over the last frame's points (src/ORBmatcher.cc:1353-1443) on the candidate lists
__global__ void __launch_bounds__(RESOLVE_THREADS) k_proj_resolve(const uint32_t* __restrict__ cand, const int* __restrict__ cand_n, int cap,
                                                                  int n_last, int n_cur, const uint8_t* __restrict__ observed,
                                                                  const uint8_t* __restrict__ occupied_in, int max_dist,
                                                                  int* __restrict__ cur_match, int* __restrict__ accepted,
                                                                  int* __restrict__ nmatches_out)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int s_count;
    const int t = threadIdx.x;
    int* taken = reinterpret_cast<int*>(smem);                 // lowest blocking query that takes the feature; -1 = occupied from the start
    unsigned char* occ0 = smem + 4 * (size_t)n_cur;            // mvpMapPoints[i2] && Observations() > 0 on entry (:1403-1405)
    unsigned char* obs = occ0 + (size_t)n_cur;
    size_t used = (5 * (size_t)n_cur + (size_t)n_last + 1) & ~(size_t)1;
    short* take = reinterpret_cast<short*>(smem + used);       // feature the query currently takes, -1 = none
    used = (used + 2 * (size_t)n_last + 15) & ~(size_t)15;
    for (int i = t; i < n_cur; i += RESOLVE_THREADS) { const unsigned char o = occupied_in[i]; occ0[i] = o; taken[i] = o ? -1 : INT_MAX; cur_match[i] = -1; }
    for (int i = t; i < n_last; i += RESOLVE_THREADS) { obs[i] = observed ? observed[i] : 1; take[i] = -1; }
    if (t == 0) s_count = 0;
    const ResolveStage S = stage_lists(smem, used, cand, cand_n, cap, n_last);
    for (int round = 0; round <= n_last; ++round) {
        bool changed = false;
        for (int i = t; i < n_last; i += RESOLVE_THREADS) {
            const int cnt = S.cnt[i];
            const uint32_t* c = S.base + S.off[i];
            unsigned best = 0xFFFFFFFFu;                       // (distance << 16) | list position: the earliest wins ties
            for (int p = 0; p < cnt; ++p) {
                const uint32_t e = c[p];
                if (taken[e >> CAND_DIST_BITS] < i) continue;
                best = min(best, ((e & ((1u << CAND_DIST_BITS) - 1u)) << 16) | (unsigned)p);
            }
            int tk = -1;                                       // bestDist < 256 by construction; max_dist = TH_HIGH or ORBdist
            if (best != 0xFFFFFFFFu && (int)(best >> 16) <= max_dist) tk = (int)(c[best & 0xFFFFu] >> CAND_DIST_BITS);
            if (tk != (int)take[i]) { take[i] = (short)tk; changed = true; }
        }
        if (!__syncthreads_or(changed)) break;
        for (int f = t; f < n_cur; f += RESOLVE_THREADS) taken[f] = occ0[f] ? -1 : INT_MAX;
        __syncthreads();
        for (int i = t; i < n_last; i += RESOLVE_THREADS)
            if (take[i] >= 0 && obs[i]) atomicMin(&taken[take[i]], i);          // an unobserved point leaves the feature free (:1431)
        __syncthreads();
    }
    int mine = 0;
    for (int i = t; i < n_last; i += RESOLVE_THREADS) {
        const int tk = take[i];
        accepted[i] = tk;
        if (tk >= 0) { atomicMax(&cur_match[tk], i); ++mine; }                  // the last point in index order keeps the feature
    }
    if (mine) atomicAdd(&s_count, mine);
    __syncthreads();
    if (t == 0) *nmatches_out = s_count;
}

// SearchByProjection(Frame, vector<MapPoint*>): the loop over the local map points (src/ORBmatcher.cc:50-122) on the
// candidate lists.  Best and second best keep the octave of the keypoint they came from (bestLevel / bestLevel2); the
// sequential scan's second best is the minimum of (distance, position) over everything but the best.
__global__ void __launch_bounds__(RESOLVE_THREADS) k_local_resolve(const uint32_t* __restrict__ cand, const int* __restrict__ cand_n, int cap,
                                                                   int n_points, int n_cur, const uint8_t* __restrict__ observed,
                                                                   const uint8_t* __restrict__ occupied_in, const float* __restrict__ keys,
                                                                   float nnratio, int* __restrict__ cur_match, int* __restrict__ nmatches_out)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int s_count;
    const int t = threadIdx.x;
    int* taken = reinterpret_cast<int*>(smem);                 // lowest blocking point that takes the feature; -1 = occupied from the start
    unsigned char* occ0 = smem + 4 * (size_t)n_cur;            // F.mvpMapPoints[idx] && Observations() > 0 on entry (:89-91)
    unsigned char* oct = occ0 + (size_t)n_cur;                 // F.mvKeysUn[idx].octave
    unsigned char* obs = oct + (size_t)n_cur;
    size_t used = (6 * (size_t)n_cur + (size_t)n_points + 1) & ~(size_t)1;
    short* take = reinterpret_cast<short*>(smem + used);       // feature the point currently takes, -1 = none
    used = (used + 2 * (size_t)n_points + 15) & ~(size_t)15;
    for (int i = t; i < n_cur; i += RESOLVE_THREADS) {
        const unsigned char o = occupied_in[i];
        occ0[i] = o; taken[i] = o ? -1 : INT_MAX; oct[i] = (unsigned char)key_octave(keys, i); cur_match[i] = -1;
    }
    for (int i = t; i < n_points; i += RESOLVE_THREADS) { obs[i] = observed[i]; take[i] = -1; }
    if (t == 0) s_count = 0;
    const ResolveStage S = stage_lists(smem, used, cand, cand_n, cap, n_points);
    for (int round = 0; round <= n_points; ++round) {          // the fixed point of the in-order rule, see k_proj_resolve
        bool changed = false;
        for (int i = t; i < n_points; i += RESOLVE_THREADS) {
            const int cnt = S.cnt[i];
            const uint32_t* c = S.base + S.off[i];
            unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;
            for (int p = 0; p < cnt; ++p) {
                const uint32_t e = c[p];
                if (taken[e >> CAND_DIST_BITS] < i) continue;
                const unsigned key = ((e & ((1u << CAND_DIST_BITS) - 1u)) << 16) | (unsigned)p;
                if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
            }
            int tk = -1;
            if (k1 != 0xFFFFFFFFu && (int)(k1 >> 16) <= 100) { // TH_HIGH (:113); a distance of 256 never replaces the initial 256
                const int bestDist = (int)(k1 >> 16);
                const int bestIdx = (int)(c[k1 & 0xFFFFu] >> CAND_DIST_BITS);
                int bestDist2 = 256, bestLevel2 = -1;
                if (k2 != 0xFFFFFFFFu && (int)(k2 >> 16) < 256) { bestDist2 = (int)(k2 >> 16); bestLevel2 = oct[c[k2 & 0xFFFFu] >> CAND_DIST_BITS]; }
                if (!((int)oct[bestIdx] == bestLevel2 && (float)bestDist > __fmul_rn(nnratio, (float)bestDist2))) tk = bestIdx;    // :115-116
            }
            if (tk != (int)take[i]) { take[i] = (short)tk; changed = true; }
        }
        if (!__syncthreads_or(changed)) break;
        for (int f = t; f < n_cur; f += RESOLVE_THREADS) taken[f] = occ0[f] ? -1 : INT_MAX;
        __syncthreads();
        for (int i = t; i < n_points; i += RESOLVE_THREADS)
            if (take[i] >= 0 && obs[i]) atomicMin(&taken[take[i]], i);          // an unobserved point does not reserve its feature (:89-91)
        __syncthreads();
    }
    int mine = 0;
    for (int i = t; i < n_points; i += RESOLVE_THREADS) {
        const int tk = take[i];
        if (tk >= 0) { atomicMax(&cur_match[tk], i); ++mine; }                  // the last point in index order keeps the feature
    }
    if (mine) atomicAdd(&s_count, mine);
    __syncthreads();
    if (t == 0) *nmatches_out = s_count;
}

// ---- SearchByBoW (src/ORBmatcher.cc:159-288, :522-655) -----------------------------------------------------------
// A query = one valid feature of side 1 inside a vocabulary node both sides share; its candidates = the features of
// side 2 in that node, in the FeatureVector's order.  Queries are listed by the host in the reference's visiting
// order (nodes ascending, then the node's index list).
struct BowQuery { int idx1, start, len; };                    // candidates = indices2[start .. start + len)

__global__ void __launch_bounds__(256) k_bow_candidates(const BowQuery* __restrict__ Q, int nq, const uint8_t* __restrict__ desc1,
                                                        const uint8_t* __restrict__ desc2, const uint32_t* __restrict__ indices2,
                                                        uint32_t* __restrict__ cand, int cap, int* __restrict__ cand_n)
{
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (q >= nq) return;
    const BowQuery B = Q[q];
    const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)B.idx1);
    const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)B.idx1 + 1);
    for (int p = lane; p < B.len; p += 32) {
        const uint32_t i2 = indices2[B.start + p];
        const int d = hamming_regs(a0, a1, reinterpret_cast<const uint4*>(desc2) + 2 * (size_t)i2);
        cand[(size_t)q * cap + p] = (i2 << CAND_DIST_BITS) | (uint32_t)d;
    }
    if (lane == 0) cand_n[q] = B.len;
}

// the in-order part: vbMatched2 / vpMapPointMatches exclusion, best + second, TH_LOW and ratio tests.  The exclusion
// only couples queries of the same vocabulary node (a feature lies in one node), so each warp of the CTA walks the
// queries of its own nodes in order: group g = queries [gstart[g], gstart[g + 1]).
__global__ void __launch_bounds__(RESOLVE_THREADS) k_bow_resolve(const uint32_t* __restrict__ cand, const int* __restrict__ cand_n, int cap,
                                                                 int nq, int n2, const uint8_t* __restrict__ valid2,
                                                                 const int* __restrict__ gstart, int ngroups, float nnratio,
                                                                 int th_low_excl, int* __restrict__ accepted, int* __restrict__ nmatches_out)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ int s_nmatches;
    const int t = threadIdx.x;
    unsigned char* blocked = smem;                             // vbMatched2[idx2] || no usable map point
    for (int i = t; i < n2; i += RESOLVE_THREADS) blocked[i] = valid2[i] ? 0 : 1;
    for (int i = t; i < nq; i += RESOLVE_THREADS) accepted[i] = -1;
    if (t == 0) s_nmatches = 0;
    const ResolveStage S = stage_lists(smem, ((size_t)n2 + 15) & ~(size_t)15, cand, cand_n, cap, nq);
    const int lane = t & 31, warp = t >> 5;
    int nmatches = 0;
    for (int g = warp; g < ngroups; g += RESOLVE_THREADS / 32) {
        const int q1 = gstart[g + 1];
        for (int q = gstart[g]; q < q1; ++q) {
            const int cnt = S.cnt[q];
            const uint32_t* c = S.base + S.off[q];
            unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;
            for (int p = lane; p < cnt; p += 32) {
                const uint32_t e = c[p];
                if (blocked[e >> CAND_DIST_BITS]) continue;
                const unsigned key = ((e & ((1u << CAND_DIST_BITS) - 1u)) << 16) | (unsigned)p;
                if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
            }
            const unsigned b1 = __reduce_min_sync(0xffffffffu, k1);
            const unsigned b2 = __reduce_min_sync(0xffffffffu, k1 == b1 ? k2 : k1);
            // bestDist1 / bestDist2 start at 256 and only take smaller values (:199-201)
            const int bestDist1 = b1 == 0xFFFFFFFFu ? 256 : min((int)(b1 >> 16), 256);
            const int bestDist2 = b2 == 0xFFFFFFFFu ? 256 : min((int)(b2 >> 16), 256);
            if (bestDist1 < th_low_excl && (float)bestDist1 < __fmul_rn(nnratio, (float)bestDist2)) {
                const int bestIdx2 = (int)(c[b1 & 0xFFFFu] >> CAND_DIST_BITS);
                __syncwarp();
                if (lane == 0) { blocked[bestIdx2] = 1; accepted[q] = bestIdx2; }
                nmatches++;
                __syncwarp();
            }
        }
    }
    if (lane == 0 && nmatches) atomicAdd(&s_nmatches, nmatches);
    __syncthreads();
    if (t == 0) *nmatches_out = s_nmatches;
}

// Both steps in one launch for ordinary vocabularies (a node holds a handful of features): one WARP per shared node.  The
// node's features of side 2 (descriptor, index, "still free" flag) are staged in the warp's slice of shared memory once;
// the node's queries are loaded 32 at a time, one per lane, and broadcast by shuffle; each query is one XOR / POPC pass
// over the slice and two warp minima.  Nothing goes through global candidate lists or a single CTA.  Nodes with more than
// BOW_NODE_CAP features of side 2 read them from memory with the "free" flags in the uploaded copy of valid2 (a feature
// lies in one node, so only this warp touches its flag).  The host keeps the two-kernel path above for degenerate
// vocabularies (one node holding everything), where the distance evaluations need the whole GPU.
#define BOW_NODE_CAP 256
__device__ __forceinline__ uint4 shfl4(const uint4& v, int src)
{
    return make_uint4(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src), __shfl_sync(0xffffffffu, v.z, src),
                      __shfl_sync(0xffffffffu, v.w, src));
}
__device__ __forceinline__ int hamming_rr(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(128) k_bow_fused(const BowQuery* __restrict__ Q, const int* __restrict__ gstart, int ngroups,
                                                   const uint8_t* __restrict__ desc1, const uint8_t* __restrict__ desc2,
                                                   const uint32_t* __restrict__ indices2, uint8_t* __restrict__ free2, float nnratio,
                                                   int th_low_excl, int* __restrict__ accepted)
{
    __shared__ uint4 s_lo[4][BOW_NODE_CAP], s_hi[4][BOW_NODE_CAP];       // the two halves apart: conflict-free LDS.128
    __shared__ uint32_t s_i2[4][BOW_NODE_CAP];
    __shared__ uint8_t s_free[4][BOW_NODE_CAP];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int g = blockIdx.x * 4 + w;
    if (g >= ngroups) return;
    const int q0 = gstart[g], q1 = gstart[g + 1];
    if (q0 >= q1) return;
    const int start = Q[q0].start, len = Q[q0].len;            // every query of a node searches the same list
    const bool staged = len <= BOW_NODE_CAP;
    if (staged) {
        for (int p = lane; p < len; p += 32) {
            const uint32_t i2 = indices2[start + p];
            s_i2[w][p] = i2;
            s_lo[w][p] = __ldg(reinterpret_cast<const uint4*>(desc2) + 2 * (size_t)i2);
            s_hi[w][p] = __ldg(reinterpret_cast<const uint4*>(desc2) + 2 * (size_t)i2 + 1);
            s_free[w][p] = free2[i2];
        }
        __syncwarp();
    }
    for (int qb = q0; qb < q1; qb += 32) {
        uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
        if (qb + lane < q1) {
            const int idx1 = Q[qb + lane].idx1;
            a0 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)idx1);
            a1 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)idx1 + 1);
        }
        const int m = min(32, q1 - qb);
        for (int j = 0; j < m; ++j) {
            const uint4 b0 = shfl4(a0, j), b1 = shfl4(a1, j);
            unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;
            if (staged) {
                for (int p = lane; p < len; p += 32) {
                    if (!s_free[w][p]) continue;
                    const unsigned key = ((unsigned)hamming_rr(b0, b1, s_lo[w][p], s_hi[w][p]) << 16) | (unsigned)p;
                    if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
                }
            } else {
                for (int p = lane; p < len; p += 32) {
                    const uint32_t i2 = indices2[start + p];
                    if (!free2[i2]) continue;
                    const unsigned key = ((unsigned)hamming_regs(b0, b1, reinterpret_cast<const uint4*>(desc2) + 2 * (size_t)i2) << 16) | (unsigned)p;
                    if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
                }
            }
            const unsigned m1 = __reduce_min_sync(0xffffffffu, k1);
            const unsigned m2 = __reduce_min_sync(0xffffffffu, k1 == m1 ? k2 : k1);
            const int bestDist1 = m1 == 0xFFFFFFFFu ? 256 : (int)(m1 >> 16);       // bestDist1 / bestDist2 start at 256 (:199-201)
            const int bestDist2 = m2 == 0xFFFFFFFFu ? 256 : (int)(m2 >> 16);
            int res = -1;
            if (bestDist1 < th_low_excl && (float)bestDist1 < __fmul_rn(nnratio, (float)bestDist2)) {
                const int bp = (int)(m1 & 0xFFFFu);
                if (staged) {
                    res = (int)s_i2[w][bp];
                    if (lane == 0) s_free[w][bp] = 0;
                } else {
                    res = (int)indices2[start + bp];
                    if (lane == 0) free2[res] = 0;
                }
                __syncwarp();
            }
            if (lane == 0) accepted[qb + j] = res;
        }
    }
}

// The ordinary case -- every shared node has at most BOW_NODE_CAP features of side 2 and at most BOW_NODE_QUERIES queries --
// without any cross-lane reduction in the order-dependent part.  One CTA per node (its queries dealt to the warps 32 at a time):
//   1. the node's features of side 2 are staged in shared memory; the distance of every (query, feature) pair goes into a
//      shared-memory matrix, lane = query, the feature's descriptor read as a broadcast;
//   2. the reference's in-order rule "a feature accepted by an earlier query is gone" is solved as a fixed point instead of
//      a walk: every query (lane = query, all at once) takes its best / second best among the features that no EARLIER
//      query currently takes and decides; then taken[p] = the lowest query index that takes p is rebuilt; repeat until no
//      decision changes.  After round k the first k queries hold their final decision (query 0 is never blocked, and a
//      query whose predecessors are final sees exactly the sequential state), so the loop ends after at most nq + 1 rounds
//      and at a state in which every decision is consistent with all earlier ones -- the sequential result.  In practice a
//      handful of rounds: a decision only changes when a predecessor took one of its two best features.
// A round costs ~len shared-memory reads per lane and no shuffles; the warp-per-node walk above pays two warp reductions
// and ~0.4 us per query, one query after the other.
#define BOW_NODE_QUERIES 256
#define BOW_DIST_ENTRIES (BOW_NODE_QUERIES * 32 + 3 * BOW_NODE_QUERIES + 64)   // nq * (len + 3) with nq * ceil(len / 32) <= 256

#define BOW_ROUND_WARPS 4          // a node's queries are dealt to the warps 32 at a time; an ordinary node keeps one warp busy

__global__ void __launch_bounds__(32 * BOW_ROUND_WARPS) k_bow_rounds(const BowQuery* __restrict__ Q, const int* __restrict__ gstart,
                                                                     const uint8_t* __restrict__ desc1, const uint8_t* __restrict__ desc2,
                                                                     const uint32_t* __restrict__ indices2, const uint8_t* __restrict__ valid2,
                                                                     float nnratio, int th_low_excl, int* __restrict__ accepted)
{
    __shared__ uint4 s_lo[BOW_NODE_CAP], s_hi[BOW_NODE_CAP];
    __shared__ uint32_t s_i2[BOW_NODE_CAP];
    __shared__ int s_taken[BOW_NODE_CAP];                      // lowest query index that currently takes the feature; -1 = never free
    __shared__ uint8_t s_valid[BOW_NODE_CAP];
    __shared__ unsigned short s_dist[BOW_DIST_ENTRIES];
    __shared__ short s_take[BOW_NODE_QUERIES];                 // list position the query currently takes, -1 = none
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = blockIdx.x;
    const int q0 = gstart[g], nq = gstart[g + 1] - q0;
    if (nq <= 0) return;
    const int start = Q[q0].start, len = Q[q0].len;            // every query of a node searches the same list
    int stride = (len + 1) & ~1;                               // row pitch of the distance matrix in 16-bit entries:
    if (((stride >> 1) & 1) == 0) stride += 2;                 // an odd number of words, so lane = query is conflict-free
    for (int p = tid; p < len; p += 32 * BOW_ROUND_WARPS) {
        const uint32_t i2 = indices2[start + p];
        s_i2[p] = i2;
        s_lo[p] = __ldg(reinterpret_cast<const uint4*>(desc2) + 2 * (size_t)i2);
        s_hi[p] = __ldg(reinterpret_cast<const uint4*>(desc2) + 2 * (size_t)i2 + 1);
        const uint8_t v = valid2[i2];
        s_valid[p] = v;
        s_taken[p] = v ? INT_MAX : -1;
    }
    __syncthreads();
    for (int qb = 32 * warp; qb < nq; qb += 32 * BOW_ROUND_WARPS) {
        const int q = qb + lane;
        if (q < nq) {
            const int idx1 = Q[q0 + q].idx1;
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)idx1);
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)idx1 + 1);
            unsigned short* row = s_dist + q * stride;
#pragma unroll 4
            for (int p = 0; p < len; ++p) row[p] = (unsigned short)hamming_rr(a0, a1, s_lo[p], s_hi[p]);
            s_take[q] = -1;
        }
    }
    __syncthreads();
    for (int round = 0; round <= nq; ++round) {
        bool changed = false;
        for (int qb = 32 * warp; qb < nq; qb += 32 * BOW_ROUND_WARPS) {
            const int q = qb + lane;
            if (q >= nq) continue;
            const unsigned short* row = s_dist + q * stride;
            unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;       // (distance << 16) | list position: the earliest wins ties
#pragma unroll 4
            for (int p = 0; p < len; ++p) {
                const unsigned key = s_taken[p] < q ? 0xFFFFFFFFu : (((unsigned)row[p] << 16) | (unsigned)p);
                const unsigned hi = max(k1, key);              // keep the two smallest of {k1, k2, key}
                k1 = min(k1, key); k2 = min(k2, hi);
            }
            const int bestDist1 = k1 == 0xFFFFFFFFu ? 256 : (int)(k1 >> 16);       // bestDist1 / bestDist2 start at 256 (:199-201)
            const int bestDist2 = k2 == 0xFFFFFFFFu ? 256 : (int)(k2 >> 16);
            const int take = (bestDist1 < th_low_excl && (float)bestDist1 < __fmul_rn(nnratio, (float)bestDist2)) ? (int)(k1 & 0xFFFFu) : -1;
            if (take != (int)s_take[q]) { s_take[q] = (short)take; changed = true; }
        }
        if (!__syncthreads_or(changed)) break;
        for (int p = tid; p < len; p += 32 * BOW_ROUND_WARPS) s_taken[p] = s_valid[p] ? INT_MAX : -1;
        __syncthreads();
        for (int q = tid; q < nq; q += 32 * BOW_ROUND_WARPS)
            if (s_take[q] >= 0) atomicMin(&s_taken[s_take[q]], q);
        __syncthreads();
    }
    for (int q = tid; q < nq; q += 32 * BOW_ROUND_WARPS) accepted[q0 + q] = s_take[q] >= 0 ? (int)s_i2[s_take[q]] : -1;
}

// ---- Frame::isInFrustum (src/Frame.cc:269-325) ---------------------------------------------------------------------
// glibc 2.39 logf (sysdeps/ieee754/flt-32/e_logf.c, the routine the reference object calls from MapPoint::PredictScale)
// followed operation by operation in double; tests/test_local_map.py sweeps float bit patterns against the host's logf
// (the rounding to float absorbs whether the host build contracts the multiply-adds, checked over every normal float).
__constant__ double c_logf_tab[32] = {
    0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2, 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2, 0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2,
    0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3, 0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3, 0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3,
    0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4, 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4, 0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5,
    0x1p+0, 0x0p+0, 0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5, 0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4,
    0x1.b2036576afce6p-1, 0x1.526e57720db08p-3, 0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3, 0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2,
    0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2};

__device__ __forceinline__ float orb_logf(float x)
{
    uint32_t ix = __float_as_uint(x);
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {       // x < 0x1p-126, inf or nan
        if (ix * 2 == 0) return __int_as_float(0xff800000);    // log(0) = -inf
        if (ix == 0x7f800000u) return x;                       // log(inf) = inf
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return __int_as_float(0x7fc00000);
        ix = __float_as_uint(__fmul_rn(x, 0x1p23f));           // subnormal: normalise
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const int k = (int)tmp >> 23;
    const uint32_t iz = ix - (tmp & (0x1ffu << 23));
    const double invc = c_logf_tab[2 * i], logc = c_logf_tab[2 * i + 1];
    const double z = (double)__uint_as_float(iz);
    const double r = __dadd_rn(__dmul_rn(z, invc), -1.0);
    const double y0 = __dadd_rn(logc, __dmul_rn((double)k, 0x1.62e42fefa39efp-1));
    const double r2 = __dmul_rn(r, r);
    double y = __dadd_rn(__dmul_rn(0x1.5575b0be00b6ap-2, r), -0x1.ffffef20a4123p-2);
    y = __dadd_rn(__dmul_rn(-0x1.00ea348b88334p-2, r2), y);
    y = __dadd_rn(__dmul_rn(y, r2), __dadd_rn(y0, r));
    return (float)y;
}

__global__ void k_debug_logf(const float* __restrict__ x, int n, float* __restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = orb_logf(x[i]);
}

struct FrustumCam { float R[9], t[3], Ow[3], fx, fy, cx, cy, bf, minX, maxX, minY, maxY, cos_limit, log_scale, th; int nlevels; };
struct TrackOut { uint8_t* in_view; float *proj_x, *proj_y, *proj_xr, *view_cos; int* scale_level; };

__global__ void __launch_bounds__(1024) k_frustum(FrustumCam C, const uint8_t* __restrict__ candidate, const float* __restrict__ world_pos,
                                                  const float* __restrict__ normal, const float* __restrict__ max_distance,
                                                  const float* __restrict__ min_distance, const float* __restrict__ scale, int n, TrackOut T,
                                                  AreaQuery* __restrict__ Q, int* __restrict__ n_in_view, GridDev G, int with_grid)
{
    if (with_grid) { GRID_RIDER_BLOCK(G) }
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool in = false;
    if (i < n) {
        AreaQuery q;
        q.x = q.y = q.r = q.ur = 0.f; q.min_level = q.max_level = -1; q.active = 0; q.check_ur = 1;
        float u = 0.f, v = 0.f, ur = 0.f, viewCos = 0.f;
        int level = 0;
        if (candidate[i]) {
            const float P[3] = {world_pos[3 * (size_t)i], world_pos[3 * (size_t)i + 1], world_pos[3 * (size_t)i + 2]};
            float Pc[3];
#pragma unroll
            for (int r = 0; r < 3; ++r) {                      // mRcw*P+mtcw: cv::gemm's 3x3 float case (:277)
                const float t = __fadd_rn(__fadd_rn(__fmul_rn(C.R[3 * r], P[0]), __fmul_rn(C.R[3 * r + 1], P[1])), __fmul_rn(C.R[3 * r + 2], P[2]));
                Pc[r] = (float)__dadd_rn(__dmul_rn((double)t, 1.0), __dmul_rn((double)C.t[r], 1.0));
            }
            if (!(Pc[2] < 0.0f)) {                             // :283
                const float invz = __fdiv_rn(1.0f, Pc[2]);
                u = __fadd_rn(__fmul_rn(__fmul_rn(C.fx, Pc[0]), invz), C.cx);
                v = __fadd_rn(__fmul_rn(__fmul_rn(C.fy, Pc[1]), invz), C.cy);
                if (!(u < C.minX || u > C.maxX) && !(v < C.minY || v > C.maxY)) {
                    const float maxDistance = __fmul_rn(1.2f, max_distance[i]);        // src/MapPoint.cc:373-383
                    const float minDistance = __fmul_rn(0.8f, min_distance[i]);
                    const float PO[3] = {__fsub_rn(P[0], C.Ow[0]), __fsub_rn(P[1], C.Ow[1]), __fsub_rn(P[2], C.Ow[2])};
                    // cv::norm: squares summed in double, sqrt in double, narrowed by "const float dist" (:299)
                    double s2 = __dmul_rn((double)PO[0], (double)PO[0]);
                    s2 = __dadd_rn(s2, __dmul_rn((double)PO[1], (double)PO[1]));
                    s2 = __dadd_rn(s2, __dmul_rn((double)PO[2], (double)PO[2]));
                    const float dist = (float)__dsqrt_rn(s2);
                    if (!(dist < minDistance || dist > maxDistance)) {
                        // cv::Mat::dot: products summed in double; "/dist" divides in double (:307)
                        double d = __dmul_rn((double)PO[0], (double)normal[3 * (size_t)i]);
                        d = __dadd_rn(d, __dmul_rn((double)PO[1], (double)normal[3 * (size_t)i + 1]));
                        d = __dadd_rn(d, __dmul_rn((double)PO[2], (double)normal[3 * (size_t)i + 2]));
                        viewCos = (float)__ddiv_rn(d, (double)dist);
                        if (!(viewCos < C.cos_limit)) {
                            // MapPoint::PredictScale (src/MapPoint.cc:402-417)
                            const float ratio = __fdiv_rn(max_distance[i], dist);
                            const float c = ceilf(__fdiv_rn(orb_logf(ratio), C.log_scale));
                            int nScale = (fabsf(c) < 2147483648.f) ? (int)c : INT_MIN;     // cvttss2si of inf / nan
                            if (nScale < 0) nScale = 0;
                            else if (nScale >= C.nlevels) nScale = C.nlevels - 1;
                            level = nScale;
                            ur = __fsub_rn(u, __fmul_rn(C.bf, invz));
                            in = true;
                        }
                    }
                }
            }
        }
        if (!in) { u = v = ur = viewCos = 0.f; level = 0; }     // the reference leaves the members of such a point untouched
        // a point on the camera plane or at the camera centre yields 0*inf or 0/0; x86 writes its default NaN (sign set)
        if (u != u) u = __int_as_float(0xffc00000);
        if (v != v) v = __int_as_float(0xffc00000);
        if (ur != ur) ur = __int_as_float(0xffc00000);
        if (viewCos != viewCos) viewCos = __int_as_float(0xffc00000);
        T.in_view[i] = in ? 1 : 0;
        T.proj_x[i] = u; T.proj_y[i] = v; T.proj_xr[i] = ur; T.view_cos[i] = viewCos; T.scale_level[i] = level;
        if (in) {                                              // the window of src/ORBmatcher.cc:62-67
            float r = (double)viewCos > 0.998 ? 2.5f : 4.0f;
            if (C.th != 1.0f) r = __fmul_rn(r, C.th);
            q.x = u; q.y = v; q.r = __fmul_rn(r, scale[level]);
            q.min_level = level - 1; q.max_level = level; q.ur = ur; q.active = 1;
        }
        Q[i] = q;
    }
    const int cnt = __syncthreads_count(in);
    if (threadIdx.x == 0 && cnt) atomicAdd(n_in_view, cnt);
}

// ---- the search of ORBmatcher::Fuse (src/ORBmatcher.cc:825-957) -----------------------------------------------------
// mode flags: the keyframe-projection loops of ORBmatcher differ in small print only
//   two_stage   Pc = R2*(R*P + t) + t2                      SearchBySim3 (:1158-1160, :1238-1240)
//   inv_double  invz = 1.0/z (a double division)             Fuse(KF, Scw) :1018, SearchBySim3 :1166; "1/z" elsewhere
//   cam_dist    distance = norm(Pc) instead of norm(P - Ow)  SearchBySim3 :1181
//   view_check  PO.dot(Pn) < 0.5*dist rejects                Fuse :883, Fuse(Scw) :1041, SearchByProjection(KF, Scw) :351
struct FuseCam { float R[9], t[3], R2[9], t2[3], Ow[3], fx, fy, cx, cy, bf, minX, maxX, minY, maxY, log_scale, th; int nlevels;
                 int two_stage, inv_double, cam_dist, view_check; };
struct FuseQuery { float u, v, ur, r; int level; int active; };

__global__ void k_fuse_queries(FuseCam C, const uint8_t* __restrict__ present, const float* __restrict__ world_pos,
                               const float* __restrict__ normal, const float* __restrict__ max_distance,
                               const float* __restrict__ min_distance, const float* __restrict__ scale, int n, FuseQuery* __restrict__ Q,
                               AreaQuery* __restrict__ AQ)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    FuseQuery q;
    q.u = q.v = q.ur = q.r = 0.f; q.level = 0; q.active = 0;
    if (present[i]) {
        const float P[3] = {world_pos[3 * (size_t)i], world_pos[3 * (size_t)i + 1], world_pos[3 * (size_t)i + 2]};
        float Pc[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {                          // Rcw*p3Dw + tcw: cv::gemm's 3x3 float case (:854)
            const float t = __fadd_rn(__fadd_rn(__fmul_rn(C.R[3 * r], P[0]), __fmul_rn(C.R[3 * r + 1], P[1])), __fmul_rn(C.R[3 * r + 2], P[2]));
            Pc[r] = (float)__dadd_rn(__dmul_rn((double)t, 1.0), __dmul_rn((double)C.t[r], 1.0));
        }
        if (C.two_stage) {                                     // sR21*p3Dc1 + t21
            const float A[3] = {Pc[0], Pc[1], Pc[2]};
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const float t = __fadd_rn(__fadd_rn(__fmul_rn(C.R2[3 * r], A[0]), __fmul_rn(C.R2[3 * r + 1], A[1])), __fmul_rn(C.R2[3 * r + 2], A[2]));
                Pc[r] = (float)__dadd_rn(__dmul_rn((double)t, 1.0), __dmul_rn((double)C.t2[r], 1.0));
            }
        }
        if (!(Pc[2] < 0.0f)) {                                 // :857
            const float invz = C.inv_double ? (float)__ddiv_rn(1.0, (double)Pc[2]) : __fdiv_rn(1.0f, Pc[2]);
            const float x = __fmul_rn(Pc[0], invz), y = __fmul_rn(Pc[1], invz);
            const float u = __fadd_rn(__fmul_rn(C.fx, x), C.cx), v = __fadd_rn(__fmul_rn(C.fy, y), C.cy);
            if (u >= C.minX && u < C.maxX && v >= C.minY && v < C.maxY) {             // KeyFrame::IsInImage, int bounds
                const float maxDistance = __fmul_rn(1.2f, max_distance[i]);
                const float minDistance = __fmul_rn(0.8f, min_distance[i]);
                float PO[3];
                if (C.cam_dist) { PO[0] = Pc[0]; PO[1] = Pc[1]; PO[2] = Pc[2]; }
                else { PO[0] = __fsub_rn(P[0], C.Ow[0]); PO[1] = __fsub_rn(P[1], C.Ow[1]); PO[2] = __fsub_rn(P[2], C.Ow[2]); }
                double s2 = __dmul_rn((double)PO[0], (double)PO[0]);
                s2 = __dadd_rn(s2, __dmul_rn((double)PO[1], (double)PO[1]));
                s2 = __dadd_rn(s2, __dmul_rn((double)PO[2], (double)PO[2]));
                const float dist3D = (float)__dsqrt_rn(s2);
                if (!(dist3D < minDistance || dist3D > maxDistance)) {
                    bool seen = true;
                    if (C.view_check) {
                        double d = __dmul_rn((double)PO[0], (double)normal[3 * (size_t)i]);
                        d = __dadd_rn(d, __dmul_rn((double)PO[1], (double)normal[3 * (size_t)i + 1]));
                        d = __dadd_rn(d, __dmul_rn((double)PO[2], (double)normal[3 * (size_t)i + 2]));
                        seen = !(d < __dmul_rn(0.5, (double)dist3D));                  // :883
                    }
                    if (seen) {
                        const float ratio = __fdiv_rn(max_distance[i], dist3D);        // PredictScale(dist3D, pKF), src/MapPoint.cc:385-400
                        const float c = ceilf(__fdiv_rn(orb_logf(ratio), C.log_scale));
                        int nScale = (fabsf(c) < 2147483648.f) ? (int)c : INT_MIN;
                        if (nScale < 0) nScale = 0;
                        else if (nScale >= C.nlevels) nScale = C.nlevels - 1;
                        q.u = u; q.v = v; q.ur = __fsub_rn(u, __fmul_rn(C.bf, invz));
                        q.r = __fmul_rn(C.th, scale[nScale]);
                        q.level = nScale; q.active = 1;
                    }
                }
            }
        }
    }
    Q[i] = q;
    if (AQ) {                                                  // the same query for k_area_candidates (SearchByProjection(KF, Scw))
        AreaQuery a;
        a.x = q.u; a.y = q.v; a.r = q.r; a.min_level = q.level - 1; a.max_level = q.level; a.active = q.active; a.ur = 0.f; a.check_ur = 0;
        AQ[i] = a;
    }
}

// one warp per map point: KeyFrame::GetFeaturesInArea's visiting order (columns of the window, cells top to bottom, index
// order inside a cell) with the level and chi-square gates of :897-928; the smallest distance wins, the first one on ties
__global__ void __launch_bounds__(256) k_fuse_match(GridDev G, const FuseQuery* __restrict__ Q, const uint8_t* __restrict__ qdesc, int nq,
                                                    const float* __restrict__ u_right, const float* __restrict__ inv_sigma2,
                                                    int chi2_gate, unsigned max_dist, int* __restrict__ best_idx)
{
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (q >= nq) return;
    const FuseQuery A = Q[q];
    unsigned long long best = ~0ull;                           // (dist << 40) | (visit position << 16) | ... the index rides below
    if (A.active) {
        const float x = A.u, y = A.v, r = A.r;
        const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, G.winX), r), G.invW)));
        const int nMaxCellX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, G.winX), r), G.invW)));
        const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, G.winY), r), G.invH)));
        const int nMaxCellY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, G.winY), r), G.invH)));
        if (nMinCellX < GRID_COLS && nMaxCellX >= 0 && nMinCellY < GRID_ROWS && nMaxCellY >= 0 && nMinCellY <= nMaxCellY) {
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(qdesc) + 2 * (size_t)q);
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(qdesc) + 2 * (size_t)q + 1);
            int visited = 0;
            for (int ix = nMinCellX; ix <= nMaxCellX; ++ix) {
                const int p0 = G.cell_start[ix * GRID_ROWS + nMinCellY], p1 = G.cell_start[ix * GRID_ROWS + nMaxCellY + 1];
                for (int pb = p0; pb < p1; pb += 32) {
                    const int p = pb + lane;
                    if (p < p1) {
                        const int i2 = G.sorted[p];
                        const float kx = G.keys[7 * (size_t)i2], ky = G.keys[7 * (size_t)i2 + 1];
                        const int oct = key_octave(G.keys, i2);
                        bool ok = fabsf(__fsub_rn(kx, x)) < r && fabsf(__fsub_rn(ky, y)) < r;       // src/KeyFrame.cc:598-602
                        if (oct < A.level - 1 || oct > A.level) ok = false;                           // :901-902
                        if (ok && chi2_gate) {
                            const float ex = __fsub_rn(x, kx), ey = __fsub_rn(y, ky);
                            const float kr = u_right[i2];
                            float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
                            double lim = 5.99;
                            if (kr >= 0.f) { const float er = __fsub_rn(A.ur, kr); e2 = __fadd_rn(e2, __fmul_rn(er, er)); lim = 7.8; }
                            if ((double)__fmul_rn(e2, inv_sigma2[oct]) > lim) ok = false;            // :914, :926
                        }
                        if (ok) {
                            const unsigned d = (unsigned)hamming_regs(a0, a1, reinterpret_cast<const uint4*>(G.desc) + 2 * (size_t)i2);
                            best = min(best, ((unsigned long long)d << 40) | ((unsigned long long)(unsigned)(visited + p - pb) << 16) | (unsigned)i2);
                        }
                    }
                    visited += 32;
                }
            }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) best_idx[q] = (best != ~0ull && (unsigned)(best >> 40) <= max_dist) ? (int)(best & 0xFFFFu) : -1;   // TH_LOW (:943) / TH_HIGH
}

// SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, ORBdist): the projection of the keyframe's map points
// (src/ORBmatcher.cc:1496-1528).  No depth-sign test here, and "1.0/z" is a double division (:1503).
struct RelocCam { float R[9], t[3], Ow[3], fx, fy, cx, cy, minX, maxX, minY, maxY, log_scale, th; int nlevels; };

__global__ void k_reloc_queries(RelocCam C, const uint8_t* __restrict__ valid, const float* __restrict__ world_pos,
                                const float* __restrict__ max_distance, const float* __restrict__ min_distance,
                                const float* __restrict__ scale, int n, AreaQuery* __restrict__ Q)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    AreaQuery q;
    q.x = q.y = q.r = q.ur = 0.f; q.min_level = q.max_level = -1; q.active = 0; q.check_ur = 0;
    if (valid[i]) {
        const float P[3] = {world_pos[3 * (size_t)i], world_pos[3 * (size_t)i + 1], world_pos[3 * (size_t)i + 2]};
        float c[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            const float t = __fadd_rn(__fadd_rn(__fmul_rn(C.R[3 * r], P[0]), __fmul_rn(C.R[3 * r + 1], P[1])), __fmul_rn(C.R[3 * r + 2], P[2]));
            c[r] = (float)__dadd_rn(__dmul_rn((double)t, 1.0), __dmul_rn((double)C.t[r], 1.0));
        }
        const float invzc = (float)__ddiv_rn(1.0, (double)c[2]);
        const float u = __fadd_rn(__fmul_rn(__fmul_rn(C.fx, c[0]), invzc), C.cx);
        const float v = __fadd_rn(__fmul_rn(__fmul_rn(C.fy, c[1]), invzc), C.cy);
        if (!(u < C.minX || u > C.maxX) && !(v < C.minY || v > C.maxY)) {
            const float PO[3] = {__fsub_rn(P[0], C.Ow[0]), __fsub_rn(P[1], C.Ow[1]), __fsub_rn(P[2], C.Ow[2])};
            double s2 = __dmul_rn((double)PO[0], (double)PO[0]);
            s2 = __dadd_rn(s2, __dmul_rn((double)PO[1], (double)PO[1]));
            s2 = __dadd_rn(s2, __dmul_rn((double)PO[2], (double)PO[2]));
            const float dist3D = (float)__dsqrt_rn(s2);
            const float maxDistance = __fmul_rn(1.2f, max_distance[i]);
            const float minDistance = __fmul_rn(0.8f, min_distance[i]);
            if (!(dist3D < minDistance || dist3D > maxDistance)) {
                const float ratio = __fdiv_rn(max_distance[i], dist3D);
                const float cl = ceilf(__fdiv_rn(orb_logf(ratio), C.log_scale));
                int nScale = (fabsf(cl) < 2147483648.f) ? (int)cl : INT_MIN;
                if (nScale < 0) nScale = 0;
                else if (nScale >= C.nlevels) nScale = C.nlevels - 1;
                q.x = u; q.y = v; q.r = __fmul_rn(C.th, scale[nScale]);
                q.min_level = nScale - 1; q.max_level = nScale + 1;
                q.active = 1;
            }
        }
    }
    Q[i] = q;
}

// ---- SearchForTriangulation (src/ORBmatcher.cc:657-820) -----------------------------------------------------------
// Queries as for SearchByBoW: the features of pKF1 without a map point inside a vocabulary node both keyframes share.
// vbMatched2 is never set in the reference (:680, :722), so the queries are independent: one warp per query scans the
// node's features of pKF2 and keeps, among those that pass the static tests (no map point, stereo flag, distance <=
// TH_LOW, distance to the epipole, CheckDistEpipolarLine), the smallest distance -- the LAST one on ties, because the
// reference's running test is "dist > bestDist -> skip" (:734).
struct TriCam { float F[9]; float ex, ey; int only_stereo; };

// one entry per keyframe pair; the table travels with the inputs, blockIdx.y picks the pair: the neighbours of
// CreateNewMapPoints are one launch
struct TriJobDev {
    const BowQuery* Q; int nq; TriCam C;
    const float* keys1; const uint8_t* desc1; const float* ur1;
    const float* keys2; const uint8_t* desc2; const float* ur2; const uint8_t* has_mp2; const uint32_t* indices2;
    const float* scale2; const float* sigma2; int* result;
};

__global__ void __launch_bounds__(256) k_tri_match(const TriJobDev* __restrict__ jobs)
{
    const TriJobDev& J = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int q = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int nq = J.nq;
    if (q >= nq) return;
    const BowQuery* __restrict__ Q = J.Q;
    const TriCam C = J.C;
    const float* __restrict__ keys1 = J.keys1; const uint8_t* __restrict__ desc1 = J.desc1; const float* __restrict__ ur1 = J.ur1;
    const float* __restrict__ keys2 = J.keys2; const uint8_t* __restrict__ desc2 = J.desc2; const float* __restrict__ ur2 = J.ur2;
    const uint8_t* __restrict__ has_mp2 = J.has_mp2; const uint32_t* __restrict__ indices2 = J.indices2;
    const float* __restrict__ scale2 = J.scale2; const float* __restrict__ sigma2 = J.sigma2; int* __restrict__ result = J.result;
    const BowQuery B = Q[q];
    const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)B.idx1);
    const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(desc1) + 2 * (size_t)B.idx1 + 1);
    const float x1 = keys1[7 * (size_t)B.idx1], y1 = keys1[7 * (size_t)B.idx1 + 1];
    const bool stereo1 = ur1[B.idx1] >= 0.f;
    // epipolar line l = x1' F12 (:143-145), float products summed left to right
    const float la = __fadd_rn(__fadd_rn(__fmul_rn(x1, C.F[0]), __fmul_rn(y1, C.F[3])), C.F[6]);
    const float lb = __fadd_rn(__fadd_rn(__fmul_rn(x1, C.F[1]), __fmul_rn(y1, C.F[4])), C.F[7]);
    const float lc = __fadd_rn(__fadd_rn(__fmul_rn(x1, C.F[2]), __fmul_rn(y1, C.F[5])), C.F[8]);
    const float den = __fadd_rn(__fmul_rn(la, la), __fmul_rn(lb, lb));
    unsigned best = 0xFFFFFFFFu;
    for (int p = lane; p < B.len; p += 32) {
        const uint32_t i2 = indices2[B.start + p];
        if (has_mp2[i2]) continue;                             // :722
        const bool stereo2 = ur2[i2] >= 0.f;
        if (C.only_stereo && !stereo2) continue;
        const int d = hamming_regs(a0, a1, reinterpret_cast<const uint4*>(desc2) + 2 * (size_t)i2);
        if (d > 50) continue;                                  // TH_LOW (:734)
        const float x2 = keys2[7 * (size_t)i2], y2 = keys2[7 * (size_t)i2 + 1];
        const int oct = __float_as_int(keys2[7 * (size_t)i2 + 5]);
        if (!stereo1 && !stereo2) {                            // :739-745: too close to the epipole
            const float dx = __fsub_rn(C.ex, x2), dy = __fsub_rn(C.ey, y2);
            if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.f, scale2[oct])) continue;
        }
        if (den == 0.f) continue;                              // CheckDistEpipolarLine (:147-156)
        const float num = __fadd_rn(__fadd_rn(__fmul_rn(la, x2), __fmul_rn(lb, y2)), lc);
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if (!((double)dsqr < __dmul_rn(3.84, (double)sigma2[oct]))) continue;
        best = min(best, ((unsigned)d << 16) | (unsigned)(0xFFFF - p));
    }
    best = __reduce_min_sync(0xffffffffu, best);
    if (lane == 0) result[q] = best != 0xFFFFFFFFu ? (int)indices2[B.start + (0xFFFF - (int)(best & 0xFFFFu))] : -1;
}

// ---- host side ---------------------------------------------------------------------------------------------------
namespace {

#define CUS(call)                                                                                         \
    do {                                                                                                  \
        cudaError_t _e = (call);                                                                          \
        if (_e != cudaSuccess) {                                                                          \
            char _b[256];                                                                                 \
            snprintf(_b, sizeof(_b), "%s at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(_e)); \
            orbgpu_internal_set_error(_b);                                                                \
            return ORBGPU_ERR_CUDA;                                                                       \
        }                                                                                                 \
    } while (0)

// Per host thread: one device block, one pinned block and one stream, grown on demand and kept for the next call
// (a call is then: pack inputs into the pinned block, ONE H2D copy, the kernels, ONE D2H copy, one synchronise).
struct Workspace {
    int device = -1;
    uint8_t* d = nullptr; size_t dcap = 0;
    uint8_t* h = nullptr; size_t hcap = 0;
    cudaStream_t st = nullptr;
    bool attr_set = false;
    int reserve(int dev, size_t dbytes, size_t hbytes)
    {
        CUS(cudaSetDevice(dev));
        if (dev != device) {
            if (d) cudaFree(d);
            if (st) cudaStreamDestroy(st);
            d = nullptr; dcap = 0; st = nullptr; attr_set = false; device = dev;
        }
        if (!st) CUS(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        if (!attr_set) {
            CUS(cudaFuncSetAttribute(k_grid_build, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_BUILD_SMEM));
            CUS(cudaFuncSetAttribute(k_proj_queries, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_BUILD_SMEM));
            CUS(cudaFuncSetAttribute(k_frustum, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_BUILD_SMEM));
            CUS(cudaFuncSetAttribute(k_init_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, RESOLVE_SMEM));
            CUS(cudaFuncSetAttribute(k_proj_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, RESOLVE_SMEM));
            CUS(cudaFuncSetAttribute(k_bow_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, RESOLVE_SMEM));
            CUS(cudaFuncSetAttribute(k_local_resolve, cudaFuncAttributeMaxDynamicSharedMemorySize, RESOLVE_SMEM));
            attr_set = true;
        }
        if (dbytes > dcap) {
            if (d) cudaFree(d);
            d = nullptr; dcap = 0;
            const size_t want = dbytes + dbytes / 4;
            if (cudaMalloc(&d, want) != cudaSuccess) { cudaGetLastError(); return ORBGPU_ERR_NOMEM; }
            dcap = want;
        }
        if (hbytes > hcap) {
            if (h) cudaFreeHost(h);
            h = nullptr; hcap = 0;
            const size_t want = hbytes + hbytes / 4;
            if (cudaMallocHost(&h, want) != cudaSuccess) { cudaGetLastError(); return ORBGPU_ERR_NOMEM; }
            hcap = want;
        }
        return ORBGPU_OK;
    }
};
thread_local Workspace g_ws;      // released with the process (no CUDA calls from thread-exit destructors)

struct Layout {                   // 256-byte aligned carve-out shared by the device block and its pinned mirror
    size_t off = 0;
    size_t add(size_t bytes) { const size_t o = off; off = (off + bytes + 255) & ~(size_t)255; return o; }
};

// src/ORBmatcher.cc:1601-1642 on bin sizes
void three_maxima(const int* size, int L, int& ind1, int& ind2, int& ind3)
{
    int top[3] = {0, 0, 0};
    int* ind[3] = {&ind1, &ind2, &ind3};
    for (int i = 0; i < L; ++i) {
        const int s = size[i];
        int rank = 0;
        while (rank < 3 && s <= top[rank]) ++rank;
        if (rank == 3) continue;
        for (int k = 2; k > rank; --k) { top[k] = top[k - 1]; *ind[k] = *ind[k - 1]; }
        top[rank] = s; *ind[rank] = i;
    }
    if ((float)top[1] < 0.1f * (float)top[0]) { ind2 = -1; ind3 = -1; }
    else if ((float)top[2] < 0.1f * (float)top[0]) { ind3 = -1; }
}

inline int rot_bin(float a1, float a2)                        // src/ORBmatcher.cc:474-480
{
    const float factor = 1.0f / 30;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)roundf(rot * factor);
    if (bin == 30) bin = 0;
    return bin;
}

void set_grid(GridDev& G, uint8_t* d, size_t o_keys, size_t o_desc, size_t o_start, size_t o_sorted, int n, const float* bounds)
{
    G.keys = reinterpret_cast<const float*>(d + o_keys); G.desc = d + o_desc; G.n = n;
    G.minX = bounds[0]; G.minY = bounds[2];
    G.winX = bounds[0]; G.winY = bounds[2];
    G.invW = static_cast<float>(GRID_COLS) / static_cast<float>(bounds[1] - bounds[0]);       // src/Frame.cc:155-156
    G.invH = static_cast<float>(GRID_ROWS) / static_cast<float>(bounds[3] - bounds[2]);
    G.cell_start = reinterpret_cast<int*>(d + o_start); G.sorted = reinterpret_cast<int*>(d + o_sorted);
}

}  // namespace

// cudaMemcpyAsync of a few KB takes the driver's inline path and costs the calling thread 5-8 us more than a 32 KB copy
// (measured on B200: enqueue 15.2 -> 7.5 us for the 14 KB of a resident SearchByBoW); small uploads are padded up to this
// size with whatever follows them in the pinned block (the device block is at least as large)
static const size_t SEARCH_MIN_UPLOAD = (size_t)32 << 10;

// ORBGPU_TRACE_SEARCH=1: host-clock phases of the BoW-gated searches on stderr (where a call's microseconds go)
struct PhaseTrace {
    bool on; const char* what; int k = 0; double t[8]; const char* name[8];
    explicit PhaseTrace(const char* w) : what(w) { static const bool env = getenv("ORBGPU_TRACE_SEARCH") != nullptr; on = env; mark("entry"); }
    void mark(const char* n)
    {
        if (!on || k >= 8) return;
        timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
        t[k] = ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; name[k++] = n;
    }
    ~PhaseTrace()
    {
        if (!on) return;
        fprintf(stderr, "[orbgpu] %s:", what);
        for (int i = 1; i < k; ++i) fprintf(stderr, " %s %.1f", name[i], t[i] - t[i - 1]);
        fprintf(stderr, " | total %.1f us\n", t[k - 1] - t[0]);
    }
};

// ORBGPU_TRACE_SEARCH=1, device side: time between marks on one stream (CUDA events), printed by the destructor
struct StreamTrace {
    bool on; cudaStream_t st; int k = 0; cudaEvent_t ev[10]; const char* name[10];
    StreamTrace(bool enabled, cudaStream_t s) : on(enabled), st(s) { mark("begin"); }
    void mark(const char* n)
    {
        if (!on || k >= 10) return;
        cudaEventCreate(&ev[k]); cudaEventRecord(ev[k], st); name[k++] = n;
    }
    ~StreamTrace()
    {
        if (!on) return;
        cudaEventSynchronize(ev[k - 1]);
        fprintf(stderr, "[orbgpu]   device:");
        for (int i = 1; i < k; ++i) { float ms = 0.f; cudaEventElapsedTime(&ms, ev[i - 1], ev[i]); fprintf(stderr, " %s %.1f", name[i], ms * 1e3); }
        fprintf(stderr, " us\n");
        for (int i = 0; i < k; ++i) cudaEventDestroy(ev[i]);
    }
};

extern "C" {

int orbgpu_search_for_initialization(int device, const orbgpu_keypoint* keys1_un, const uint8_t* desc1, int n1,
                                     const orbgpu_keypoint* keys2_un, const uint8_t* desc2, int n2, const float* bounds4,
                                     float* prev_matched, int window_size, float nnratio, int check_orientation,
                                     int32_t* matches12, int* nmatches)
{
    if (!keys1_un || !desc1 || !keys2_un || !desc2 || !bounds4 || !prev_matched || !matches12 || !nmatches || n1 < 0 || n2 < 0)
        return ORBGPU_ERR_ARG;
    if (!(bounds4[1] > bounds4[0]) || !(bounds4[3] > bounds4[2])) return ORBGPU_ERR_ARG;
    if (n1 > GRID_MAX_KEYS || n2 > GRID_MAX_KEYS || (size_t)n1 * (size_t)n2 > ((size_t)1 << 26)) return ORBGPU_ERR_UNSUPPORTED;
    *nmatches = 0;
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBGPU_OK;
    // device block: [inputs | outputs | scratch]; the pinned block mirrors inputs and outputs
    Layout L;
    const size_t o_keys = L.add(sizeof(orbgpu_keypoint) * (size_t)n2), o_desc = L.add(32 * (size_t)n2);
    const size_t o_q = L.add(sizeof(AreaQuery) * (size_t)n1), o_qdesc = L.add(32 * (size_t)n1);
    const size_t in_bytes = L.off;
    const size_t o_m12 = L.add(4 * (size_t)n1), o_acc = L.add(4 * (size_t)n1), o_nm = L.add(4);
    const size_t io_bytes = L.off;
    const size_t o_start = L.add(4 * (GRID_CELLS + 1)), o_sorted = L.add(4 * (size_t)n2), o_cn = L.add(4 * (size_t)n1);
    const size_t o_cand = L.add(4 * (size_t)n1 * n2);
    Workspace& W = g_ws;
    int rc = W.reserve(device, L.off, io_bytes);
    if (rc != ORBGPU_OK) return rc;
    memcpy(W.h + o_keys, keys2_un, sizeof(orbgpu_keypoint) * (size_t)n2);
    memcpy(W.h + o_desc, desc2, 32 * (size_t)n2);
    memcpy(W.h + o_qdesc, desc1, 32 * (size_t)n1);
    AreaQuery* hq = reinterpret_cast<AreaQuery*>(W.h + o_q);   // level-0 keypoints of F1 around vbPrevMatched (:418-425)
    for (int i = 0; i < n1; ++i) {
        AreaQuery& q = hq[i];
        q.x = prev_matched[2 * i]; q.y = prev_matched[2 * i + 1]; q.r = (float)window_size;
        q.min_level = q.max_level = keys1_un[i].octave;
        q.active = keys1_un[i].octave > 0 ? 0 : 1;
        q.ur = 0.f; q.check_ur = 0;
    }
    CUS(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.st));
    GridDev G;
    set_grid(G, W.d, o_keys, o_desc, o_start, o_sorted, n2, bounds4);
    uint32_t* d_cand = reinterpret_cast<uint32_t*>(W.d + o_cand);
    int* d_cn = reinterpret_cast<int*>(W.d + o_cn);
    k_grid_build<<<1, 1024, 4 * (size_t)G.n, W.st>>>(G);
    k_area_candidates<<<(n1 + 7) / 8, 256, 0, W.st>>>(G, reinterpret_cast<const AreaQuery*>(W.d + o_q), W.d + o_qdesc, n1, nullptr, d_cand,
                                                      n2, d_cn);
    k_init_resolve<<<1, RESOLVE_THREADS, RESOLVE_SMEM, W.st>>>(d_cand, d_cn, n2, n1, n2, nnratio, reinterpret_cast<int*>(W.d + o_m12),
                                                               reinterpret_cast<int*>(W.d + o_acc), reinterpret_cast<int*>(W.d + o_nm));
    CUS(cudaGetLastError());
    CUS(cudaMemcpyAsync(W.h + in_bytes, W.d + in_bytes, io_bytes - in_bytes, cudaMemcpyDeviceToHost, W.st));
    CUS(cudaStreamSynchronize(W.st));
    memcpy(matches12, W.h + o_m12, 4 * (size_t)n1);
    const int* acc = reinterpret_cast<const int*>(W.h + o_acc);
    int nm = *reinterpret_cast<const int*>(W.h + o_nm);
    if (check_orientation) {                                   // :487-511, bins in push order = ascending i1
        int bin_of[30] = {0};
        std::vector<int> bin((size_t)n1, -1);
        for (int i1 = 0; i1 < n1; ++i1)
            if (acc[i1] >= 0) { bin[i1] = rot_bin(keys1_un[i1].angle, keys2_un[acc[i1]].angle); bin_of[bin[i1]]++; }
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(bin_of, 30, ind1, ind2, ind3);
        for (int i1 = 0; i1 < n1; ++i1) {
            if (bin[i1] < 0 || bin[i1] == ind1 || bin[i1] == ind2 || bin[i1] == ind3) continue;
            if (matches12[i1] >= 0) { matches12[i1] = -1; nm--; }
        }
    }
    for (int i1 = 0; i1 < n1; ++i1)                            // :514-517
        if (matches12[i1] >= 0) { prev_matched[2 * i1] = keys2_un[matches12[i1]].x; prev_matched[2 * i1 + 1] = keys2_un[matches12[i1]].y; }
    *nmatches = nm;
    return ORBGPU_OK;
}

int orbgpu_search_by_projection(int device, const orbgpu_projection_args* a, int32_t* cur_match, int* nmatches)
{
    if (!a || !cur_match || !nmatches || a->n_last < 0 || a->n_cur < 0) return ORBGPU_ERR_ARG;
    if (!a->world_pos || !a->mp_desc || !a->mp_valid || !a->mp_observed || !a->last_octave || !a->last_angle || !a->Tcw_last ||
        !a->cur_keys_un || !a->cur_desc || !a->cur_u_right || !a->cur_occupied || !a->Tcw_cur || !a->scale_factors || a->nlevels < 1)
        return ORBGPU_ERR_ARG;
    if (!(a->bounds[1] > a->bounds[0]) || !(a->bounds[3] > a->bounds[2])) return ORBGPU_ERR_ARG;
    const int n1 = a->n_last, n2 = a->n_cur;
    if (n1 > GRID_MAX_KEYS || n2 > GRID_MAX_KEYS || (size_t)n1 * (size_t)n2 > ((size_t)1 << 26)) return ORBGPU_ERR_UNSUPPORTED;
    PhaseTrace tr("search_by_projection");
    for (int i = 0; i < n1; ++i)
        if (a->mp_valid[i] && (a->last_octave[i] < 0 || a->last_octave[i] >= a->nlevels)) return ORBGPU_ERR_ARG;
    *nmatches = 0;
    for (int i = 0; i < n2; ++i) cur_match[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBGPU_OK;
    // camera motion between the frames (:1339-1350); the cv::Mat expressions restated with cv::gemm's arithmetic
    ProjCam C;
    const float* Tc = a->Tcw_cur; const float* Tl = a->Tcw_last;
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) C.R[3 * r + c] = Tc[4 * r + c]; C.t[r] = Tc[4 * r + 3]; }
    float twc[3], tlc[3];
    for (int i = 0; i < 3; ++i) {                              // -Rcw.t() * tcw: cv::MatExpr negates a transposed COPY,
        volatile float t = Tc[i] * C.t[0];                     // so this is gemm's 3x3 float case with alpha = -1
        t = t + Tc[4 + i] * C.t[1];                            // (volatile: keep the host compiler from contracting)
        t = t + Tc[8 + i] * C.t[2];
        twc[i] = (float)((double)t * -1.0);
    }
    for (int i = 0; i < 3; ++i) {                              // Rlw * twc + tlw: 3x3 float case
        const float t = Tl[4 * i] * twc[0] + Tl[4 * i + 1] * twc[1] + Tl[4 * i + 2] * twc[2];
        tlc[i] = (float)((double)t * 1.0 + (double)Tl[4 * i + 3] * 1.0);
    }
    C.forward = (tlc[2] > a->b && !a->mono) ? 1 : 0;
    C.backward = (-tlc[2] > a->b && !a->mono) ? 1 : 0;
    C.fx = a->fx; C.fy = a->fy; C.cx = a->cx; C.cy = a->cy; C.bf = a->bf; C.th = a->th;
    C.minX = a->bounds[0]; C.maxX = a->bounds[1]; C.minY = a->bounds[2]; C.maxY = a->bounds[3];

    Layout L;
    const size_t o_keys = L.add(sizeof(orbgpu_keypoint) * (size_t)n2), o_desc = L.add(32 * (size_t)n2);
    const size_t o_pos = L.add(12 * (size_t)n1), o_qdesc = L.add(32 * (size_t)n1), o_valid = L.add((size_t)n1), o_obs = L.add((size_t)n1);
    const size_t o_occ = L.add((size_t)n2), o_oct = L.add(4 * (size_t)n1), o_scale = L.add(4 * (size_t)a->nlevels), o_ur = L.add(4 * (size_t)n2);
    const size_t in_bytes = L.off;
    const size_t o_cm = L.add(4 * (size_t)n2), o_acc = L.add(4 * (size_t)n1), o_nm = L.add(4);
    const size_t io_bytes = L.off;
    const size_t o_start = L.add(4 * (GRID_CELLS + 1)), o_sorted = L.add(4 * (size_t)n2), o_cn = L.add(4 * (size_t)n1);
    const size_t o_q = L.add(sizeof(AreaQuery) * (size_t)n1), o_cand = L.add(4 * (size_t)n1 * n2);
    Workspace& W = g_ws;
    int rc = W.reserve(device, L.off, io_bytes);
    if (rc != ORBGPU_OK) return rc;
    memcpy(W.h + o_keys, a->cur_keys_un, sizeof(orbgpu_keypoint) * (size_t)n2);
    memcpy(W.h + o_desc, a->cur_desc, 32 * (size_t)n2);
    memcpy(W.h + o_pos, a->world_pos, 12 * (size_t)n1);
    memcpy(W.h + o_qdesc, a->mp_desc, 32 * (size_t)n1);
    memcpy(W.h + o_valid, a->mp_valid, (size_t)n1);
    memcpy(W.h + o_obs, a->mp_observed, (size_t)n1);
    memcpy(W.h + o_occ, a->cur_occupied, (size_t)n2);
    memcpy(W.h + o_oct, a->last_octave, 4 * (size_t)n1);
    memcpy(W.h + o_scale, a->scale_factors, 4 * (size_t)a->nlevels);
    memcpy(W.h + o_ur, a->cur_u_right, 4 * (size_t)n2);
    tr.mark("stage");
    StreamTrace dt(tr.on, W.st);
    CUS(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.st));
    dt.mark("upload");
    GridDev G;
    set_grid(G, W.d, o_keys, o_desc, o_start, o_sorted, n2, a->bounds);
    uint32_t* d_cand = reinterpret_cast<uint32_t*>(W.d + o_cand);
    int* d_cn = reinterpret_cast<int*>(W.d + o_cn);
    AreaQuery* d_q = reinterpret_cast<AreaQuery*>(W.d + o_q);
    // the queries and, in one block more, the grid of the current frame
    k_proj_queries<<<(n1 + 1023) / 1024 + 1, 1024, 4 * (size_t)G.n, W.st>>>(C, reinterpret_cast<const float*>(W.d + o_pos), W.d + o_valid,
                                                                             reinterpret_cast<const int*>(W.d + o_oct),
                                                                             reinterpret_cast<const float*>(W.d + o_scale), n1, d_q, G);
    dt.mark("grid+queries");
    k_area_candidates<<<(n1 + 7) / 8, 256, 0, W.st>>>(G, d_q, W.d + o_qdesc, n1, reinterpret_cast<const float*>(W.d + o_ur), d_cand, n2, d_cn);
    dt.mark("candidates");
    k_proj_resolve<<<1, RESOLVE_THREADS, RESOLVE_SMEM, W.st>>>(d_cand, d_cn, n2, n1, n2, W.d + o_obs, W.d + o_occ, 100, reinterpret_cast<int*>(W.d + o_cm),
                                                               reinterpret_cast<int*>(W.d + o_acc), reinterpret_cast<int*>(W.d + o_nm));
    dt.mark("resolve");
    CUS(cudaGetLastError());
    CUS(cudaMemcpyAsync(W.h + in_bytes, W.d + in_bytes, io_bytes - in_bytes, cudaMemcpyDeviceToHost, W.st));
    dt.mark("download");
    tr.mark("enqueue");
    CUS(cudaStreamSynchronize(W.st));
    tr.mark("wait");
    memcpy(cur_match, W.h + o_cm, 4 * (size_t)n2);
    const int* acc = reinterpret_cast<const int*>(W.h + o_acc);
    int nm = *reinterpret_cast<const int*>(W.h + o_nm);
    if (a->check_orientation) {                                // :1446-1466: every entry of a losing bin is reset, even twice
        int bin_of[30] = {0};
        std::vector<int> bin((size_t)n1, -1);
        for (int i = 0; i < n1; ++i)
            if (acc[i] >= 0) { bin[i] = rot_bin(a->last_angle[i], a->cur_keys_un[acc[i]].angle); bin_of[bin[i]]++; }
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(bin_of, 30, ind1, ind2, ind3);
        for (int i = 0; i < n1; ++i) {
            if (bin[i] < 0 || bin[i] == ind1 || bin[i] == ind2 || bin[i] == ind3) continue;
            cur_match[acc[i]] = -2;
            nm--;
        }
    }
    *nmatches = nm;
    return ORBGPU_OK;
}

int orbgpu_search_local_points(int device, const orbgpu_local_points_args* a, int32_t* cur_match, int* nmatches)
{
    if (!a || !cur_match || !nmatches || a->n_points < 0 || a->n_cur < 0 || a->nlevels < 1 || a->nlevels > 255) return ORBGPU_ERR_ARG;
    if (!a->in_view || !a->proj_x || !a->proj_y || !a->proj_xr || !a->scale_level || !a->view_cos || !a->mp_desc || !a->mp_observed ||
        !a->cur_keys_un || !a->cur_desc || !a->cur_u_right || !a->cur_occupied || !a->scale_factors) return ORBGPU_ERR_ARG;
    if (!(a->bounds[1] > a->bounds[0]) || !(a->bounds[3] > a->bounds[2])) return ORBGPU_ERR_ARG;
    const int n1 = a->n_points, n2 = a->n_cur;
    if (n1 > GRID_MAX_KEYS || n2 > GRID_MAX_KEYS || (size_t)n1 * (size_t)n2 > ((size_t)1 << 26)) return ORBGPU_ERR_UNSUPPORTED;
    for (int i = 0; i < n1; ++i)
        if (a->in_view[i] && (a->scale_level[i] < 0 || a->scale_level[i] >= a->nlevels)) return ORBGPU_ERR_ARG;
    for (int i = 0; i < n2; ++i)
        if (a->cur_keys_un[i].octave < 0 || a->cur_keys_un[i].octave > 255) return ORBGPU_ERR_ARG;
    *nmatches = 0;
    for (int i = 0; i < n2; ++i) cur_match[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBGPU_OK;
    Layout L;
    const size_t o_keys = L.add(sizeof(orbgpu_keypoint) * (size_t)n2), o_desc = L.add(32 * (size_t)n2);
    const size_t o_q = L.add(sizeof(AreaQuery) * (size_t)n1), o_qdesc = L.add(32 * (size_t)n1), o_obs = L.add((size_t)n1);
    const size_t o_occ = L.add((size_t)n2), o_ur = L.add(4 * (size_t)n2);
    const size_t in_bytes = L.off;
    const size_t o_cm = L.add(4 * (size_t)n2), o_nm = L.add(4);
    const size_t io_bytes = L.off;
    const size_t o_start = L.add(4 * (GRID_CELLS + 1)), o_sorted = L.add(4 * (size_t)n2), o_cn = L.add(4 * (size_t)n1);
    const size_t o_cand = L.add(4 * (size_t)n1 * n2);
    Workspace& W = g_ws;
    int rc = W.reserve(device, L.off, io_bytes);
    if (rc != ORBGPU_OK) return rc;
    memcpy(W.h + o_keys, a->cur_keys_un, sizeof(orbgpu_keypoint) * (size_t)n2);
    memcpy(W.h + o_desc, a->cur_desc, 32 * (size_t)n2);
    memcpy(W.h + o_qdesc, a->mp_desc, 32 * (size_t)n1);
    memcpy(W.h + o_obs, a->mp_observed, (size_t)n1);
    memcpy(W.h + o_occ, a->cur_occupied, (size_t)n2);
    memcpy(W.h + o_ur, a->cur_u_right, 4 * (size_t)n2);
    AreaQuery* hq = reinterpret_cast<AreaQuery*>(W.h + o_q);
    const bool bFactor = a->th != 1.0;                         // :48
    for (int i = 0; i < n1; ++i) {
        AreaQuery& q = hq[i];
        q.x = q.y = q.r = q.ur = 0.f; q.min_level = q.max_level = -1; q.active = 0; q.check_ur = 1;
        if (!a->in_view[i]) continue;                          // :53-57
        const int lvl = a->scale_level[i];
        volatile float r = (double)a->view_cos[i] > 0.998 ? 2.5f : 4.0f;        // RadiusByViewingCos (:127-133)
        if (bFactor) r = r * a->th;
        r = r * a->scale_factors[lvl];                         // the window of :67 and the stereo gate of :96
        q.x = a->proj_x[i]; q.y = a->proj_y[i]; q.r = r;
        q.min_level = lvl - 1; q.max_level = lvl;
        q.ur = a->proj_xr[i];
        q.active = 1;
    }
    CUS(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.st));
    GridDev G;
    set_grid(G, W.d, o_keys, o_desc, o_start, o_sorted, n2, a->bounds);
    uint32_t* d_cand = reinterpret_cast<uint32_t*>(W.d + o_cand);
    int* d_cn = reinterpret_cast<int*>(W.d + o_cn);
    k_grid_build<<<1, 1024, 4 * (size_t)G.n, W.st>>>(G);
    k_area_candidates<<<(n1 + 7) / 8, 256, 0, W.st>>>(G, reinterpret_cast<const AreaQuery*>(W.d + o_q), W.d + o_qdesc, n1,
                                                      reinterpret_cast<const float*>(W.d + o_ur), d_cand, n2, d_cn);
    k_local_resolve<<<1, RESOLVE_THREADS, RESOLVE_SMEM, W.st>>>(d_cand, d_cn, n2, n1, n2, W.d + o_obs, W.d + o_occ, G.keys, a->nnratio,
                                                                reinterpret_cast<int*>(W.d + o_cm), reinterpret_cast<int*>(W.d + o_nm));
    CUS(cudaGetLastError());
    CUS(cudaMemcpyAsync(W.h + in_bytes, W.d + in_bytes, io_bytes - in_bytes, cudaMemcpyDeviceToHost, W.st));
    CUS(cudaStreamSynchronize(W.st));
    memcpy(cur_match, W.h + o_cm, 4 * (size_t)n2);
    *nmatches = *reinterpret_cast<const int*>(W.h + o_nm);
    return ORBGPU_OK;
}

int orbgpu_search_local_map(int device, const orbgpu_local_map_args* a, const orbgpu_track_fields* out, int* n_in_view,
                            int32_t* cur_match, int* nmatches)
{
    if (!a || !out || !n_in_view || !cur_match || !nmatches || a->n_points < 0 || a->n_cur < 0 || a->nlevels < 1 || a->nlevels > 255)
        return ORBGPU_ERR_ARG;
    if (!a->candidate || !a->world_pos || !a->normal || !a->max_distance || !a->min_distance || !a->mp_desc || !a->mp_observed || !a->Tcw ||
        !a->scale_factors || !out->in_view || !out->proj_x || !out->proj_y || !out->proj_xr || !out->scale_level || !out->view_cos)
        return ORBGPU_ERR_ARG;
    if (a->n_cur > 0 && (!a->cur_keys_un || !a->cur_desc || !a->cur_u_right || !a->cur_occupied)) return ORBGPU_ERR_ARG;
    if (!(a->bounds[1] > a->bounds[0]) || !(a->bounds[3] > a->bounds[2])) return ORBGPU_ERR_ARG;
    const int n1 = a->n_points, n2 = a->n_cur;
    if (n1 > GRID_MAX_KEYS || n2 > GRID_MAX_KEYS || (size_t)n1 * (size_t)n2 > ((size_t)1 << 26)) return ORBGPU_ERR_UNSUPPORTED;
    for (int i = 0; i < n2; ++i)
        if (a->cur_keys_un[i].octave < 0 || a->cur_keys_un[i].octave > 255) return ORBGPU_ERR_ARG;
    *nmatches = 0; *n_in_view = 0;
    for (int i = 0; i < n2; ++i) cur_match[i] = -1;
    if (n1 == 0) return ORBGPU_OK;
    // pose matrices as Frame::UpdatePoseMatrices leaves them (src/Frame.cc:261-267); mOw = -mRcw.t()*mtcw is the same
    // cv::MatExpr as at src/ORBmatcher.cc:1341: gemm of a transposed copy with alpha = -1 (float products, float sums)
    FrustumCam C;
    const float* Tc = a->Tcw;
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) C.R[3 * r + c] = Tc[4 * r + c]; C.t[r] = Tc[4 * r + 3]; }
    for (int i = 0; i < 3; ++i) {
        volatile float t = Tc[i] * C.t[0];
        t = t + Tc[4 + i] * C.t[1];
        t = t + Tc[8 + i] * C.t[2];
        C.Ow[i] = (float)((double)t * -1.0);
    }
    C.fx = a->fx; C.fy = a->fy; C.cx = a->cx; C.cy = a->cy; C.bf = a->bf;
    C.minX = a->bounds[0]; C.maxX = a->bounds[1]; C.minY = a->bounds[2]; C.maxY = a->bounds[3];
    C.cos_limit = a->viewing_cos_limit; C.log_scale = a->log_scale_factor; C.th = a->th; C.nlevels = a->nlevels;

    Layout L;
    const size_t o_keys = L.add(sizeof(orbgpu_keypoint) * (size_t)n2), o_desc = L.add(32 * (size_t)n2);
    const size_t o_qdesc = L.add(32 * (size_t)n1), o_obs = L.add((size_t)n1), o_occ = L.add((size_t)n2), o_ur = L.add(4 * (size_t)n2);
    const size_t o_cand = L.add((size_t)n1), o_pos = L.add(12 * (size_t)n1), o_nrm = L.add(12 * (size_t)n1);
    const size_t o_maxd = L.add(4 * (size_t)n1), o_mind = L.add(4 * (size_t)n1), o_scale = L.add(4 * (size_t)a->nlevels);
    const size_t in_bytes = L.off;
    const size_t o_cm = L.add(4 * (size_t)n2), o_nm = L.add(8);
    const size_t o_iv = L.add((size_t)n1), o_px = L.add(4 * (size_t)n1), o_py = L.add(4 * (size_t)n1), o_pxr = L.add(4 * (size_t)n1);
    const size_t o_vc = L.add(4 * (size_t)n1), o_lvl = L.add(4 * (size_t)n1);
    const size_t io_bytes = L.off;
    const size_t o_start = L.add(4 * (GRID_CELLS + 1)), o_sorted = L.add(4 * (size_t)n2), o_cn = L.add(4 * (size_t)n1);
    const size_t o_q = L.add(sizeof(AreaQuery) * (size_t)n1), o_cl = L.add(4 * (size_t)n1 * (size_t)(n2 > 0 ? n2 : 1));
    PhaseTrace tr("search_local_map");
    Workspace& W = g_ws;
    int rc = W.reserve(device, L.off, io_bytes);
    if (rc != ORBGPU_OK) return rc;
    if (n2) {
        memcpy(W.h + o_keys, a->cur_keys_un, sizeof(orbgpu_keypoint) * (size_t)n2);
        memcpy(W.h + o_desc, a->cur_desc, 32 * (size_t)n2);
        memcpy(W.h + o_occ, a->cur_occupied, (size_t)n2);
        memcpy(W.h + o_ur, a->cur_u_right, 4 * (size_t)n2);
    }
    memcpy(W.h + o_qdesc, a->mp_desc, 32 * (size_t)n1);
    memcpy(W.h + o_obs, a->mp_observed, (size_t)n1);
    memcpy(W.h + o_cand, a->candidate, (size_t)n1);
    memcpy(W.h + o_pos, a->world_pos, 12 * (size_t)n1);
    memcpy(W.h + o_nrm, a->normal, 12 * (size_t)n1);
    memcpy(W.h + o_maxd, a->max_distance, 4 * (size_t)n1);
    memcpy(W.h + o_mind, a->min_distance, 4 * (size_t)n1);
    memcpy(W.h + o_scale, a->scale_factors, 4 * (size_t)a->nlevels);
    tr.mark("stage");
    StreamTrace dt(tr.on, W.st);
    CUS(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.st));
    dt.mark("upload");
    CUS(cudaMemsetAsync(W.d + o_nm, 0, 8, W.st));
    TrackOut T;
    T.in_view = W.d + o_iv; T.proj_x = reinterpret_cast<float*>(W.d + o_px); T.proj_y = reinterpret_cast<float*>(W.d + o_py);
    T.proj_xr = reinterpret_cast<float*>(W.d + o_pxr); T.view_cos = reinterpret_cast<float*>(W.d + o_vc);
    T.scale_level = reinterpret_cast<int*>(W.d + o_lvl);
    AreaQuery* d_q = reinterpret_cast<AreaQuery*>(W.d + o_q);
    int* d_nm = reinterpret_cast<int*>(W.d + o_nm);
    GridDev G;
    memset(&G, 0, sizeof(G));
    if (n2) set_grid(G, W.d, o_keys, o_desc, o_start, o_sorted, n2, a->bounds);
    // isInFrustum for every point and, in one block more, the grid of the frame (when there is a frame to search)
    k_frustum<<<(n1 + 1023) / 1024 + (n2 ? 1 : 0), 1024, 4 * (size_t)n2, W.st>>>(
        C, W.d + o_cand, reinterpret_cast<const float*>(W.d + o_pos), reinterpret_cast<const float*>(W.d + o_nrm),
        reinterpret_cast<const float*>(W.d + o_maxd), reinterpret_cast<const float*>(W.d + o_mind),
        reinterpret_cast<const float*>(W.d + o_scale), n1, T, d_q, d_nm + 1, G, n2 ? 1 : 0);
    if (n2) {
        uint32_t* d_cl = reinterpret_cast<uint32_t*>(W.d + o_cl);
        int* d_cn = reinterpret_cast<int*>(W.d + o_cn);
        dt.mark("frustum+grid");
        k_area_candidates<<<(n1 + 7) / 8, 256, 0, W.st>>>(G, d_q, W.d + o_qdesc, n1, reinterpret_cast<const float*>(W.d + o_ur), d_cl, n2, d_cn);
        dt.mark("candidates");
        k_local_resolve<<<1, RESOLVE_THREADS, RESOLVE_SMEM, W.st>>>(d_cl, d_cn, n2, n1, n2, W.d + o_obs, W.d + o_occ, G.keys, a->nnratio,
                                                                    reinterpret_cast<int*>(W.d + o_cm), d_nm);
    }
    CUS(cudaGetLastError());
    dt.mark("resolve");
    CUS(cudaMemcpyAsync(W.h + in_bytes, W.d + in_bytes, io_bytes - in_bytes, cudaMemcpyDeviceToHost, W.st));
    dt.mark("download");
    tr.mark("enqueue");
    CUS(cudaStreamSynchronize(W.st));
    tr.mark("wait");
    if (n2) memcpy(cur_match, W.h + o_cm, 4 * (size_t)n2);
    *nmatches = reinterpret_cast<const int*>(W.h + o_nm)[0];
    *n_in_view = reinterpret_cast<const int*>(W.h + o_nm)[1];
    memcpy(out->in_view, W.h + o_iv, (size_t)n1);
    memcpy(out->proj_x, W.h + o_px, 4 * (size_t)n1);
    memcpy(out->proj_y, W.h + o_py, 4 * (size_t)n1);
    memcpy(out->proj_xr, W.h + o_pxr, 4 * (size_t)n1);
    memcpy(out->view_cos, W.h + o_vc, 4 * (size_t)n1);
    memcpy(out->scale_level, W.h + o_lvl, 4 * (size_t)n1);
    return ORBGPU_OK;
}

/* test hook: the device logf of MapPoint::PredictScale on host arrays */
int orbgpu_debug_logf(int device, const float* x, int n, float* out)
{
    if (!x || !out || n < 0) return ORBGPU_ERR_ARG;
    if (n == 0) return ORBGPU_OK;
    Workspace& W = g_ws;
    int rc = W.reserve(device, 8 * (size_t)n, 16);
    if (rc != ORBGPU_OK) return rc;
    float* d_x = reinterpret_cast<float*>(W.d); float* d_o = d_x + n;
    CUS(cudaMemcpyAsync(d_x, x, 4 * (size_t)n, cudaMemcpyHostToDevice, W.st));
    k_debug_logf<<<(n + 255) / 256, 256, 0, W.st>>>(d_x, n, d_o);
    CUS(cudaGetLastError());
    CUS(cudaMemcpyAsync(out, d_o, 4 * (size_t)n, cudaMemcpyDeviceToHost, W.st));
    CUS(cudaStreamSynchronize(W.st));
    return ORBGPU_OK;
}

// One engine for the keyframe-projection loops (Fuse, Fuse with Sim3, SearchByProjection with Sim3, the two passes of
// SearchBySim3).  C carries the pose(s) and the mode flags; occupied != NULL switches from "best keypoint per point"
// (best[n_points]) to the occupancy-ordered assignment of SearchByProjection(KF, Scw) (best[n_kf] = point index).
// One keyframe-projection problem.  Alone (job == nullptr) it reserves the workspace, runs and waits.  As part of a batch it is
// called twice: phase 1 validates, presets the outputs and sizes the job's region; phase 2 stages and enqueues it at job->base
// without waiting (the caller has reserved the workspace for all jobs and collects the results after one stream wait).
struct FuseJob {
    size_t base = 0, bytes = 0, o_best = 0;
    int nout = 0;
    bool active = false;
};

static int fuse_run(int device, const orbgpu_fuse_args* a, FuseCam C, int chi2_gate, int max_dist, const uint8_t* occupied,
                    int32_t* best, int* nmatches, FuseJob* job = nullptr, int phase = 0)
{
    if (!a || !best || a->n_points < 0 || a->n_kf < 0 || a->nlevels < 1) return ORBGPU_ERR_ARG;
    if (!a->present || !a->world_pos || !a->max_distance || !a->min_distance || !a->mp_desc || !a->scale_factors) return ORBGPU_ERR_ARG;
    if ((C.view_check && !a->normal) || (chi2_gate && !a->inv_level_sigma2)) return ORBGPU_ERR_ARG;
    if (a->n_kf > 0 && (!a->keys_un || !a->desc || (chi2_gate && !a->u_right))) return ORBGPU_ERR_ARG;
    if (!(a->bounds[1] > a->bounds[0]) || !(a->bounds[3] > a->bounds[2])) return ORBGPU_ERR_ARG;
    const int n1 = a->n_points, n2 = a->n_kf;
    if (n1 > (1 << 20) || n2 > GRID_MAX_KEYS) return ORBGPU_ERR_UNSUPPORTED;
    if (occupied && (n1 > GRID_MAX_KEYS || (size_t)n1 * (size_t)n2 > ((size_t)1 << 26))) return ORBGPU_ERR_UNSUPPORTED;
    for (int i = 0; i < n2; ++i)
        if (a->keys_un[i].octave < 0 || a->keys_un[i].octave >= a->nlevels) return ORBGPU_ERR_ARG;
    const int nout = occupied ? n2 : n1;
    if (phase != 2) {
        for (int i = 0; i < nout; ++i) best[i] = -1;
        if (nmatches) *nmatches = 0;
    }
    if (n1 == 0 || n2 == 0) return ORBGPU_OK;
    C.fx = a->fx; C.fy = a->fy; C.cx = a->cx; C.cy = a->cy; C.bf = a->bf;
    // the keyframe holds the image bounds as ints (include/KeyFrame.h:185-188, src/KeyFrame.cc:42-43)
    C.minX = (float)(int)a->bounds[0]; C.maxX = (float)(int)a->bounds[1]; C.minY = (float)(int)a->bounds[2]; C.maxY = (float)(int)a->bounds[3];
    C.log_scale = a->log_scale_factor; C.th = a->th; C.nlevels = a->nlevels;
    Layout L;
    const size_t o_keys = L.add(sizeof(orbgpu_keypoint) * (size_t)n2), o_desc = L.add(32 * (size_t)n2), o_ur = L.add(4 * (size_t)n2);
    const size_t o_qdesc = L.add(32 * (size_t)n1), o_pres = L.add((size_t)n1), o_pos = L.add(12 * (size_t)n1), o_nrm = L.add(12 * (size_t)n1);
    const size_t o_maxd = L.add(4 * (size_t)n1), o_mind = L.add(4 * (size_t)n1), o_occ = L.add((size_t)n2);
    const size_t o_scale = L.add(4 * (size_t)a->nlevels), o_sig = L.add(4 * (size_t)a->nlevels);
    const size_t in_bytes = L.off;
    const size_t o_best = L.add(4 * (size_t)nout), o_nm = L.add(4);
    const size_t io_bytes = L.off;
    const size_t o_start = L.add(4 * (GRID_CELLS + 1)), o_sorted = L.add(4 * (size_t)n2), o_q = L.add(sizeof(FuseQuery) * (size_t)n1);
    const size_t o_aq = L.add(occupied ? sizeof(AreaQuery) * (size_t)n1 : 0), o_cn = L.add(occupied ? 4 * (size_t)n1 : 0);
    const size_t o_acc = L.add(occupied ? 4 * (size_t)n1 : 0), o_cand = L.add(occupied ? 4 * (size_t)n1 * n2 : 0);
    Workspace& W = g_ws;
    if (job && phase == 1) { job->bytes = L.off; job->o_best = o_best; job->nout = nout; job->active = true; return ORBGPU_OK; }
    if (!job) {
        const int rc = W.reserve(device, L.off, io_bytes);
        if (rc != ORBGPU_OK) return rc;
    }
    uint8_t* const D = W.d + (job ? job->base : 0);
    uint8_t* const H = W.h + (job ? job->base : 0);
    memcpy(H + o_keys, a->keys_un, sizeof(orbgpu_keypoint) * (size_t)n2);
    memcpy(H + o_desc, a->desc, 32 * (size_t)n2);
    if (a->u_right) memcpy(H + o_ur, a->u_right, 4 * (size_t)n2); else memset(H + o_ur, 0, 4 * (size_t)n2);
    memcpy(H + o_qdesc, a->mp_desc, 32 * (size_t)n1);
    memcpy(H + o_pres, a->present, (size_t)n1);
    memcpy(H + o_pos, a->world_pos, 12 * (size_t)n1);
    if (a->normal) memcpy(H + o_nrm, a->normal, 12 * (size_t)n1);
    memcpy(H + o_maxd, a->max_distance, 4 * (size_t)n1);
    memcpy(H + o_mind, a->min_distance, 4 * (size_t)n1);
    if (occupied) memcpy(H + o_occ, occupied, (size_t)n2);
    memcpy(H + o_scale, a->scale_factors, 4 * (size_t)a->nlevels);
    if (a->inv_level_sigma2) memcpy(H + o_sig, a->inv_level_sigma2, 4 * (size_t)a->nlevels);
    CUS(cudaMemcpyAsync(D, H, in_bytes, cudaMemcpyHostToDevice, W.st));
    GridDev G;
    set_grid(G, D, o_keys, o_desc, o_start, o_sorted, n2, a->bounds);
    G.winX = C.minX; G.winY = C.minY;                          // src/KeyFrame.cc:574-586 subtract the int members
    FuseQuery* d_q = reinterpret_cast<FuseQuery*>(D + o_q);
    AreaQuery* d_aq = occupied ? reinterpret_cast<AreaQuery*>(D + o_aq) : nullptr;
    k_grid_build<<<1, 1024, 4 * (size_t)G.n, W.st>>>(G);
    k_fuse_queries<<<(n1 + 127) / 128, 128, 0, W.st>>>(C, D + o_pres, reinterpret_cast<const float*>(D + o_pos), reinterpret_cast<const float*>(D + o_nrm),
                                                       reinterpret_cast<const float*>(D + o_maxd), reinterpret_cast<const float*>(D + o_mind),
                                                       reinterpret_cast<const float*>(D + o_scale), n1, d_q, d_aq);
    if (!occupied) {
        k_fuse_match<<<(n1 + 7) / 8, 256, 0, W.st>>>(G, d_q, D + o_qdesc, n1, reinterpret_cast<const float*>(D + o_ur),
                                                     reinterpret_cast<const float*>(D + o_sig), chi2_gate, (unsigned)max_dist,
                                                     reinterpret_cast<int*>(D + o_best));
    } else {                                                   // vpMatched[idx] blocks a keypoint; an assignment blocks it for later points (:375, :397)
        uint32_t* d_cand = reinterpret_cast<uint32_t*>(D + o_cand);
        int* d_cn = reinterpret_cast<int*>(D + o_cn);
        k_area_candidates<<<(n1 + 7) / 8, 256, 0, W.st>>>(G, d_aq, D + o_qdesc, n1, nullptr, d_cand, n2, d_cn);
        k_proj_resolve<<<1, RESOLVE_THREADS, RESOLVE_SMEM, W.st>>>(d_cand, d_cn, n2, n1, n2, nullptr, D + o_occ, max_dist,
                                                                   reinterpret_cast<int*>(D + o_best), reinterpret_cast<int*>(D + o_acc),
                                                                   reinterpret_cast<int*>(D + o_nm));
    }
    CUS(cudaGetLastError());
    CUS(cudaMemcpyAsync(H + o_best, D + o_best, io_bytes - o_best, cudaMemcpyDeviceToHost, W.st));
    if (job) return ORBGPU_OK;
    CUS(cudaStreamSynchronize(W.st));
    memcpy(best, H + o_best, 4 * (size_t)nout);
    if (nmatches && occupied) *nmatches = *reinterpret_cast<const int*>(H + o_nm);
    return ORBGPU_OK;
}

static void cam_from_pose(FuseCam& C, const float* T16)
{
    memset(&C, 0, sizeof(C));
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) C.R[3 * r + c] = T16[4 * r + c]; C.t[r] = T16[4 * r + 3]; }
}

// "Decompose Scw" (src/ORBmatcher.cc:299-303, :979-984): scw = sqrt(row0 . row0) (cv::Mat::dot sums in double), Rcw = sRcw/scw and
// tcw = t/scw through convertTo with scale 1/scw (each element times (float)(1.0/scw), in float), Ow = -Rcw.t()*tcw
static void cam_from_sim3(FuseCam& C, const float* S16)
{
    memset(&C, 0, sizeof(C));
    double d = 0;
    for (int c = 0; c < 3; ++c) d += (double)S16[c] * (double)S16[c];
    const float scw = (float)sqrt(d);
    const float inv = (float)(1.0 / (double)scw);
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) { volatile float v = S16[4 * r + c] * inv; C.R[3 * r + c] = v; }
        volatile float v = S16[4 * r + 3] * inv; C.t[r] = v;
    }
    for (int i = 0; i < 3; ++i) {
        volatile float t = C.R[i] * C.t[0];
        t = t + C.R[3 + i] * C.t[1];
        t = t + C.R[6 + i] * C.t[2];
        C.Ow[i] = (float)((double)t * -1.0);
    }
}

int orbgpu_fuse_search(int device, const orbgpu_fuse_args* a, int32_t* best_idx)
{
    if (!a || !a->Tcw) return ORBGPU_ERR_ARG;
    FuseCam C;
    cam_from_pose(C, a->Tcw);
    for (int r = 0; r < 3; ++r) C.Ow[r] = a->Ow[r];
    C.view_check = 1;
    return fuse_run(device, a, C, 1, 50, nullptr, best_idx, nullptr);
}

int orbgpu_fuse_search_batch(int device, const orbgpu_fuse_args* args, int n_jobs, int32_t* best_idx)
{
    if (n_jobs < 0) return ORBGPU_ERR_ARG;
    if (n_jobs == 0) return ORBGPU_OK;
    if (!args || !best_idx) return ORBGPU_ERR_ARG;
    std::vector<FuseJob> jobs((size_t)n_jobs);
    std::vector<FuseCam> cams((size_t)n_jobs);
    std::vector<size_t> out_off((size_t)n_jobs);
    size_t total = 0, out = 0;
    for (int p = 0; p < n_jobs; ++p) {
        const orbgpu_fuse_args* a = args + p;
        if (!a->Tcw || a->n_points < 0) return ORBGPU_ERR_ARG;
        FuseCam& C = cams[(size_t)p];
        cam_from_pose(C, a->Tcw);
        for (int r = 0; r < 3; ++r) C.Ow[r] = a->Ow[r];
        C.view_check = 1;
        out_off[(size_t)p] = out;
        const int rc = fuse_run(device, a, C, 1, 50, nullptr, best_idx + out, nullptr, &jobs[(size_t)p], 1);
        if (rc != ORBGPU_OK) return rc;
        out += (size_t)a->n_points;
        jobs[(size_t)p].base = total;
        total += jobs[(size_t)p].bytes;
    }
    if (total == 0) return ORBGPU_OK;
    Workspace& W = g_ws;
    int rc = W.reserve(device, total, total);
    if (rc != ORBGPU_OK) return rc;
    for (int p = 0; p < n_jobs; ++p) {
        if (!jobs[(size_t)p].active) continue;
        rc = fuse_run(device, args + p, cams[(size_t)p], 1, 50, nullptr, best_idx + out_off[(size_t)p], nullptr, &jobs[(size_t)p], 2);
        if (rc != ORBGPU_OK) { cudaStreamSynchronize(W.st); return rc; }
    }
    CUS(cudaStreamSynchronize(W.st));
    for (int p = 0; p < n_jobs; ++p)
        if (jobs[(size_t)p].active)
            memcpy(best_idx + out_off[(size_t)p], W.h + jobs[(size_t)p].base + jobs[(size_t)p].o_best, 4 * (size_t)jobs[(size_t)p].nout);
    return ORBGPU_OK;
}

int orbgpu_fuse_search_sim3(int device, const orbgpu_fuse_args* a, const float* Scw, int32_t* best_idx)
{
    if (!a || !Scw) return ORBGPU_ERR_ARG;
    FuseCam C;
    cam_from_sim3(C, Scw);
    C.view_check = 1; C.inv_double = 1;
    return fuse_run(device, a, C, 0, 50, nullptr, best_idx, nullptr);
}

int orbgpu_search_by_projection_sim3(int device, const orbgpu_fuse_args* a, const float* Scw, const uint8_t* matched_in,
                                     int32_t* matched, int* nmatches)
{
    if (!a || !Scw || !matched_in || !nmatches) return ORBGPU_ERR_ARG;
    FuseCam C;
    cam_from_sim3(C, Scw);
    C.view_check = 1;
    return fuse_run(device, a, C, 0, 50, matched_in, matched, nmatches);
}

int orbgpu_search_by_sim3(int device, const orbgpu_fuse_args* a12, const orbgpu_fuse_args* a21, float s12, const float* R12,
                          const float* t12, int32_t* match12, int* nfound)
{
    if (!a12 || !a21 || !R12 || !t12 || !match12 || !nfound || !a12->Tcw || !a21->Tcw) return ORBGPU_ERR_ARG;
    if (a12->n_points != a21->n_kf || a21->n_points != a12->n_kf) return ORBGPU_ERR_ARG;
    // "Transformation between cameras" (:1117-1120): sR12 = s12*R12 and sR21 = (1.0/s12)*R12.t() are convertTo with a
    // scale (element times (float)scale, in float); t21 = -sR21*t12 is gemm's 3x3 float case with alpha = -1
    float sR12[9], sR21[9], t21[3];
    const float inv = (float)(1.0 / (double)s12);
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            volatile float v = R12[3 * r + c] * (float)(double)s12; sR12[3 * r + c] = v;
            volatile float w = R12[3 * c + r] * inv; sR21[3 * r + c] = w;
        }
    for (int i = 0; i < 3; ++i) {
        volatile float t = sR21[3 * i] * t12[0];
        t = t + sR21[3 * i + 1] * t12[1];
        t = t + sR21[3 * i + 2] * t12[2];
        t21[i] = (float)((double)t * -1.0);
    }
    const int N1 = a12->n_points, N2 = a21->n_points;
    std::vector<int32_t> m1((size_t)(N1 > 0 ? N1 : 1)), m2((size_t)(N2 > 0 ? N2 : 1));
    FuseCam C;
    cam_from_pose(C, a12->Tcw);                                // KF1 -> KF2 (:1147-1225)
    for (int k = 0; k < 9; ++k) C.R2[k] = sR21[k];
    for (int k = 0; k < 3; ++k) C.t2[k] = t21[k];
    C.two_stage = 1; C.inv_double = 1; C.cam_dist = 1;
    int rc = fuse_run(device, a12, C, 0, 100, nullptr, m1.data(), nullptr);
    if (rc != ORBGPU_OK) return rc;
    cam_from_pose(C, a21->Tcw);                                // KF2 -> KF1 (:1227-1305)
    for (int k = 0; k < 9; ++k) C.R2[k] = sR12[k];
    for (int k = 0; k < 3; ++k) C.t2[k] = t12[k];
    C.two_stage = 1; C.inv_double = 1; C.cam_dist = 1;
    rc = fuse_run(device, a21, C, 0, 100, nullptr, m2.data(), nullptr);
    if (rc != ORBGPU_OK) return rc;
    int nf = 0;                                                // "Check agreement" (:1307-1323)
    for (int i1 = 0; i1 < N1; ++i1) {
        match12[i1] = -1;
        const int idx2 = m1[i1];
        if (idx2 >= 0 && m2[idx2] == i1) { match12[i1] = idx2; nf++; }
    }
    *nfound = nf;
    return ORBGPU_OK;
}

int orbgpu_search_by_projection_keyframe(int device, const orbgpu_reloc_projection_args* a, int32_t* cur_match, int* nmatches)
{
    if (!a || !cur_match || !nmatches || a->n_kf < 0 || a->n_cur < 0 || a->nlevels < 1) return ORBGPU_ERR_ARG;
    if (!a->mp_valid || !a->world_pos || !a->max_distance || !a->min_distance || !a->mp_desc || !a->kf_angle || !a->cur_keys_un ||
        !a->cur_desc || !a->cur_occupied || !a->Tcw_cur || !a->scale_factors) return ORBGPU_ERR_ARG;
    if (!(a->bounds[1] > a->bounds[0]) || !(a->bounds[3] > a->bounds[2])) return ORBGPU_ERR_ARG;
    const int n1 = a->n_kf, n2 = a->n_cur;
    if (n1 > GRID_MAX_KEYS || n2 > GRID_MAX_KEYS || (size_t)n1 * (size_t)n2 > ((size_t)1 << 26)) return ORBGPU_ERR_UNSUPPORTED;
    *nmatches = 0;
    for (int i = 0; i < n2; ++i) cur_match[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBGPU_OK;
    RelocCam C;
    const float* Tc = a->Tcw_cur;
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) C.R[3 * r + c] = Tc[4 * r + c]; C.t[r] = Tc[4 * r + 3]; }
    for (int i = 0; i < 3; ++i) {                              // Ow = -Rcw.t()*tcw (:1478), the cv::MatExpr of :1341
        volatile float t = Tc[i] * C.t[0];
        t = t + Tc[4 + i] * C.t[1];
        t = t + Tc[8 + i] * C.t[2];
        C.Ow[i] = (float)((double)t * -1.0);
    }
    C.fx = a->fx; C.fy = a->fy; C.cx = a->cx; C.cy = a->cy;
    C.minX = a->bounds[0]; C.maxX = a->bounds[1]; C.minY = a->bounds[2]; C.maxY = a->bounds[3];
    C.log_scale = a->log_scale_factor; C.th = a->th; C.nlevels = a->nlevels;
    Layout L;
    const size_t o_keys = L.add(sizeof(orbgpu_keypoint) * (size_t)n2), o_desc = L.add(32 * (size_t)n2), o_occ = L.add((size_t)n2);
    const size_t o_pos = L.add(12 * (size_t)n1), o_qdesc = L.add(32 * (size_t)n1), o_valid = L.add((size_t)n1);
    const size_t o_maxd = L.add(4 * (size_t)n1), o_mind = L.add(4 * (size_t)n1), o_scale = L.add(4 * (size_t)a->nlevels);
    const size_t in_bytes = L.off;
    const size_t o_cm = L.add(4 * (size_t)n2), o_acc = L.add(4 * (size_t)n1), o_nm = L.add(4);
    const size_t io_bytes = L.off;
    const size_t o_start = L.add(4 * (GRID_CELLS + 1)), o_sorted = L.add(4 * (size_t)n2), o_cn = L.add(4 * (size_t)n1);
    const size_t o_q = L.add(sizeof(AreaQuery) * (size_t)n1), o_cand = L.add(4 * (size_t)n1 * n2);
    Workspace& W = g_ws;
    int rc = W.reserve(device, L.off, io_bytes);
    if (rc != ORBGPU_OK) return rc;
    memcpy(W.h + o_keys, a->cur_keys_un, sizeof(orbgpu_keypoint) * (size_t)n2);
    memcpy(W.h + o_desc, a->cur_desc, 32 * (size_t)n2);
    memcpy(W.h + o_occ, a->cur_occupied, (size_t)n2);
    memcpy(W.h + o_pos, a->world_pos, 12 * (size_t)n1);
    memcpy(W.h + o_qdesc, a->mp_desc, 32 * (size_t)n1);
    memcpy(W.h + o_valid, a->mp_valid, (size_t)n1);
    memcpy(W.h + o_maxd, a->max_distance, 4 * (size_t)n1);
    memcpy(W.h + o_mind, a->min_distance, 4 * (size_t)n1);
    memcpy(W.h + o_scale, a->scale_factors, 4 * (size_t)a->nlevels);
    CUS(cudaMemcpyAsync(W.d, W.h, in_bytes, cudaMemcpyHostToDevice, W.st));
    GridDev G;
    set_grid(G, W.d, o_keys, o_desc, o_start, o_sorted, n2, a->bounds);
    uint32_t* d_cand = reinterpret_cast<uint32_t*>(W.d + o_cand);
    int* d_cn = reinterpret_cast<int*>(W.d + o_cn);
    AreaQuery* d_q = reinterpret_cast<AreaQuery*>(W.d + o_q);
    k_grid_build<<<1, 1024, 4 * (size_t)G.n, W.st>>>(G);
    k_reloc_queries<<<(n1 + 127) / 128, 128, 0, W.st>>>(C, W.d + o_valid, reinterpret_cast<const float*>(W.d + o_pos),
                                                        reinterpret_cast<const float*>(W.d + o_maxd), reinterpret_cast<const float*>(W.d + o_mind),
                                                        reinterpret_cast<const float*>(W.d + o_scale), n1, d_q);
    k_area_candidates<<<(n1 + 7) / 8, 256, 0, W.st>>>(G, d_q, W.d + o_qdesc, n1, nullptr, d_cand, n2, d_cn);
    // every assigned feature holds a map point afterwards (:1545-1546): "observed" is all ones
    k_proj_resolve<<<1, RESOLVE_THREADS, RESOLVE_SMEM, W.st>>>(d_cand, d_cn, n2, n1, n2, nullptr, W.d + o_occ, a->orb_dist < 255 ? a->orb_dist : 255,
                                                               reinterpret_cast<int*>(W.d + o_cm), reinterpret_cast<int*>(W.d + o_acc),
                                                               reinterpret_cast<int*>(W.d + o_nm));
    CUS(cudaGetLastError());
    CUS(cudaMemcpyAsync(W.h + in_bytes, W.d + in_bytes, io_bytes - in_bytes, cudaMemcpyDeviceToHost, W.st));
    CUS(cudaStreamSynchronize(W.st));
    memcpy(cur_match, W.h + o_cm, 4 * (size_t)n2);
    const int* acc = reinterpret_cast<const int*>(W.h + o_acc);
    int nm = *reinterpret_cast<const int*>(W.h + o_nm);
    if (a->check_orientation) {                                // :1566-1587
        int bin_of[30] = {0};
        std::vector<int> bin((size_t)n1, -1);
        for (int i = 0; i < n1; ++i)
            if (acc[i] >= 0) { bin[i] = rot_bin(a->kf_angle[i], a->cur_keys_un[acc[i]].angle); bin_of[bin[i]]++; }
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(bin_of, 30, ind1, ind2, ind3);
        for (int i = 0; i < n1; ++i) {
            if (bin[i] < 0 || bin[i] == ind1 || bin[i] == ind2 || bin[i] == ind3) continue;
            cur_match[acc[i]] = -2;
            nm--;
        }
    }
    *nmatches = nm;
    return ORBGPU_OK;
}

}  // extern "C"

// ---- the two BoW-gated searches on "side views" ---------------------------------------------------------------------
// A side is what the engine reads of one KeyFrame / Frame.  The arrays that gate the search (FeatureVector, map-point
// flags, angles) are read on the host; keypoints, descriptors, mvuRight and the FeatureVector's index list are either
// host arrays uploaded with the call (orbgpu_search_by_bow, orbgpu_search_for_triangulation) or already resident in HBM
// (orbgpu_kf_* below: a keyframe's mvKeysUn / mDescriptors / mvuRight / mFeatVec never change after construction).
struct SideView {
    int n = 0, n_nodes = 0;
    const uint32_t* node_ids = nullptr; const int32_t* node_start = nullptr; const uint32_t* indices = nullptr;   // host
    const uint8_t* flags = nullptr;                  // host: "valid" (SearchByBoW) / "has map point" (SearchForTriangulation)
    const float* angle = nullptr; int angle_stride = 1;                            // host, floats
    const float* u_right = nullptr;                  // host (only_stereo test of side 1)
    const orbgpu_keypoint* h_keys = nullptr; const uint8_t* h_desc = nullptr;      // uploaded per call ...
    const float* d_keys = nullptr; const uint8_t* d_desc = nullptr;                // ... or resident
    const float* d_ur = nullptr; const uint32_t* d_indices = nullptr;
    bool resident() const { return d_desc != nullptr; }
    int nidx() const { return n_nodes > 0 ? node_start[n_nodes] : 0; }
};

static int check_feature_vector(int n, const uint32_t* node_ids, const int32_t* node_start, const uint32_t* indices, int n_nodes)
{
    if (n_nodes < 0 || (n_nodes > 0 && (!node_ids || !node_start || !indices))) return ORBGPU_ERR_ARG;
    if (n_nodes > 0 && node_start[0] != 0) return ORBGPU_ERR_ARG;
    for (int i = 0; i < n_nodes; ++i) {
        if (node_start[i] > node_start[i + 1] || (i && node_ids[i - 1] >= node_ids[i])) return ORBGPU_ERR_ARG;
        for (int c = node_start[i]; c < node_start[i + 1]; ++c) if (indices[c] >= (uint32_t)n) return ORBGPU_ERR_ARG;
    }
    return ORBGPU_OK;
}

// One SearchForTriangulation problem, split so that several (LocalMapping::CreateNewMapPoints runs one per neighbouring
// keyframe, src/LocalMapping.cc:239-270) can be enqueued back to back on the workspace stream and waited for once.
struct TriProblem {
    SideView s1, s2;
    const float* F12 = nullptr; const float* Cw1 = nullptr; const float* T2w = nullptr;
    float fx = 0, fy = 0, cx = 0, cy = 0;            // pKF2
    const float* h_scale2 = nullptr; const float* h_sigma2 = nullptr;              // uploaded per call ...
    const float* d_scale2 = nullptr; const float* d_sigma2 = nullptr;              // ... or resident
    int only_stereo = 0, check_orientation = 0;
};
struct TriJob {
    TriProblem P;
    std::vector<BowQuery> hq;
    TriCam C;
    int nq = 0;
    size_t base = 0, in_bytes = 0;                   // the job's inputs inside the workspace: [base, base + in_bytes)
    size_t out = 0;                                  // its results (one int per query) in the result area shared by all jobs
    size_t o_k1 = 0, o_d1 = 0, o_u1 = 0, o_k2 = 0, o_d2 = 0, o_u2 = 0, o_m2 = 0, o_i2 = 0, o_q = 0, o_sc = 0, o_sg = 0;
};

// the merge walk over the two FeatureVectors, the epipole; lays the job's region out (inputs already validated)
static int tri_prepare(TriJob& J, int32_t* matches12, int* nmatches)
{
    const TriProblem& P = J.P;
    const SideView &s1 = P.s1, &s2 = P.s2;
    std::vector<BowQuery>& hq = J.hq;
    const int n1 = s1.n, n2 = s2.n;
    *nmatches = 0;
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    // the merge walk over the two FeatureVectors (:689-696, :782-789) -> queries in visiting order
    int it1 = 0, it2 = 0;
    while (it1 != s1.n_nodes && it2 != s2.n_nodes) {
        if (s1.node_ids[it1] == s2.node_ids[it2]) {
            const int st = s2.node_start[it2], len = s2.node_start[it2 + 1] - st;
            for (int c = s1.node_start[it1]; c < s1.node_start[it1 + 1]; ++c) {
                const int idx1 = (int)s1.indices[c];
                if (s1.flags[idx1] || len == 0) continue;                              // :703-706
                if (P.only_stereo && !(s1.u_right[idx1] >= 0)) continue;               // :708-712
                BowQuery q; q.idx1 = idx1; q.start = st; q.len = len;
                hq.push_back(q);
            }
            ++it1; ++it2;
        } else if (s1.node_ids[it1] < s2.node_ids[it2]) {
            while (it1 != s1.n_nodes && s1.node_ids[it1] < s2.node_ids[it2]) ++it1;
        } else {
            while (it2 != s2.n_nodes && s2.node_ids[it2] < s1.node_ids[it1]) ++it2;
        }
    }
    J.nq = (int)hq.size();
    if (J.nq == 0) return ORBGPU_OK;
    const int nidx2 = s2.nidx();
    if (nidx2 > 0xFFFF) return ORBGPU_ERR_UNSUPPORTED;
    TriCam& C = J.C;
    // epipole of pKF1's centre in pKF2 (:664-670): C2 = R2w*Cw + t2w is cv::gemm's 3x3 float case
    float C2[3];
    for (int r = 0; r < 3; ++r) {
        volatile float t = P.T2w[4 * r] * P.Cw1[0];            // (volatile: keep the host compiler from contracting)
        t = t + P.T2w[4 * r + 1] * P.Cw1[1];
        t = t + P.T2w[4 * r + 2] * P.Cw1[2];
        C2[r] = (float)((double)t * 1.0 + (double)P.T2w[4 * r + 3] * 1.0);
    }
    {
        const float invz = 1.0f / C2[2];
        volatile float px = P.fx * C2[0]; px = px * invz; px = px + P.cx;
        volatile float py = P.fy * C2[1]; py = py * invz; py = py + P.cy;
        C.ex = px; C.ey = py;
    }
    for (int i = 0; i < 9; ++i) C.F[i] = P.F12[i];
    C.only_stereo = P.only_stereo ? 1 : 0;

    Layout L;
    if (!s1.resident()) { J.o_k1 = L.add(sizeof(orbgpu_keypoint) * (size_t)n1); J.o_d1 = L.add(32 * (size_t)n1); J.o_u1 = L.add(4 * (size_t)n1); }
    if (!s2.resident()) {
        J.o_k2 = L.add(sizeof(orbgpu_keypoint) * (size_t)n2); J.o_d2 = L.add(32 * (size_t)n2); J.o_u2 = L.add(4 * (size_t)n2);
        J.o_i2 = L.add(4 * (size_t)nidx2);
    }
    if (!P.d_scale2) { J.o_sc = L.add(4 * (size_t)ORBGPU_MAX_LEVELS); J.o_sg = L.add(4 * (size_t)ORBGPU_MAX_LEVELS); }
    J.o_m2 = L.add((size_t)n2); J.o_q = L.add(sizeof(BowQuery) * (size_t)J.nq);
    J.in_bytes = L.off;
    return ORBGPU_OK;
}

// launch == false: stage the job's inputs in the pinned mirror (tri_run uploads the inputs of all jobs with one copy);
// launch == true: the job's entry of the device job table (device pointers of its operands, results into the result area
// shared by all jobs)
static int tri_enqueue(Workspace& W, const TriJob& J, int nlevels, bool launch, TriJobDev* table)
{
    if (J.nq == 0) { if (launch) { memset(table, 0, sizeof(*table)); } return ORBGPU_OK; }
    const TriProblem& P = J.P;
    const SideView &s1 = P.s1, &s2 = P.s2;
    const int n1 = s1.n, n2 = s2.n;
    uint8_t* h = W.h + J.base;
    uint8_t* d = W.d + J.base;
    const float *k1 = s1.d_keys, *u1 = s1.d_ur, *k2 = s2.d_keys, *u2 = s2.d_ur, *sc = P.d_scale2, *sg = P.d_sigma2;
    const uint8_t *d1 = s1.d_desc, *d2 = s2.d_desc;
    const uint32_t* i2 = s2.d_indices;
    if (!s1.resident()) {
        if (!launch) {
            memcpy(h + J.o_k1, s1.h_keys, sizeof(orbgpu_keypoint) * (size_t)n1);
            memcpy(h + J.o_d1, s1.h_desc, 32 * (size_t)n1);
            memcpy(h + J.o_u1, s1.u_right, 4 * (size_t)n1);
        }
        k1 = reinterpret_cast<const float*>(d + J.o_k1); d1 = d + J.o_d1; u1 = reinterpret_cast<const float*>(d + J.o_u1);
    }
    if (!s2.resident()) {
        if (!launch) {
            memcpy(h + J.o_k2, s2.h_keys, sizeof(orbgpu_keypoint) * (size_t)n2);
            memcpy(h + J.o_d2, s2.h_desc, 32 * (size_t)n2);
            memcpy(h + J.o_u2, s2.u_right, 4 * (size_t)n2);
            memcpy(h + J.o_i2, s2.indices, 4 * (size_t)s2.nidx());
        }
        k2 = reinterpret_cast<const float*>(d + J.o_k2); d2 = d + J.o_d2; u2 = reinterpret_cast<const float*>(d + J.o_u2);
        i2 = reinterpret_cast<const uint32_t*>(d + J.o_i2);
    }
    if (!P.d_scale2) {
        if (!launch) {
            memcpy(h + J.o_sc, P.h_scale2, 4 * (size_t)nlevels);
            memcpy(h + J.o_sg, P.h_sigma2, 4 * (size_t)nlevels);
        }
        sc = reinterpret_cast<const float*>(d + J.o_sc); sg = reinterpret_cast<const float*>(d + J.o_sg);
    }
    if (!launch) {
        memcpy(h + J.o_m2, s2.flags, (size_t)n2);
        memcpy(h + J.o_q, J.hq.data(), sizeof(BowQuery) * (size_t)J.nq);
        return ORBGPU_OK;
    }
    TriJobDev& T = *table;
    T.Q = reinterpret_cast<const BowQuery*>(d + J.o_q); T.nq = J.nq; T.C = J.C;
    T.keys1 = k1; T.desc1 = d1; T.ur1 = u1; T.keys2 = k2; T.desc2 = d2; T.ur2 = u2; T.has_mp2 = d + J.o_m2; T.indices2 = i2;
    T.scale2 = sc; T.sigma2 = sg; T.result = reinterpret_cast<int*>(W.d + J.out);
    return ORBGPU_OK;
}

// after the stream has drained: results out of the mirror, the rotation histogram on the host
static void tri_finish(const Workspace& W, const TriJob& J, int32_t* matches12, int* nmatches)
{
    if (J.nq == 0) return;
    const SideView &s1 = J.P.s1, &s2 = J.P.s2;
    const std::vector<BowQuery>& hq = J.hq;
    const int nq = J.nq;
    const int* res = reinterpret_cast<const int*>(W.h + J.out);                  // matches12 is all -1 since tri_prepare
    int nm = 0;
    for (int q = 0; q < nq; ++q) { matches12[hq[q].idx1] = res[q]; nm += res[q] >= 0; }
    if (J.P.check_orientation) {                               // :760-776, :794-813
        int bin_of[30] = {0};
        std::vector<int> bin((size_t)nq, -1);
        for (int q = 0; q < nq; ++q) {
            const int i1 = hq[q].idx1;
            if (matches12[i1] >= 0) {
                bin[q] = rot_bin(s1.angle[(size_t)i1 * s1.angle_stride], s2.angle[(size_t)matches12[i1] * s2.angle_stride]);
                bin_of[bin[q]]++;
            }
        }
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(bin_of, 30, ind1, ind2, ind3);
        for (int q = 0; q < nq; ++q) {
            if (bin[q] < 0 || bin[q] == ind1 || bin[q] == ind2 || bin[q] == ind3) continue;
            matches12[hq[q].idx1] = -1;
            nm--;
        }
    }
    *nmatches = nm;
}

// All jobs in ONE upload, one kernel each, ONE download and one wait (a copy costs ~10 us of latency on its own, whatever its
// size: twenty jobs with their own copies spent most of their time in them).  Job p writes s1.n entries of matches12 right
// after those of job p - 1.
static int tri_run(int device, std::vector<TriJob>& jobs, const int* nlevels, int32_t* matches12, int* nmatches)
{
    const size_t n_jobs = jobs.size();
    PhaseTrace tr("search_for_triangulation");
    std::vector<size_t> out_off(n_jobs);
    size_t total_in = (sizeof(TriJobDev) * n_jobs + 255) & ~(size_t)255, out = 0, total_q = 0;      // the job table first
    int max_q = 0;
    for (size_t p = 0; p < n_jobs; ++p) {
        TriJob& J = jobs[p];
        out_off[p] = out;
        const int rc = tri_prepare(J, matches12 + out, nmatches + p);
        if (rc != ORBGPU_OK) return rc;
        out += (size_t)J.P.s1.n;
        J.base = total_in;
        total_in += (J.in_bytes + 255) & ~(size_t)255;
        total_q += (size_t)J.nq;
        max_q = std::max(max_q, J.nq);
    }
    if (total_q == 0) return ORBGPU_OK;
    const size_t up_bytes = std::max(total_in, SEARCH_MIN_UPLOAD);
    size_t o = up_bytes;
    for (size_t p = 0; p < n_jobs; ++p) { jobs[p].out = o; o += 4 * (size_t)jobs[p].nq; }
    tr.mark("walk");
    Workspace& W = g_ws;
    int rc = W.reserve(device, o, o);
    if (rc != ORBGPU_OK) return rc;
    TriJobDev* table = reinterpret_cast<TriJobDev*>(W.h);
    for (size_t p = 0; p < n_jobs; ++p) { tri_enqueue(W, jobs[p], nlevels[p], false, nullptr); tri_enqueue(W, jobs[p], nlevels[p], true, table + p); }
    tr.mark("stage");
    CUS(cudaMemcpyAsync(W.d, W.h, up_bytes, cudaMemcpyHostToDevice, W.st));
    k_tri_match<<<dim3((unsigned)((max_q + 7) / 8), (unsigned)n_jobs), 256, 0, W.st>>>(reinterpret_cast<const TriJobDev*>(W.d));
    CUS(cudaGetLastError());
    CUS(cudaMemcpyAsync(W.h + up_bytes, W.d + up_bytes, 4 * total_q, cudaMemcpyDeviceToHost, W.st));
    tr.mark("enqueue");
    CUS(cudaStreamSynchronize(W.st));
    tr.mark("wait");
    for (size_t p = 0; p < n_jobs; ++p) tri_finish(W, jobs[p], matches12 + out_off[p], nmatches + p);
    tr.mark("finish");
    return ORBGPU_OK;
}

static void view_of(SideView& v, const orbgpu_triangulation_side& s)
{
    v.n = s.n; v.n_nodes = s.n_nodes; v.node_ids = s.node_ids; v.node_start = s.node_start; v.indices = s.indices;
    v.flags = s.has_map_point; v.u_right = s.u_right; v.h_keys = s.keys_un; v.h_desc = s.desc;
    v.angle = s.keys_un ? &s.keys_un[0].angle : nullptr; v.angle_stride = (int)(sizeof(orbgpu_keypoint) / sizeof(float));
}

extern "C" {

int orbgpu_search_for_triangulation_batch(int device, const orbgpu_triangulation_args* args, int n_pairs, int32_t* matches12,
                                          int* nmatches)
{
    if (n_pairs < 0) return ORBGPU_ERR_ARG;
    if (n_pairs == 0) return ORBGPU_OK;
    if (!args || !matches12 || !nmatches) return ORBGPU_ERR_ARG;
    std::vector<TriJob> jobs((size_t)n_pairs);
    std::vector<int> nlevels((size_t)n_pairs);
    for (int p = 0; p < n_pairs; ++p) {
        const orbgpu_triangulation_args* a = args + p;
        if (a->nlevels < 1 || a->nlevels > ORBGPU_MAX_LEVELS || !a->scale_factors2 || !a->level_sigma2) return ORBGPU_ERR_ARG;
        const orbgpu_triangulation_side* S[2] = {&a->kf1, &a->kf2};
        for (int k = 0; k < 2; ++k) {
            const orbgpu_triangulation_side* s = S[k];
            if (s->n < 0 || (s->n > 0 && (!s->keys_un || !s->desc || !s->has_map_point || !s->u_right))) return ORBGPU_ERR_ARG;
            if (s->n > GRID_MAX_KEYS) return ORBGPU_ERR_UNSUPPORTED;
            const int rc = check_feature_vector(s->n, s->node_ids, s->node_start, s->indices, s->n_nodes);
            if (rc != ORBGPU_OK) return rc;
        }
        for (int i = 0; i < a->kf2.n; ++i)
            if (a->kf2.keys_un[i].octave < 0 || a->kf2.keys_un[i].octave >= a->nlevels) return ORBGPU_ERR_ARG;
        TriProblem& P = jobs[(size_t)p].P;
        view_of(P.s1, a->kf1); view_of(P.s2, a->kf2);
        P.F12 = a->F12; P.Cw1 = a->Cw1; P.T2w = a->T2w; P.fx = a->fx; P.fy = a->fy; P.cx = a->cx; P.cy = a->cy;
        P.h_scale2 = a->scale_factors2; P.h_sigma2 = a->level_sigma2;
        P.only_stereo = a->only_stereo; P.check_orientation = a->check_orientation;
        nlevels[(size_t)p] = a->nlevels;
    }
    return tri_run(device, jobs, nlevels.data(), matches12, nmatches);
}

int orbgpu_search_for_triangulation(int device, const orbgpu_triangulation_args* a, int32_t* matches12, int* nmatches)
{
    return orbgpu_search_for_triangulation_batch(device, a, a ? 1 : 0, matches12, nmatches);
}

}  // extern "C"

// SearchByBoW on two side views (inputs already validated)
static int bow_run(int device, const SideView& s1, const SideView& s2, float nnratio, int check_orientation, int strict_th_low,
                   int32_t* match12, int32_t* match21, int* nmatches)
{
    const int n1 = s1.n, n2 = s2.n;
    PhaseTrace tr("search_by_bow");
    *nmatches = 0;
    for (int i = 0; i < n1; ++i) match12[i] = -1;
    for (int i = 0; i < n2; ++i) match21[i] = -1;
    // the merge walk over the two FeatureVectors (:176-180, :254-266) -> queries in visiting order
    std::vector<BowQuery> hq;
    std::vector<int> gstart;                                   // query ranges per shared node
    int cap = 0, it1 = 0, it2 = 0;
    while (it1 != s1.n_nodes && it2 != s2.n_nodes) {
        if (s1.node_ids[it1] == s2.node_ids[it2]) {
            const int st = s2.node_start[it2], len = s2.node_start[it2 + 1] - st;
            gstart.push_back((int)hq.size());
            for (int a = s1.node_start[it1]; a < s1.node_start[it1 + 1]; ++a) {
                const int idx1 = (int)s1.indices[a];
                if (!s1.flags[idx1] || len == 0) continue;
                BowQuery q; q.idx1 = idx1; q.start = st; q.len = len;
                hq.push_back(q);
                cap = len > cap ? len : cap;
            }
            ++it1; ++it2;
        } else if (s1.node_ids[it1] < s2.node_ids[it2]) {
            while (it1 != s1.n_nodes && s1.node_ids[it1] < s2.node_ids[it2]) ++it1;
        } else {
            while (it2 != s2.n_nodes && s2.node_ids[it2] < s1.node_ids[it1]) ++it2;
        }
    }
    const int nq = (int)hq.size(), ngroups = (int)gstart.size();
    gstart.push_back(nq);
    if (nq == 0) return ORBGPU_OK;
    // one warp per node (k_bow_fused) unless some node is so large that its distance evaluations want the whole GPU
    long long heaviest = 0;
    for (int g = 0; g < ngroups; ++g)
        if (gstart[(size_t)g + 1] > gstart[(size_t)g])
            heaviest = std::max(heaviest, (long long)(gstart[(size_t)g + 1] - gstart[(size_t)g]) * ((hq[(size_t)gstart[(size_t)g]].len + 31) / 32));
    const bool fused = heaviest <= 256;
    if (nq > GRID_MAX_KEYS || (size_t)nq * (size_t)cap > ((size_t)1 << 26)) return ORBGPU_ERR_UNSUPPORTED;
    const int nidx2 = s2.nidx();
    Layout L;
    size_t o_d1 = 0, o_d2 = 0, o_i2 = 0;
    if (!s1.resident()) o_d1 = L.add(32 * (size_t)n1);
    if (!s2.resident()) { o_d2 = L.add(32 * (size_t)n2); o_i2 = L.add(4 * (size_t)nidx2); }
    const size_t o_v2 = L.add((size_t)n2);
    const size_t o_q = L.add(sizeof(BowQuery) * (size_t)nq), o_g = L.add(4 * (size_t)(ngroups + 1));
    const size_t in_bytes = L.off;
    const size_t o_acc = L.add(4 * (size_t)nq), o_nm = L.add(4);
    const size_t io_bytes = L.off;
    size_t o_cn = 0, o_cand = 0;
    if (!fused) { o_cn = L.add(4 * (size_t)nq); o_cand = L.add(4 * (size_t)nq * cap); }
    tr.mark("walk");
    Workspace& W = g_ws;
    int rc = W.reserve(device, std::max(L.off, SEARCH_MIN_UPLOAD), std::max(io_bytes, SEARCH_MIN_UPLOAD));
    if (rc != ORBGPU_OK) return rc;
    tr.mark("reserve");
    const uint8_t *d1 = s1.d_desc, *d2 = s2.d_desc;
    const uint32_t* i2 = s2.d_indices;
    if (!s1.resident()) { memcpy(W.h + o_d1, s1.h_desc, 32 * (size_t)n1); d1 = W.d + o_d1; }
    if (!s2.resident()) {
        memcpy(W.h + o_d2, s2.h_desc, 32 * (size_t)n2);
        memcpy(W.h + o_i2, s2.indices, 4 * (size_t)nidx2);
        d2 = W.d + o_d2; i2 = reinterpret_cast<const uint32_t*>(W.d + o_i2);
    }
    memcpy(W.h + o_v2, s2.flags, (size_t)n2);
    memcpy(W.h + o_q, hq.data(), sizeof(BowQuery) * (size_t)nq);
    memcpy(W.h + o_g, gstart.data(), 4 * (size_t)(ngroups + 1));
    tr.mark("stage");
    CUS(cudaMemcpyAsync(W.d, W.h, std::max(in_bytes, SEARCH_MIN_UPLOAD), cudaMemcpyHostToDevice, W.st));
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    if (tr.on) { cudaEventCreate(&ev0); cudaEventCreate(&ev1); cudaEventRecord(ev0, W.st); }
    if (fused && cap <= BOW_NODE_CAP) {
        k_bow_rounds<<<ngroups, 32 * BOW_ROUND_WARPS, 0, W.st>>>(reinterpret_cast<const BowQuery*>(W.d + o_q), reinterpret_cast<const int*>(W.d + o_g), d1, d2, i2,
                                               W.d + o_v2, nnratio, strict_th_low ? 50 : 51, reinterpret_cast<int*>(W.d + o_acc));
    } else if (fused) {
        k_bow_fused<<<(ngroups + 3) / 4, 128, 0, W.st>>>(reinterpret_cast<const BowQuery*>(W.d + o_q), reinterpret_cast<const int*>(W.d + o_g),
                                                          ngroups, d1, d2, i2, W.d + o_v2, nnratio, strict_th_low ? 50 : 51,
                                                          reinterpret_cast<int*>(W.d + o_acc));
    } else {
        uint32_t* d_cand = reinterpret_cast<uint32_t*>(W.d + o_cand);
        int* d_cn = reinterpret_cast<int*>(W.d + o_cn);
        k_bow_candidates<<<(nq + 7) / 8, 256, 0, W.st>>>(reinterpret_cast<const BowQuery*>(W.d + o_q), nq, d1, d2, i2, d_cand, cap, d_cn);
        k_bow_resolve<<<1, RESOLVE_THREADS, RESOLVE_SMEM, W.st>>>(d_cand, d_cn, cap, nq, n2, W.d + o_v2, reinterpret_cast<const int*>(W.d + o_g),
                                                                  ngroups, nnratio, strict_th_low ? 50 : 51,
                                                                  reinterpret_cast<int*>(W.d + o_acc), reinterpret_cast<int*>(W.d + o_nm));
    }
    CUS(cudaGetLastError());
    if (tr.on) cudaEventRecord(ev1, W.st);
    CUS(cudaMemcpyAsync(W.h + in_bytes, W.d + in_bytes, io_bytes - in_bytes, cudaMemcpyDeviceToHost, W.st));
    tr.mark("enqueue");
    CUS(cudaStreamSynchronize(W.st));
    tr.mark("wait");
    if (tr.on) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ev0, ev1);
        fprintf(stderr, "[orbgpu]   kernels %.1f us; nq %d groups %d longest list %d heaviest node %lld %s\n", ms * 1e3, nq, ngroups, cap,
                heaviest, fused && cap <= BOW_NODE_CAP ? "fixed-point rounds" : fused ? "one warp per node" : "two kernels");
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
    }
    const int* acc = reinterpret_cast<const int*>(W.h + o_acc);
    int nm = 0;
    for (int q = 0; q < nq; ++q)
        if (acc[q] >= 0) { match12[hq[q].idx1] = acc[q]; match21[acc[q]] = hq[q].idx1; nm++; }
    if (check_orientation) {                                   // :268-285, :628-652
        int bin_of[30] = {0};
        std::vector<int> bin((size_t)nq, -1);
        for (int q = 0; q < nq; ++q)
            if (acc[q] >= 0) {
                bin[q] = rot_bin(s1.angle[(size_t)hq[q].idx1 * s1.angle_stride], s2.angle[(size_t)acc[q] * s2.angle_stride]);
                bin_of[bin[q]]++;
            }
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(bin_of, 30, ind1, ind2, ind3);
        for (int q = 0; q < nq; ++q) {
            if (bin[q] < 0 || bin[q] == ind1 || bin[q] == ind2 || bin[q] == ind3) continue;
            match12[hq[q].idx1] = -1; match21[acc[q]] = -1;
            nm--;
        }
    }
    *nmatches = nm;
    tr.mark("finish");
    return ORBGPU_OK;
}

static int view_of_bow(SideView& v, const orbgpu_bow_side* s)
{
    if (s->n < 0 || (s->n > 0 && (!s->desc || !s->valid || !s->angle))) return ORBGPU_ERR_ARG;
    if (s->n > GRID_MAX_KEYS) return ORBGPU_ERR_UNSUPPORTED;
    const int rc = check_feature_vector(s->n, s->node_ids, s->node_start, s->indices, s->n_nodes);
    if (rc != ORBGPU_OK) return rc;
    v.n = s->n; v.n_nodes = s->n_nodes; v.node_ids = s->node_ids; v.node_start = s->node_start; v.indices = s->indices;
    v.flags = s->valid; v.angle = s->angle; v.angle_stride = 1; v.h_desc = s->desc;
    return ORBGPU_OK;
}

extern "C" int orbgpu_search_by_bow(int device, const orbgpu_bow_side* s1, const orbgpu_bow_side* s2, float nnratio, int check_orientation,
                                    int strict_th_low, int32_t* match12, int32_t* match21, int* nmatches)
{
    if (!s1 || !s2 || !match12 || !match21 || !nmatches) return ORBGPU_ERR_ARG;
    SideView v1, v2;
    int rc;
    if ((rc = view_of_bow(v1, s1)) != ORBGPU_OK || (rc = view_of_bow(v2, s2)) != ORBGPU_OK) return rc;
    return bow_run(device, v1, v2, nnratio, check_orientation, strict_th_low, match12, match21, nmatches);
}

// ---- resident keyframe store ---------------------------------------------------------------------------------------
// The reference hands every search the KeyFrame object itself; behind a C ABI that means re-staging 100-250 KB of
// keypoints / descriptors per call although a keyframe's mvKeysUn, mDescriptors, mvuRight and mFeatVec are written once in
// the constructor / ComputeBoW (src/KeyFrame.cc:32-58, :66-76) and never again.  The store keeps them in HBM; searches
// name keyframes by id and send only what changes between calls (map-point flags, poses).  What the host-side gating
// needs (FeatureVector, angles, mvuRight) is kept in host copies beside the device block.
namespace {

struct KfSlot {
    bool live = false;
    int n = 0, n_nodes = 0, nlevels = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0;
    uint8_t* d = nullptr; size_t bytes = 0;          // keys | desc | u_right | indices | scale factors | level sigma2
    size_t o_keys = 0, o_desc = 0, o_ur = 0, o_idx = 0, o_sc = 0, o_sg = 0;
    std::vector<uint32_t> node_ids, indices;
    std::vector<int32_t> node_start;
    std::vector<float> angle, u_right;
};

}  // namespace

struct orbgpu_kfstore {
    int device = 0;
    std::mutex mu;                                   // slot table and block pool; searches copy what they need under it
    std::vector<std::shared_ptr<KfSlot> > slots;   // searches hold a reference while they run
    std::vector<int> free_ids;
    // device memory comes in 32 MB slabs cut into blocks of 16 KB multiples; erased blocks are reused by size class
    // (no cudaMalloc / cudaFree on the add / erase path once a slab has room)
    static const size_t SLAB = (size_t)32 << 20, QUANTUM = (size_t)16 << 10;
    std::vector<uint8_t*> slabs;
    size_t slab_used = SLAB;
    std::map<size_t, std::vector<uint8_t*> > pool;
    int live = 0;

    uint8_t* block(size_t bytes)
    {
        std::vector<uint8_t*>& fl = pool[bytes];
        if (!fl.empty()) { uint8_t* p = fl.back(); fl.pop_back(); return p; }
        if (bytes > SLAB) return nullptr;
        if (slab_used + bytes > SLAB) {
            uint8_t* s = nullptr;
            if (cudaMalloc(&s, SLAB) != cudaSuccess) { cudaGetLastError(); return nullptr; }
            slabs.push_back(s); slab_used = 0;
        }
        uint8_t* p = slabs.back() + slab_used;
        slab_used += bytes;
        return p;
    }
};

static void side_of_slot(SideView& v, const KfSlot& s, const uint8_t* flags)
{
    v.n = s.n; v.n_nodes = s.n_nodes; v.node_ids = s.node_ids.data(); v.node_start = s.node_start.data(); v.indices = s.indices.data();
    v.flags = flags; v.angle = s.angle.data(); v.angle_stride = 1; v.u_right = s.u_right.data();
    v.d_keys = reinterpret_cast<const float*>(s.d + s.o_keys); v.d_desc = s.d + s.o_desc;
    v.d_ur = reinterpret_cast<const float*>(s.d + s.o_ur); v.d_indices = reinterpret_cast<const uint32_t*>(s.d + s.o_idx);
}

static std::shared_ptr<KfSlot> slot_of(orbgpu_kfstore* st, int id)
{
    if (id < 0 || id >= (int)st->slots.size()) return nullptr;
    return st->slots[(size_t)id];
}

extern "C" {

int orbgpu_kfstore_create(orbgpu_kfstore** out, int device)
{
    if (!out) return ORBGPU_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    CUS(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return ORBGPU_ERR_ARG;
    orbgpu_kfstore* st = new (std::nothrow) orbgpu_kfstore();
    if (!st) return ORBGPU_ERR_NOMEM;
    st->device = device;
    st->slots.reserve(1024);
    *out = st;
    return ORBGPU_OK;
}

void orbgpu_kfstore_destroy(orbgpu_kfstore* st)
{
    if (!st) return;
    cudaSetDevice(st->device);
    for (uint8_t* s : st->slabs) cudaFree(s);
    delete st;
}

int orbgpu_kfstore_size(orbgpu_kfstore* st)
{
    if (!st) return ORBGPU_ERR_ARG;
    std::lock_guard<std::mutex> g(st->mu);
    return st->live;
}

int orbgpu_kf_add(orbgpu_kfstore* st, const orbgpu_kf_data* k, int32_t* id)
{
    if (!st || !k || !id) return ORBGPU_ERR_ARG;
    *id = -1;
    if (k->n < 0 || (k->n > 0 && (!k->keys_un || !k->desc))) return ORBGPU_ERR_ARG;
    if (k->n > GRID_MAX_KEYS) return ORBGPU_ERR_UNSUPPORTED;
    if (k->nlevels < 1 || k->nlevels > ORBGPU_MAX_LEVELS) return ORBGPU_ERR_ARG;
    int rc = check_feature_vector(k->n, k->node_ids, k->node_start, k->indices, k->n_nodes);
    if (rc != ORBGPU_OK) return rc;
    for (int i = 0; i < k->n; ++i)
        if (k->keys_un[i].octave < 0 || k->keys_un[i].octave >= k->nlevels) return ORBGPU_ERR_ARG;
    const int n = k->n, nidx = k->n_nodes > 0 ? k->node_start[k->n_nodes] : 0;
    if (nidx > 0xFFFF) return ORBGPU_ERR_UNSUPPORTED;
    std::shared_ptr<KfSlot> sp = std::make_shared<KfSlot>();
    KfSlot& s = *sp;
    Layout L;
    s.o_keys = L.add(sizeof(orbgpu_keypoint) * (size_t)n); s.o_desc = L.add(32 * (size_t)n); s.o_ur = L.add(4 * (size_t)n);
    s.o_idx = L.add(4 * (size_t)nidx); s.o_sc = L.add(4 * (size_t)ORBGPU_MAX_LEVELS); s.o_sg = L.add(4 * (size_t)ORBGPU_MAX_LEVELS);
    s.bytes = (L.off + orbgpu_kfstore::QUANTUM - 1) / orbgpu_kfstore::QUANTUM * orbgpu_kfstore::QUANTUM;
    s.n = n; s.n_nodes = k->n_nodes; s.nlevels = k->nlevels; s.fx = k->fx; s.fy = k->fy; s.cx = k->cx; s.cy = k->cy;
    s.node_ids.assign(k->node_ids, k->node_ids + k->n_nodes);
    s.node_start.assign(k->node_start, k->node_start + (k->n_nodes > 0 ? k->n_nodes + 1 : 0));
    if (k->n_nodes == 0) s.node_start.assign(1, 0);
    s.indices.assign(k->indices, k->indices + nidx);
    s.angle.resize((size_t)n); s.u_right.resize((size_t)n);
    for (int i = 0; i < n; ++i) { s.angle[(size_t)i] = k->keys_un[i].angle; s.u_right[(size_t)i] = k->u_right ? k->u_right[i] : -1.f; }
    // stage the block in the calling thread's pinned mirror, one upload, wait: the block is complete when the id is handed out
    Workspace& W = g_ws;
    if ((rc = W.reserve(st->device, 0, L.off)) != ORBGPU_OK) return rc;
    memset(W.h, 0, L.off);
    if (n > 0) {
        memcpy(W.h + s.o_keys, k->keys_un, sizeof(orbgpu_keypoint) * (size_t)n);
        memcpy(W.h + s.o_desc, k->desc, 32 * (size_t)n);
        memcpy(W.h + s.o_ur, s.u_right.data(), 4 * (size_t)n);
    }
    if (nidx > 0) memcpy(W.h + s.o_idx, s.indices.data(), 4 * (size_t)nidx);
    if (k->scale_factors) memcpy(W.h + s.o_sc, k->scale_factors, 4 * (size_t)k->nlevels);     // (the block was zeroed above)
    if (k->level_sigma2) memcpy(W.h + s.o_sg, k->level_sigma2, 4 * (size_t)k->nlevels);
    {
        std::lock_guard<std::mutex> g(st->mu);
        s.d = st->block(s.bytes);
    }
    if (!s.d) return ORBGPU_ERR_NOMEM;
    cudaError_t ce = cudaMemcpyAsync(s.d, W.h, L.off, cudaMemcpyHostToDevice, W.st);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(W.st);
    std::lock_guard<std::mutex> g(st->mu);
    if (ce != cudaSuccess) {
        st->pool[s.bytes].push_back(s.d);
        orbgpu_internal_set_error(cudaGetErrorString(ce));
        return ORBGPU_ERR_CUDA;
    }
    s.live = true;
    int slot;
    if (!st->free_ids.empty()) { slot = st->free_ids.back(); st->free_ids.pop_back(); st->slots[(size_t)slot] = sp; }
    else { slot = (int)st->slots.size(); st->slots.push_back(sp); }
    st->live++;
    *id = slot;
    return ORBGPU_OK;
}

int orbgpu_kf_erase(orbgpu_kfstore* st, int32_t id)
{
    if (!st) return ORBGPU_ERR_ARG;
    std::lock_guard<std::mutex> g(st->mu);
    std::shared_ptr<KfSlot> s = slot_of(st, id);
    if (!s) return ORBGPU_ERR_STATE;
    st->pool[s->bytes].push_back(s->d);
    st->slots[(size_t)id].reset();
    st->free_ids.push_back(id);
    st->live--;
    return ORBGPU_OK;
}

// A search takes a reference to the slots it names under the lock, so orbgpu_kf_add / orbgpu_kf_erase from other threads
// cannot move or free their host tables under it.  Erasing a keyframe while a search that names it is in flight is still
// the caller's race (its device block may be handed to the next keyframe), as deleting the KeyFrame would be in the reference.
int orbgpu_kf_search_by_bow(orbgpu_kfstore* st, int32_t id1, const uint8_t* valid1, int32_t id2, const uint8_t* valid2,
                            const orbgpu_bow_side* frame2, float nnratio, int check_orientation, int strict_th_low,
                            int32_t* match12, int32_t* match21, int* nmatches)
{
    if (!st || !valid1 || !match12 || !match21 || !nmatches) return ORBGPU_ERR_ARG;
    if ((id2 >= 0) == (frame2 != nullptr)) return ORBGPU_ERR_ARG;          // side 2: a stored keyframe OR a Frame's arrays
    if (id2 >= 0 && !valid2) return ORBGPU_ERR_ARG;
    std::shared_ptr<KfSlot> k1, k2;
    {
        std::lock_guard<std::mutex> g(st->mu);
        k1 = slot_of(st, id1);
        if (id2 >= 0) k2 = slot_of(st, id2);
        if (!k1 || (id2 >= 0 && !k2)) return ORBGPU_ERR_STATE;
    }
    SideView v1, v2;
    side_of_slot(v1, *k1, valid1);
    if (id2 >= 0) side_of_slot(v2, *k2, valid2);
    else { const int rc = view_of_bow(v2, frame2); if (rc != ORBGPU_OK) return rc; }
    return bow_run(st->device, v1, v2, nnratio, check_orientation, strict_th_low, match12, match21, nmatches);
}

int orbgpu_kf_search_for_triangulation_batch(orbgpu_kfstore* st, const orbgpu_kf_triangulation_job* jobs_in, int n_jobs,
                                             int32_t* matches12, int* nmatches)
{
    if (!st || n_jobs < 0) return ORBGPU_ERR_ARG;
    if (n_jobs == 0) return ORBGPU_OK;
    if (!jobs_in || !matches12 || !nmatches) return ORBGPU_ERR_ARG;
    std::vector<TriJob> jobs((size_t)n_jobs);
    std::vector<std::shared_ptr<KfSlot> > k((size_t)2 * n_jobs);
    std::vector<int> nlevels((size_t)n_jobs);
    {
        std::lock_guard<std::mutex> g(st->mu);
        for (int p = 0; p < n_jobs; ++p) {
            const orbgpu_kf_triangulation_job& j = jobs_in[p];
            if (!j.has_map_point1 || !j.has_map_point2) return ORBGPU_ERR_ARG;
            k[(size_t)2 * p] = slot_of(st, j.id1); k[(size_t)2 * p + 1] = slot_of(st, j.id2);
            if (!k[(size_t)2 * p] || !k[(size_t)2 * p + 1]) return ORBGPU_ERR_STATE;
        }
    }
    for (int p = 0; p < n_jobs; ++p) {
        const orbgpu_kf_triangulation_job& j = jobs_in[p];
        const KfSlot &k1 = *k[(size_t)2 * p], &k2 = *k[(size_t)2 * p + 1];
        TriProblem& P = jobs[(size_t)p].P;
        side_of_slot(P.s1, k1, j.has_map_point1); side_of_slot(P.s2, k2, j.has_map_point2);
        P.F12 = j.F12; P.Cw1 = j.Cw1; P.T2w = j.T2w; P.fx = k2.fx; P.fy = k2.fy; P.cx = k2.cx; P.cy = k2.cy;
        P.d_scale2 = reinterpret_cast<const float*>(k2.d + k2.o_sc); P.d_sigma2 = reinterpret_cast<const float*>(k2.d + k2.o_sg);
        P.only_stereo = j.only_stereo; P.check_orientation = j.check_orientation;
        nlevels[(size_t)p] = k2.nlevels;
    }
    return tri_run(st->device, jobs, nlevels.data(), matches12, nmatches);
}

int orbgpu_kf_search_for_triangulation(orbgpu_kfstore* st, const orbgpu_kf_triangulation_job* job, int32_t* matches12, int* nmatches)
{
    return orbgpu_kf_search_for_triangulation_batch(st, job, job ? 1 : 0, matches12, nmatches);
}

}  // extern "C"
