// orb_search.cu -- the grid-gated searches between two frames:
//   Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea     src/Frame.cc:230-245, 327-391
//   ORBmatcher::SearchForInitialization                             src/ORBmatcher.cc:405-520
//   ORBmatcher::SearchByProjection(Frame&, const Frame&, th, mono)  src/ORBmatcher.cc:1328-1470
//
// Device layout.  The 64 x 48 grid of the searched frame is a counting sort of its keypoint indices by cell id
// ix * 48 + iy with the index order kept inside a cell (= the push_back order of AssignFeaturesToGrid); the cells
// (ix, iyMin..iyMax) of one window column are then ONE contiguous range, so a window is nMaxCellX - nMinCellX + 1
// ranges visited in the reference's order.  Three steps per call:
//   k_grid_build        one CTA: cell ids, per-cell counts, exclusive scan, in-cell rank by index
//   k_area_candidates   one warp per query: walks the window ranges, keeps order with ballot prefix sums, evaluates the
//                       Hamming distance of every candidate and stores (index, distance) lists
//   k_init_resolve / k_proj_resolve   one warp walks the queries in index order and applies the reference's
//                       order-dependent rules (vMatchedDistance / vnMatches21 stealing; mvpMapPoints occupancy) --
//                       a true sequential dependence in the reference, kept as such
// The rotation histogram + ComputeThreeMaxima (src/ORBmatcher.cc:1601-1642) finish on the host, O(matches).
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
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

__global__ void __launch_bounds__(1024) k_grid_build(GridDev G)
{
    __shared__ int s_cnt[GRID_CELLS];
    __shared__ short s_cell[GRID_MAX_KEYS];
    __shared__ int s_part[1024];
    const int t = threadIdx.x;
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
    // exclusive scan of 3072 counts: 3 per thread
    const int c0 = s_cnt[3 * t], c1 = s_cnt[3 * t + 1], c2 = s_cnt[3 * t + 2];
    s_part[t] = c0 + c1 + c2;
    __syncthreads();
    for (int s = 1; s < 1024; s <<= 1) {
        const int v = t >= s ? s_part[t - s] : 0;
        __syncthreads();
        s_part[t] += v;
        __syncthreads();
    }
    const int base = s_part[t] - (c0 + c1 + c2);
    s_cnt[3 * t] = base; s_cnt[3 * t + 1] = base + c0; s_cnt[3 * t + 2] = base + c0 + c1;
    G.cell_start[3 * t] = base; G.cell_start[3 * t + 1] = base + c0; G.cell_start[3 * t + 2] = base + c0 + c1;
    if (t == 1023) G.cell_start[GRID_CELLS] = s_part[1023];
    __syncthreads();
    // position inside the cell = number of earlier keypoints of the same cell (keeps the push_back order)
    for (int i = t; i < G.n; i += 1024) {
        const int c = s_cell[i];
        if (c < 0) continue;
        int rank = 0;
        for (int j = 0; j < i; ++j) rank += s_cell[j] == c;
        G.sorted[s_cnt[c] + rank] = i;
    }
}

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
        const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, G.minX), r), G.invW)));
        const int nMaxCellX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, G.minX), r), G.invW)));
        const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, G.minY), r), G.invH)));
        const int nMaxCellY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, G.minY), r), G.invH)));
        if (nMinCellX < GRID_COLS && nMaxCellX >= 0 && nMinCellY < GRID_ROWS && nMaxCellY >= 0) {
            const bool bCheckLevels = (A.min_level > 0) || (A.max_level >= 0);
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(qdesc) + 2 * (size_t)q);
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(qdesc) + 2 * (size_t)q + 1);
            uint32_t* out = cand + (size_t)q * cap;
            for (int ix = nMinCellX; ix <= nMaxCellX; ++ix) {
                if (nMinCellY > nMaxCellY) break;
                const int p0 = G.cell_start[ix * GRID_ROWS + nMinCellY], p1 = G.cell_start[ix * GRID_ROWS + nMaxCellY + 1];
                for (int pb = p0; pb < p1; pb += 32) {
                    const int p = pb + lane;
                    bool ok = p < p1;
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

// two smallest keys of a warp (keys are distinct: the candidate position is part of the key)
__device__ __forceinline__ void warp_two_min(unsigned& k1, unsigned& k2)
{
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        const unsigned o1 = __shfl_xor_sync(0xffffffffu, k1, s), o2 = __shfl_xor_sync(0xffffffffu, k2, s);
        const unsigned lo = min(k1, o1), hi = max(k1, o1);
        k2 = min(hi, min(k2, o2));
        k1 = lo;
    }
}

// SearchForInitialization, the loop over i1 (src/ORBmatcher.cc:418-484) on precomputed candidate lists.
// accepted[i1] = bestIdx2 at the moment i1 was accepted (feeds rotHist), else -1.
__global__ void __launch_bounds__(32) k_init_resolve(const uint32_t* __restrict__ cand, const int* __restrict__ cand_n, int cap, int n1,
                                                     int n2, float nnratio, int* matched_dist, int* matches21, int* matches12,
                                                     int* accepted, int* nmatches_out)
{
    const int lane = threadIdx.x;
    for (int i = lane; i < n2; i += 32) { matched_dist[i] = INT_MAX; matches21[i] = -1; }
    for (int i = lane; i < n1; i += 32) { matches12[i] = -1; accepted[i] = -1; }
    __syncwarp();
    int nmatches = 0;
    for (int i1 = 0; i1 < n1; ++i1) {
        const int cnt = cand_n[i1];
        if (cnt == 0) continue;
        const uint32_t* c = cand + (size_t)i1 * cap;
        unsigned k1 = 0xFFFFFFFFu, k2 = 0xFFFFFFFFu;           // (dist << 16) | position: earliest candidate wins ties
        for (int p = lane; p < cnt; p += 32) {
            const uint32_t e = c[p];
            const int i2 = (int)(e >> CAND_DIST_BITS), dist = (int)(e & ((1u << CAND_DIST_BITS) - 1u));
            if (matched_dist[i2] <= dist) continue;            // :444-445
            const unsigned key = ((unsigned)dist << 16) | (unsigned)p;
            if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
        }
        warp_two_min(k1, k2);
        if (lane == 0 && k1 != 0xFFFFFFFFu) {
            const int bestDist = (int)(k1 >> 16);
            const int bestDist2 = k2 == 0xFFFFFFFFu ? INT_MAX : (int)(k2 >> 16);
            if (bestDist <= 50 && (float)bestDist < __fmul_rn((float)bestDist2, nnratio)) {      // TH_LOW, :459-461
                const int bestIdx2 = (int)(c[k1 & 0xFFFFu] >> CAND_DIST_BITS);
                if (matches21[bestIdx2] >= 0) { matches12[matches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                matches21[bestIdx2] = i1;
                matched_dist[bestIdx2] = bestDist;
                accepted[i1] = bestIdx2;
                nmatches++;
            }
        }
        __syncwarp();
    }
    if (lane == 0) *nmatches_out = nmatches;
}

// SearchByProjection(Frame, Frame): projection of the last frame's map points (src/ORBmatcher.cc:1361-1391)
struct ProjCam { float R[9], t[3], fx, fy, cx, cy, bf, minX, maxX, minY, maxY, th; int forward, backward; };

__global__ void k_proj_queries(ProjCam C, const float* __restrict__ world_pos, const uint8_t* __restrict__ valid,
                               const int* __restrict__ last_octave, const float* __restrict__ scale, int n, AreaQuery* __restrict__ Q)
{
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

// the loop over the last frame's points (src/ORBmatcher.cc:1353-1443) on precomputed candidate lists
__global__ void __launch_bounds__(32) k_proj_resolve(const uint32_t* __restrict__ cand, const int* __restrict__ cand_n, int cap, int n_last,
                                                     int n_cur, const uint8_t* __restrict__ observed, uint8_t* occupied, int* cur_match,
                                                     int* accepted, int* nmatches_out)
{
    const int lane = threadIdx.x;
    for (int i = lane; i < n_cur; i += 32) cur_match[i] = -1;
    for (int i = lane; i < n_last; i += 32) accepted[i] = -1;
    __syncwarp();
    int nmatches = 0;
    for (int i = 0; i < n_last; ++i) {
        const int cnt = cand_n[i];
        if (cnt == 0) continue;
        const uint32_t* c = cand + (size_t)i * cap;
        unsigned best = 0xFFFFFFFFu;
        for (int p = lane; p < cnt; p += 32) {
            const uint32_t e = c[p];
            if (occupied[e >> CAND_DIST_BITS]) continue;       // :1403-1405
            best = min(best, ((e & ((1u << CAND_DIST_BITS) - 1u)) << 16) | (unsigned)p);
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
        if (lane == 0 && best != 0xFFFFFFFFu && (int)(best >> 16) <= 100) {     // bestDist < 256 by construction; TH_HIGH
            const int bestIdx2 = (int)(c[best & 0xFFFFu] >> CAND_DIST_BITS);
            cur_match[bestIdx2] = i;
            occupied[bestIdx2] = observed[i];
            accepted[i] = bestIdx2;
            nmatches++;
        }
        __syncwarp();
    }
    if (lane == 0) *nmatches_out = nmatches;
}

// ---- host side ---------------------------------------------------------------------------------------------------
namespace {

struct DevBuf {                   // every device allocation of one call, freed together
    std::vector<void*> ptrs;
    ~DevBuf() { for (void* p : ptrs) cudaFree(p); }
    template <typename T> T* get(size_t count)
    {
        void* p = nullptr;
        if (cudaMalloc(&p, sizeof(T) * (count ? count : 1)) != cudaSuccess) return nullptr;
        ptrs.push_back(p);
        return static_cast<T*>(p);
    }
};

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

// uploads the searched frame (mvKeysUn + mDescriptors) and builds its grid
int make_grid(DevBuf& B, const orbgpu_keypoint* keys, const uint8_t* desc, int n, const float* bounds, GridDev& G)
{
    float* d_keys = B.get<float>(7 * (size_t)n);
    uint8_t* d_desc = B.get<uint8_t>(32 * (size_t)n);
    int* d_start = B.get<int>(GRID_CELLS + 1);
    int* d_sorted = B.get<int>((size_t)n);
    if (!d_keys || !d_desc || !d_start || !d_sorted) return ORBGPU_ERR_NOMEM;
    CUS(cudaMemcpy(d_keys, keys, sizeof(orbgpu_keypoint) * (size_t)n, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_desc, desc, 32 * (size_t)n, cudaMemcpyHostToDevice));
    G.keys = d_keys; G.desc = d_desc; G.n = n;
    G.minX = bounds[0]; G.minY = bounds[2];
    G.invW = static_cast<float>(GRID_COLS) / static_cast<float>(bounds[1] - bounds[0]);       // src/Frame.cc:155-156
    G.invH = static_cast<float>(GRID_ROWS) / static_cast<float>(bounds[3] - bounds[2]);
    G.cell_start = d_start; G.sorted = d_sorted;
    k_grid_build<<<1, 1024>>>(G);
    CUS(cudaGetLastError());
    return ORBGPU_OK;
}

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

}  // namespace

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
    CUS(cudaSetDevice(device));
    DevBuf B;
    GridDev G;
    int rc = make_grid(B, keys2_un, desc2, n2, bounds4, G);
    if (rc != ORBGPU_OK) return rc;
    // queries: level-0 keypoints of F1 around vbPrevMatched (:418-425)
    std::vector<AreaQuery> hq((size_t)n1);
    for (int i = 0; i < n1; ++i) {
        AreaQuery& q = hq[(size_t)i];
        q.x = prev_matched[2 * i]; q.y = prev_matched[2 * i + 1]; q.r = (float)window_size;
        q.min_level = q.max_level = keys1_un[i].octave;
        q.active = keys1_un[i].octave > 0 ? 0 : 1;
        q.ur = 0.f; q.check_ur = 0;
    }
    AreaQuery* d_q = B.get<AreaQuery>((size_t)n1);
    uint8_t* d_qdesc = B.get<uint8_t>(32 * (size_t)n1);
    uint32_t* d_cand = B.get<uint32_t>((size_t)n1 * n2);
    int* d_ints = B.get<int>((size_t)n1 * 3 + (size_t)n2 * 2 + 1);
    if (!d_q || !d_qdesc || !d_cand || !d_ints) return ORBGPU_ERR_NOMEM;
    int *d_cn = d_ints, *d_m12 = d_cn + n1, *d_acc = d_m12 + n1, *d_md = d_acc + n1, *d_m21 = d_md + n2, *d_nm = d_m21 + n2;
    CUS(cudaMemcpy(d_q, hq.data(), sizeof(AreaQuery) * (size_t)n1, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_qdesc, desc1, 32 * (size_t)n1, cudaMemcpyHostToDevice));
    k_area_candidates<<<(n1 + 7) / 8, 256>>>(G, d_q, d_qdesc, n1, nullptr, d_cand, n2, d_cn);
    k_init_resolve<<<1, 32>>>(d_cand, d_cn, n2, n1, n2, nnratio, d_md, d_m21, d_m12, d_acc, d_nm);
    CUS(cudaGetLastError());
    std::vector<int> acc((size_t)n1);
    int nm = 0;
    CUS(cudaMemcpy(matches12, d_m12, sizeof(int) * (size_t)n1, cudaMemcpyDeviceToHost));
    CUS(cudaMemcpy(acc.data(), d_acc, sizeof(int) * (size_t)n1, cudaMemcpyDeviceToHost));
    CUS(cudaMemcpy(&nm, d_nm, sizeof(int), cudaMemcpyDeviceToHost));
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
    for (int i = 0; i < n1; ++i)
        if (a->mp_valid[i] && (a->last_octave[i] < 0 || a->last_octave[i] >= a->nlevels)) return ORBGPU_ERR_ARG;
    *nmatches = 0;
    for (int i = 0; i < n2; ++i) cur_match[i] = -1;
    if (n1 == 0 || n2 == 0) return ORBGPU_OK;
    CUS(cudaSetDevice(device));
    DevBuf B;
    GridDev G;
    int rc = make_grid(B, a->cur_keys_un, a->cur_desc, n2, a->bounds, G);
    if (rc != ORBGPU_OK) return rc;
    // camera motion between the frames (:1339-1350); the cv::Mat expressions restated with cv::gemm's arithmetic
    ProjCam C;
    const float* Tc = a->Tcw_cur; const float* Tl = a->Tcw_last;
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) C.R[3 * r + c] = Tc[4 * r + c]; C.t[r] = Tc[4 * r + 3]; }
    float twc[3], tlc[3];
    for (int i = 0; i < 3; ++i) {                              // -Rcw.t() * tcw: generic gemm path, double accumulation
        double s = 0;
        for (int k = 0; k < 3; ++k) s += (double)Tc[4 * k + i] * (double)C.t[k];
        twc[i] = (float)(s * -1.0);
    }
    for (int i = 0; i < 3; ++i) {                              // Rlw * twc + tlw: 3x3 float case
        const float t = Tl[4 * i] * twc[0] + Tl[4 * i + 1] * twc[1] + Tl[4 * i + 2] * twc[2];
        tlc[i] = (float)((double)t * 1.0 + (double)Tl[4 * i + 3] * 1.0);
    }
    C.forward = (tlc[2] > a->b && !a->mono) ? 1 : 0;
    C.backward = (-tlc[2] > a->b && !a->mono) ? 1 : 0;
    C.fx = a->fx; C.fy = a->fy; C.cx = a->cx; C.cy = a->cy; C.bf = a->bf; C.th = a->th;
    C.minX = a->bounds[0]; C.maxX = a->bounds[1]; C.minY = a->bounds[2]; C.maxY = a->bounds[3];

    float* d_pos = B.get<float>(3 * (size_t)n1);
    uint8_t* d_qdesc = B.get<uint8_t>(32 * (size_t)n1);
    uint8_t* d_bytes = B.get<uint8_t>(2 * (size_t)n1 + (size_t)n2);
    int* d_oct = B.get<int>((size_t)n1);
    float* d_scale = B.get<float>((size_t)a->nlevels);
    float* d_ur = B.get<float>((size_t)n2);
    AreaQuery* d_q = B.get<AreaQuery>((size_t)n1);
    uint32_t* d_cand = B.get<uint32_t>((size_t)n1 * n2);
    int* d_ints = B.get<int>((size_t)n1 * 2 + (size_t)n2 + 1);
    if (!d_pos || !d_qdesc || !d_bytes || !d_oct || !d_scale || !d_ur || !d_q || !d_cand || !d_ints) return ORBGPU_ERR_NOMEM;
    uint8_t *d_valid = d_bytes, *d_obs = d_bytes + n1, *d_occ = d_obs + n1;
    int *d_cn = d_ints, *d_acc = d_cn + n1, *d_cm = d_acc + n1, *d_nm = d_cm + n2;
    CUS(cudaMemcpy(d_pos, a->world_pos, 12 * (size_t)n1, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_qdesc, a->mp_desc, 32 * (size_t)n1, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_valid, a->mp_valid, (size_t)n1, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_obs, a->mp_observed, (size_t)n1, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_occ, a->cur_occupied, (size_t)n2, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_oct, a->last_octave, 4 * (size_t)n1, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_scale, a->scale_factors, 4 * (size_t)a->nlevels, cudaMemcpyHostToDevice));
    CUS(cudaMemcpy(d_ur, a->cur_u_right, 4 * (size_t)n2, cudaMemcpyHostToDevice));
    k_proj_queries<<<(n1 + 127) / 128, 128>>>(C, d_pos, d_valid, d_oct, d_scale, n1, d_q);
    k_area_candidates<<<(n1 + 7) / 8, 256>>>(G, d_q, d_qdesc, n1, d_ur, d_cand, n2, d_cn);
    k_proj_resolve<<<1, 32>>>(d_cand, d_cn, n2, n1, n2, d_obs, d_occ, d_cm, d_acc, d_nm);
    CUS(cudaGetLastError());
    std::vector<int> acc((size_t)n1);
    int nm = 0;
    CUS(cudaMemcpy(cur_match, d_cm, sizeof(int) * (size_t)n2, cudaMemcpyDeviceToHost));
    CUS(cudaMemcpy(acc.data(), d_acc, sizeof(int) * (size_t)n1, cudaMemcpyDeviceToHost));
    CUS(cudaMemcpy(&nm, d_nm, sizeof(int), cudaMemcpyDeviceToHost));
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

}  // extern "C"
