// orb_vocab.cu -- the DBoW2 vocabulary tree on the device: Frame::ComputeBoW / KeyFrame::ComputeBoW
// (src/Frame.cc:395-402, src/KeyFrame.cc:59-68) and Vocabulary::score for batches of keyframes.
//
// Reference: Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1194 (transform with FeatureVector), :1218-1261 (the walk
// of one descriptor: strict '<' on FORB::distance, so the first child wins ties), BowVector.cpp:34-86 (addWeight,
// addIfNotExist, normalize), FeatureVector.cpp:29-43, ScoringObject.cpp:25-72 (L1 score).
//
// Layout in HBM: the nodes are renumbered breadth first so that the children of a node are one contiguous block
// ("position" order): desc[pos] 32 B, child[pos] = (first child position, child count), node_of[pos] = the reference's
// node id, word_of[pos], weight_of[pos].  ORBvoc.txt (k = 10, L = 6: 1.1 M nodes) takes 62 MB and stays in L2 once warm.
//
// Kernels (byte / integer work, no tensor cores):
//   k_bow_descend   warp = one descriptor; lane = one child of the current node: one coalesced read of the child block,
//                   256-bit XOR + popcount, warp minimum over (distance << 20 | child index)
//   k_bow_finalize  one CTA: bitonic sort of (word << 32 | feature) and (node << 32 | feature) in shared memory, run heads
//                   by prefix scan; a word's value is the leaf weight added once per feature in double (as addWeight
//                   does); the norm is one thread's in-order sum over the words (the reference's summation order)
//   k_bow_score     warp = one keyframe: lanes binary-search the query for their words; matching terms are added by lane
//                   order through shuffles, i.e. in ascending word order as the merge walk of the reference does
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <new>
#include <string>
#include <vector>

#include "../../include/orbgpu.h"

void orbgpu_internal_set_error(const char* s);   // orbgpu_extractor.cu
#define CUV(call)                                                                                     \
    do {                                                                                              \
        cudaError_t _e = (call);                                                                      \
        if (_e != cudaSuccess) {                                                                      \
            char _b[256];                                                                             \
            snprintf(_b, sizeof(_b), "%s: %s", #call, cudaGetErrorString(_e));                        \
            orbgpu_internal_set_error(_b);                                                            \
            return ORBGPU_ERR_CUDA;                                                                   \
        }                                                                                             \
    } while (0)

#define BOW_MAX_FEATURES 16384
#define BOW_FIN_THREADS 1024
#define BOW_NO_KEY 0xFFFFFFFFFFFFFFFFull

struct VocabDev {
    const uint4* desc;        // [n_nodes][2]
    const int2* child;        // [n_nodes] (first child position, count)
    const uint32_t* node_of;  // [n_nodes]
    const uint32_t* word_of;  // [n_nodes]
    const double* weight_of;  // [n_nodes]
};

// ------------------------------------------------------------------------------------------------ the walk of one descriptor
__global__ void __launch_bounds__(256) k_bow_descend(VocabDev V, const uint4* __restrict__ feat, int n, int nid_level,
                                                     uint32_t* __restrict__ f_word, uint32_t* __restrict__ f_node,
                                                     double* __restrict__ f_weight)
{
    const int lane = threadIdx.x & 31;
    const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (f >= n) return;
    const uint4 q0 = __ldg(feat + 2 * (size_t)f), q1 = __ldg(feat + 2 * (size_t)f + 1);
    int pos = 0, level = 0;
    bool have_nid = nid_level <= 0;
    uint32_t nid = 0;                                        // root (:1225)
    for (;;) {
        const int2 ch = __ldg(V.child + pos);
        if (ch.y == 0) break;
        ++level;
        unsigned best = 0xFFFFFFFFu;
        for (int c0 = 0; c0 < ch.y; c0 += 32) {
            const int c = c0 + lane;
            if (c < ch.y) {
                const uint4 d0 = __ldg(V.desc + 2 * (size_t)(ch.x + c)), d1 = __ldg(V.desc + 2 * (size_t)(ch.x + c) + 1);
                const int d = __popc(q0.x ^ d0.x) + __popc(q0.y ^ d0.y) + __popc(q0.z ^ d0.z) + __popc(q0.w ^ d0.w) +
                              __popc(q1.x ^ d1.x) + __popc(q1.y ^ d1.y) + __popc(q1.z ^ d1.z) + __popc(q1.w ^ d1.w);
                best = min(best, ((unsigned)d << 20) | (unsigned)c);      // children per node < 2^20
            }
        }
        best = __reduce_min_sync(0xffffffffu, best);
        pos = ch.x + (int)(best & 0xFFFFFu);
        if (level == nid_level) { nid = __ldg(V.node_of + pos); have_nid = true; }
    }
    if (lane == 0) {
        f_word[f] = __ldg(V.word_of + pos);
        f_node[f] = have_nid ? nid : __ldg(V.node_of + pos);
        f_weight[f] = __ldg(V.weight_of + pos);
    }
}

// ------------------------------------------------------------------------------------- BowVector / FeatureVector of one frame
// Bitonic sort of n_pad (power of two >= 64) keys in shared memory.  Pair t of a stage with distance j is
// (i, i | j), i = t with a zero inserted at bit log2(j): for j <= 32 the 32 pairs of a warp are exactly its own block of 64
// consecutive keys, so runs of such stages need only __syncwarp; a CTA barrier is needed around the stages with j >= 64.
__device__ __forceinline__ void bitonic_stage(unsigned long long* keys, int n_pad, int k, int j)
{
    for (int t = threadIdx.x; t < (n_pad >> 1); t += blockDim.x) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const unsigned long long a = keys[i], b = keys[i | j];
        if ((a > b) == ((i & k) == 0)) { keys[i] = b; keys[i | j] = a; }
    }
}

__device__ __forceinline__ void bitonic_sort_u64(unsigned long long* keys, int n_pad)
{
    for (int k = 2; k <= n_pad; k <<= 1) {
        int j = k >> 1;
        for (; j >= 64; j >>= 1) {
            bitonic_stage(keys, n_pad, k, j);
            __syncthreads();
        }
        for (; j > 0; j >>= 1) {
            bitonic_stage(keys, n_pad, k, j);
            __syncwarp();
        }
        if (k >= 64) __syncthreads();       // the next k starts with a stage that crosses warps (k <= 32: still warp-local)
    }
    __syncthreads();
}

// exclusive prefix sum of one int per thread over the CTA (blockDim.x = 1024); *total = the sum
__device__ __forceinline__ int block_exclusive_scan(int v, int* s_warp, int* total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();                                     // s_warp may still be read from a previous scan
    if (lane == 31) s_warp[w] = inc;
    __syncthreads();
    if (w == 0) {
        int x = s_warp[lane], y = x;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, y, o);
            if (lane >= o) y += t;
        }
        s_warp[lane] = y - x;                            // exclusive warp offsets
        if (lane == 31) *total = y;
    }
    __syncthreads();
    return s_warp[w] + inc - v;
}

// norm_mode: 0 none, 1 L1, 2 L2 (GeneralScoring::mustNormalize of the vocabulary's scoring type).
// CTA = one frame: its descriptors are [frame_start[f], frame_start[f + 1]) of the flat per-descriptor arrays; its outputs go
// to the same offsets of the flat output arrays (a frame with n descriptors has at most n words / nodes / indices), its
// `start` list to out_start + frame_start[f] + f (n + 1 entries) and its three counts to out_counts + 3 f.
__global__ void __launch_bounds__(BOW_FIN_THREADS) k_bow_finalize(const int32_t* __restrict__ frame_start,
                                                                  const uint32_t* __restrict__ f_word_all,
                                                                  const uint32_t* __restrict__ f_node_all,
                                                                  const double* __restrict__ f_weight_all, int weighting, int norm_mode,
                                                                  uint32_t* out_words_all, double* out_values_all, uint32_t* out_node_ids_all,
                                                                  int32_t* out_start_all, uint32_t* out_idx_all, int* out_counts_all)
{
    extern __shared__ unsigned long long keys[];
    __shared__ int s_warp[32];
    __shared__ int s_total;
    __shared__ double s_norm;
    const int frame = blockIdx.x;
    const int base = frame_start[frame], n = frame_start[frame + 1] - base;
    const uint32_t* f_word = f_word_all + base; const uint32_t* f_node = f_node_all + base; const double* f_weight = f_weight_all + base;
    uint32_t* out_words = out_words_all + base; double* out_values = out_values_all + base;
    uint32_t* out_node_ids = out_node_ids_all + base; uint32_t* out_idx = out_idx_all + base;
    int32_t* out_start = out_start_all + base + frame;
    int* out_counts = out_counts_all + 3 * frame;
    if (n == 0) {
        if (threadIdx.x == 0) { out_start[0] = 0; out_counts[0] = out_counts[1] = out_counts[2] = 0; }
        return;
    }
    int n_pad = 64;
    while (n_pad < n) n_pad <<= 1;
    const int C = n_pad > BOW_FIN_THREADS ? n_pad / BOW_FIN_THREADS : 1;
    const int i0 = threadIdx.x * C, i1 = min(i0 + C, n_pad);
    const bool accumulate = weighting == 0 || weighting == 1;          // TF_IDF, TF: addWeight; IDF, BINARY: addIfNotExist

    // ---- BowVector: words ascending, value = weight added once per feature of the word (:1157-1159)
    for (int i = threadIdx.x; i < n_pad; i += blockDim.x)
        keys[i] = (i < n && f_weight[i] > 0.0) ? ((unsigned long long)f_word[i] << 32) | (unsigned)i : BOW_NO_KEY;
    __syncthreads();
    bitonic_sort_u64(keys, n_pad);
    int heads = 0, valid = 0;
    for (int i = i0; i < i1; ++i) {
        heads += keys[i] != BOW_NO_KEY && (i == 0 || (keys[i - 1] >> 32) != (keys[i] >> 32));
        valid += keys[i] != BOW_NO_KEY;
    }
    int r = block_exclusive_scan(heads, s_warp, &s_total);
    const int nw = s_total;
    block_exclusive_scan(valid, s_warp, &s_total);
    const int m = s_total;                                             // features that are not stopped
    // where each run starts (out_start serves as scratch here; the FeatureVector pass rewrites it): run length = next - this
    for (int i = i0; i < i1; ++i)
        if (keys[i] != BOW_NO_KEY && (i == 0 || (keys[i - 1] >> 32) != (keys[i] >> 32))) out_start[r++] = i;
    if (threadIdx.x == 0) out_start[nw] = m;
    __syncthreads();
    for (int j = threadIdx.x; j < nw; j += blockDim.x) {
        const int i = out_start[j];
        const unsigned long long key = keys[i];
        const double w = f_weight[(uint32_t)key];
        double v = w;
        if (accumulate) {
            const int cnt = out_start[j + 1] - i;
            for (int c = 1; c < cnt; ++c) v = __dadd_rn(v, w);         // addWeight: "+= w" per further feature, in double
        }
        out_words[j] = (uint32_t)(key >> 32);
        out_values[j] = v;
    }
    __syncthreads();
    double* vals = reinterpret_cast<double*>(keys);                    // the sorted keys are not needed any more
    for (int i = threadIdx.x; i < nw; i += blockDim.x) vals[i] = out_values[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        double norm = 0.0;
        // one in-order chain of DADDs (the reference's summation order); the terms of 16 words are fetched and prepared
        // ahead of the chain so that only the adds are serial
        if (norm_mode != 0) {
            int i = 0;
            for (; i + 16 <= nw; i += 16) {
                double t[16];
#pragma unroll
                for (int u = 0; u < 16; ++u) t[u] = norm_mode == 1 ? fabs(vals[i + u]) : __dmul_rn(vals[i + u], vals[i + u]);
#pragma unroll
                for (int u = 0; u < 16; ++u) norm = __dadd_rn(norm, t[u]);                    // BowVector.cpp:69-73, :75-79
            }
            for (; i < nw; ++i) norm = __dadd_rn(norm, norm_mode == 1 ? fabs(vals[i]) : __dmul_rn(vals[i], vals[i]));
            if (norm_mode == 2) norm = sqrt(norm);
        }
        s_norm = norm;
    }
    __syncthreads();
    if (norm_mode != 0) {
        const double norm = s_norm;
        if (norm > 0.0)
            for (int i = threadIdx.x; i < nw; i += blockDim.x) out_values[i] = vals[i] / norm;
    } else if (accumulate && nw > 0) {
        const double nd = (double)nw;                                                         // :1163-1169
        for (int i = threadIdx.x; i < nw; i += blockDim.x) out_values[i] = vals[i] / nd;
    }
    __syncthreads();

    // ---- FeatureVector: nodes ascending, the features of a node in ascending index order (push_back order)
    for (int i = threadIdx.x; i < n_pad; i += blockDim.x)
        keys[i] = (i < n && f_weight[i] > 0.0) ? ((unsigned long long)f_node[i] << 32) | (unsigned)i : BOW_NO_KEY;
    __syncthreads();
    bitonic_sort_u64(keys, n_pad);
    heads = 0;
    for (int i = i0; i < i1; ++i)
        heads += keys[i] != BOW_NO_KEY && (i == 0 || (keys[i - 1] >> 32) != (keys[i] >> 32));
    r = block_exclusive_scan(heads, s_warp, &s_total);
    const int nn = s_total;
    for (int i = i0; i < i1; ++i) {
        const unsigned long long key = keys[i];
        if (key == BOW_NO_KEY) continue;
        out_idx[i] = (uint32_t)key;
        if (i == 0 || (keys[i - 1] >> 32) != (key >> 32)) {
            out_node_ids[r] = (uint32_t)(key >> 32);
            out_start[r] = i;
            ++r;
        }
    }
    if (threadIdx.x == 0) {
        out_start[nn] = m;
        out_counts[0] = nw; out_counts[1] = nn; out_counts[2] = m;
    }
}

// ---------------------------------------------------------------------------------------------------- L1 score, batched
__global__ void __launch_bounds__(256) k_bow_score(const uint32_t* __restrict__ qw, const double* __restrict__ qv, int nq,
                                                   const uint32_t* __restrict__ kw, const double* __restrict__ kv,
                                                   const int32_t* __restrict__ kbegin, const int32_t* __restrict__ kend, int nk,
                                                   double* __restrict__ out, int32_t* __restrict__ common, int32_t* __restrict__ first)
{
    const int lane = threadIdx.x & 31;
    const int k = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (k >= nk) return;
    const int b = kbegin[k], e = kend[k];
    double score = 0.0;
    int ncommon = 0, first_pos = nq;
    for (int i0 = b; i0 < e; i0 += 32) {
        const int i = i0 + lane;
        bool hit = false;
        double term = 0.0;
        int lo = 0;
        if (i < e) {
            const uint32_t w = __ldg(kw + i);
            int hi = nq;                                           // first query word >= w
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (__ldg(qw + mid) < w) lo = mid + 1; else hi = mid;
            }
            if (lo < nq && __ldg(qw + lo) == w) {
                const double vi = __ldg(qv + lo), wi = __ldg(kv + i);
                term = __dsub_rn(__dsub_rn(fabs(__dsub_rn(vi, wi)), fabs(vi)), fabs(wi));   // ScoringObject.cpp:44
                hit = true;
            }
        }
        unsigned m = __ballot_sync(0xffffffffu, hit);
        if (m && ncommon == 0) first_pos = __shfl_sync(0xffffffffu, lo, __ffs(m) - 1);   // words ascend: the first hit
        ncommon += __popc(m);
        while (m) {                                                // ascending word order = the reference's merge walk
            const int src = __ffs(m) - 1;
            m &= m - 1;
            score = __dadd_rn(score, __shfl_sync(0xffffffffu, term, src));
        }
    }
    if (lane == 0) {
        out[k] = -score / 2.0;                                     // :67
        if (common) common[k] = ncommon;
        if (first) first[k] = first_pos;
    }
}

// ------------------------------------------------------------------------------------------------------------ host side
struct orbgpu_vocabulary {
    int device = 0;
    int k = 0, L = 0, scoring = 0, weighting = 0, n_nodes = 0, n_words = 0;
    cudaStream_t st = nullptr;
    void* d_tree = nullptr;                 // one allocation: desc | child | node_of | word_of | weight_of
    VocabDev dev{};
    // per-call scratch, grown on demand
    int cap = 0, fcap = 0;                  // descriptors / frames the per-call scratch holds
    uint8_t* d_call = nullptr;              // desc | f_word | f_node | f_weight | outputs
    uint8_t* h_call = nullptr;              // pinned mirror of desc (in) and the outputs (out)
    size_t score_cap = 0;
    uint8_t* d_score = nullptr;
    uint8_t* h_score = nullptr;
};

namespace {

int norm_mode_of(int scoring) { return scoring == 5 ? 0 : (scoring == 1 ? 2 : 1); }   // ScoringObject.h:74-89

size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct CallLayout {                         // offsets inside d_call / h_call for `cap` descriptors in `fcap` frames
    size_t desc, fstart, f_word, f_node, f_weight, words, values, node_ids, start, idx, counts, total;
    CallLayout(int cap, int fcap)
    {
        size_t o = 0;
        fstart = o;   o = align256(o + ((size_t)fcap + 1) * 4);
        desc = o;     o = align256(o + (size_t)cap * 32);
        f_word = o;   o = align256(o + (size_t)cap * 4);
        f_node = o;   o = align256(o + (size_t)cap * 4);
        f_weight = o; o = align256(o + (size_t)cap * 8);
        words = o;    o = align256(o + (size_t)cap * 4);
        values = o;   o = align256(o + (size_t)cap * 8);
        node_ids = o; o = align256(o + (size_t)cap * 4);
        start = o;    o = align256(o + ((size_t)cap + (size_t)fcap) * 4);
        idx = o;      o = align256(o + (size_t)cap * 4);
        counts = o;   o = align256(o + (size_t)fcap * 12);
        total = o;
    }
};

int ensure_call(orbgpu_vocabulary* v, int n, int frames)
{
    if (n <= v->cap && frames <= v->fcap) return ORBGPU_OK;
    int cap = v->cap ? v->cap : 2048, fcap = v->fcap ? v->fcap : 16;
    while (cap < n) cap <<= 1;
    while (fcap < frames) fcap <<= 1;
    if (v->d_call) { cudaFree(v->d_call); v->d_call = nullptr; }
    if (v->h_call) { cudaFreeHost(v->h_call); v->h_call = nullptr; }
    v->cap = v->fcap = 0;
    const CallLayout lay(cap, fcap);
    CUV(cudaMalloc(&v->d_call, lay.total));
    CUV(cudaHostAlloc(&v->h_call, lay.total, cudaHostAllocDefault));
    v->cap = cap; v->fcap = fcap;
    return ORBGPU_OK;
}

}  // namespace

extern "C" {

int orbgpu_vocabulary_create(orbgpu_vocabulary** out, int device, int k, int L, int scoring, int weighting, int n_nodes,
                             const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc, const double* weight)
{
    if (!out) return ORBGPU_ERR_ARG;
    *out = nullptr;
    if (n_nodes < 1 || scoring < 0 || scoring > 5 || weighting < 0 || weighting > 3) return ORBGPU_ERR_ARG;
    if (n_nodes > 1 && (!parent || !is_leaf || !desc || !weight)) return ORBGPU_ERR_ARG;
    for (int i = 1; i < n_nodes; ++i)
        if (parent[i] < 0 || parent[i] >= i) return ORBGPU_ERR_ARG;     // a node names an earlier node as its parent
    int ndev = 0;
    CUV(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return ORBGPU_ERR_ARG;
    CUV(cudaSetDevice(device));

    // children lists in order of appearance (CSR by parent), then breadth-first positions
    std::vector<int> cstart((size_t)n_nodes + 1, 0), clist((size_t)(n_nodes > 1 ? n_nodes - 1 : 0));
    for (int i = 1; i < n_nodes; ++i) ++cstart[(size_t)parent[i] + 1];
    for (int i = 0; i < n_nodes; ++i) cstart[(size_t)i + 1] += cstart[(size_t)i];
    {
        std::vector<int> fill(cstart.begin(), cstart.end() - 1);
        for (int i = 1; i < n_nodes; ++i) clist[(size_t)fill[(size_t)parent[i]]++] = i;
    }
    std::vector<uint32_t> word_id((size_t)n_nodes, 0);
    int n_words = 0;
    for (int i = 1; i < n_nodes; ++i)
        if (is_leaf[i]) word_id[(size_t)i] = (uint32_t)n_words++;
    std::vector<int> order;                                              // order[pos] = node id
    order.reserve((size_t)n_nodes);
    order.push_back(0);
    std::vector<int2> child((size_t)n_nodes);
    for (size_t head = 0; head < order.size(); ++head) {
        const int id = order[head];
        const int nc = cstart[(size_t)id + 1] - cstart[(size_t)id];
        child[head] = make_int2((int)order.size(), nc);
        if (nc >= (1 << 20)) return ORBGPU_ERR_UNSUPPORTED;
        for (int c = 0; c < nc; ++c) order.push_back(clist[(size_t)cstart[(size_t)id] + c]);
    }
    if ((int)order.size() != n_nodes) return ORBGPU_ERR_ARG;

    const size_t N = (size_t)n_nodes;
    const size_t o_desc = 0, o_child = align256(N * 32), o_node = align256(o_child + N * 8), o_word = align256(o_node + N * 4),
                 o_weight = align256(o_word + N * 4), total = align256(o_weight + N * 8);
    std::vector<uint8_t> host(total, 0);
    for (size_t pos = 0; pos < N; ++pos) {
        const int id = order[pos];
        if (id != 0) memcpy(&host[o_desc + pos * 32], desc + (size_t)id * 32, 32);
        memcpy(&host[o_child + pos * 8], &child[pos], 8);
        const uint32_t nid = (uint32_t)id;
        memcpy(&host[o_node + pos * 4], &nid, 4);
        memcpy(&host[o_word + pos * 4], &word_id[(size_t)id], 4);
        const double w = id != 0 ? weight[id] : 0.0;
        memcpy(&host[o_weight + pos * 8], &w, 8);
    }
    orbgpu_vocabulary* v = new (std::nothrow) orbgpu_vocabulary();
    if (!v) return ORBGPU_ERR_NOMEM;
    v->device = device; v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
    v->n_nodes = n_nodes; v->n_words = n_words;
    cudaError_t e = cudaStreamCreateWithFlags(&v->st, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc(&v->d_tree, total);
    if (e == cudaSuccess) e = cudaMemcpy(v->d_tree, host.data(), total, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_bow_finalize, cudaFuncAttributeMaxDynamicSharedMemorySize, BOW_MAX_FEATURES * 8);
    if (e != cudaSuccess) {
        char b[256];
        snprintf(b, sizeof(b), "orbgpu_vocabulary_create: %s", cudaGetErrorString(e));
        orbgpu_internal_set_error(b);
        orbgpu_vocabulary_destroy(v);
        return ORBGPU_ERR_CUDA;
    }
    uint8_t* base = static_cast<uint8_t*>(v->d_tree);
    v->dev.desc = reinterpret_cast<const uint4*>(base + o_desc);
    v->dev.child = reinterpret_cast<const int2*>(base + o_child);
    v->dev.node_of = reinterpret_cast<const uint32_t*>(base + o_node);
    v->dev.word_of = reinterpret_cast<const uint32_t*>(base + o_word);
    v->dev.weight_of = reinterpret_cast<const double*>(base + o_weight);
    *out = v;
    return ORBGPU_OK;
}

void orbgpu_vocabulary_destroy(orbgpu_vocabulary* v)
{
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->st) { cudaStreamSynchronize(v->st); cudaStreamDestroy(v->st); }
    if (v->d_tree) cudaFree(v->d_tree);
    if (v->d_call) cudaFree(v->d_call);
    if (v->h_call) cudaFreeHost(v->h_call);
    if (v->d_score) cudaFree(v->d_score);
    if (v->h_score) cudaFreeHost(v->h_score);
    delete v;
}

int orbgpu_vocabulary_info(const orbgpu_vocabulary* v, int* out6)
{
    if (!v || !out6) return ORBGPU_ERR_ARG;
    out6[0] = v->k; out6[1] = v->L; out6[2] = v->scoring; out6[3] = v->weighting; out6[4] = v->n_nodes; out6[5] = v->n_words;
    return ORBGPU_OK;
}

int orbgpu_vocabulary_load_text(orbgpu_vocabulary** out, int device, const char* path)
{
    if (!out) return ORBGPU_ERR_ARG;
    *out = nullptr;
    if (!path) return ORBGPU_ERR_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return ORBGPU_ERR_ARG;
    std::string text;
    {
        char buf[1 << 16];
        size_t got;
        while ((got = fread(buf, 1, sizeof(buf), f)) > 0) text.append(buf, got);
    }
    fclose(f);
    const char* p = text.c_str();
    char* q = nullptr;
    // header line: k L scoring weighting (TemplatedVocabulary.h:1350-1362)
    const char* eol = strchr(p, '\n');
    if (!eol) eol = p + text.size();
    long hdr[4] = {0, 0, 0, 0};
    for (int i = 0; i < 4; ++i) {
        const long x = strtol(p, &q, 10);
        if (q == p || q > eol) return ORBGPU_ERR_ARG;
        hdr[i] = x; p = q;
    }
    if (hdr[0] < 0 || hdr[0] > 20 || hdr[1] < 1 || hdr[1] > 10 || hdr[2] < 0 || hdr[2] > 5 || hdr[3] < 0 || hdr[3] > 3)
        return ORBGPU_ERR_ARG;
    p = *eol ? eol + 1 : eol;
    std::vector<int32_t> parent(1, 0);
    std::vector<uint8_t> leaf(1, 0), desc(32, 0);
    std::vector<double> weight(1, 0.0);
    while (*p) {
        eol = strchr(p, '\n');
        if (!eol) eol = text.c_str() + text.size();
        const char* s = p;
        while (s < eol && (*s == ' ' || *s == '\t' || *s == '\r')) ++s;
        if (s < eol) {                                                  // "parent is_leaf d0 .. d31 weight" (:1386-1405)
            const long pid = strtol(s, &q, 10);
            if (q == s || q > eol) return ORBGPU_ERR_ARG;
            s = q;
            const long isl = strtol(s, &q, 10);
            if (q == s || q > eol) return ORBGPU_ERR_ARG;
            s = q;
            uint8_t d[32];
            for (int i = 0; i < 32; ++i) {
                const long x = strtol(s, &q, 10);
                if (q == s || q > eol) return ORBGPU_ERR_ARG;
                d[i] = (uint8_t)x; s = q;                                // FORB::fromString: (unsigned char)n
            }
            const double w = strtod(s, &q);
            if (q == s || q > eol) return ORBGPU_ERR_ARG;
            if (pid < 0 || pid >= (long)parent.size()) return ORBGPU_ERR_ARG;
            parent.push_back((int32_t)pid);
            leaf.push_back(isl > 0 ? 1 : 0);
            desc.insert(desc.end(), d, d + 32);
            weight.push_back(w);
        }
        p = *eol ? eol + 1 : eol;
    }
    return orbgpu_vocabulary_create(out, device, (int)hdr[0], (int)hdr[1], (int)hdr[2], (int)hdr[3], (int)parent.size(),
                                    parent.data(), leaf.data(), desc.data(), weight.data());
}

int orbgpu_vocabulary_transform_batch(orbgpu_vocabulary* v, const uint8_t* desc, const int32_t* frame_start, int n_frames,
                                      int levelsup, uint32_t* words, double* values, uint32_t* node_ids, int32_t* start,
                                      uint32_t* idx, int* counts3)
{
    if (!v || n_frames < 0) return ORBGPU_ERR_ARG;
    if (n_frames == 0) return ORBGPU_OK;
    if (!frame_start || !counts3 || !start || frame_start[0] != 0) return ORBGPU_ERR_ARG;
    int max_n = 0;
    for (int f = 0; f < n_frames; ++f) {
        const int nf = frame_start[f + 1] - frame_start[f];
        if (nf < 0) return ORBGPU_ERR_ARG;
        if (nf > BOW_MAX_FEATURES) return ORBGPU_ERR_CAPACITY;
        max_n = nf > max_n ? nf : max_n;
    }
    const int n = frame_start[n_frames];
    if (n > 0 && (!desc || !words || !values || !node_ids || !idx)) return ORBGPU_ERR_ARG;
    for (int f = 0; f < n_frames; ++f) {
        counts3[3 * f] = counts3[3 * f + 1] = counts3[3 * f + 2] = 0;
        start[frame_start[f] + f] = 0;
    }
    CUV(cudaSetDevice(v->device));
    if (n == 0 || v->n_words == 0) return ORBGPU_OK;                     // v.clear(); fv.clear(); empty() -> return (:1132-1138)
    const int rc = ensure_call(v, n, n_frames);
    if (rc != ORBGPU_OK) return rc;
    const CallLayout lay(v->cap, v->fcap);
    memcpy(v->h_call + lay.desc, desc, (size_t)n * 32);
    memcpy(v->h_call + lay.fstart, frame_start, ((size_t)n_frames + 1) * 4);
    // frame_start and the descriptors are neighbours in the layout: one upload
    CUV(cudaMemcpyAsync(v->d_call + lay.fstart, v->h_call + lay.fstart, lay.desc + (size_t)n * 32 - lay.fstart,
                        cudaMemcpyHostToDevice, v->st));
    uint32_t* f_word = reinterpret_cast<uint32_t*>(v->d_call + lay.f_word);
    uint32_t* f_node = reinterpret_cast<uint32_t*>(v->d_call + lay.f_node);
    double* f_weight = reinterpret_cast<double*>(v->d_call + lay.f_weight);
    k_bow_descend<<<(n + 7) / 8, 256, 0, v->st>>>(v->dev, reinterpret_cast<const uint4*>(v->d_call + lay.desc), n, v->L - levelsup,
                                                  f_word, f_node, f_weight);
    int n_pad = 64;
    while (n_pad < max_n) n_pad <<= 1;
    k_bow_finalize<<<n_frames, BOW_FIN_THREADS, (size_t)n_pad * 8, v->st>>>(
        reinterpret_cast<const int32_t*>(v->d_call + lay.fstart), f_word, f_node, f_weight, v->weighting, norm_mode_of(v->scoring),
        reinterpret_cast<uint32_t*>(v->d_call + lay.words), reinterpret_cast<double*>(v->d_call + lay.values),
        reinterpret_cast<uint32_t*>(v->d_call + lay.node_ids), reinterpret_cast<int32_t*>(v->d_call + lay.start),
        reinterpret_cast<uint32_t*>(v->d_call + lay.idx), reinterpret_cast<int*>(v->d_call + lay.counts));
    CUV(cudaGetLastError());
    // the outputs are one contiguous block [words .. counts] of the layout
    CUV(cudaMemcpyAsync(v->h_call + lay.words, v->d_call + lay.words, lay.total - lay.words, cudaMemcpyDeviceToHost, v->st));
    CUV(cudaStreamSynchronize(v->st));
    const int* c = reinterpret_cast<const int*>(v->h_call + lay.counts);
    memcpy(counts3, c, (size_t)n_frames * 12);
    for (int f = 0; f < n_frames; ++f) {                                 // only what each frame produced is copied out
        const size_t b = (size_t)frame_start[f];
        memcpy(words + b, v->h_call + lay.words + b * 4, (size_t)c[3 * f] * 4);
        memcpy(values + b, v->h_call + lay.values + b * 8, (size_t)c[3 * f] * 8);
        memcpy(node_ids + b, v->h_call + lay.node_ids + b * 4, (size_t)c[3 * f + 1] * 4);
        memcpy(start + b + f, v->h_call + lay.start + (b + f) * 4, ((size_t)c[3 * f + 1] + 1) * 4);
        memcpy(idx + b, v->h_call + lay.idx + b * 4, (size_t)c[3 * f + 2] * 4);
    }
    return ORBGPU_OK;
}

int orbgpu_vocabulary_transform(orbgpu_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* words,
                                double* values, uint32_t* node_ids, int32_t* start, uint32_t* idx, int* counts3)
{
    if (!v || !counts3 || n < 0) return ORBGPU_ERR_ARG;
    if (n > 0 && (!desc || !words || !values || !node_ids || !start || !idx)) return ORBGPU_ERR_ARG;
    if (n > BOW_MAX_FEATURES) return ORBGPU_ERR_CAPACITY;
    int32_t one_start = 0;
    const int32_t frame_start[2] = {0, n};
    return orbgpu_vocabulary_transform_batch(v, desc, frame_start, 1, levelsup, words, values, node_ids, start ? start : &one_start,
                                             idx, counts3);
}

int orbgpu_vocabulary_score(orbgpu_vocabulary* v, const uint32_t* qw, const double* qv, int nq, const uint32_t* kw,
                            const double* kv, const int32_t* kstart, int nk, double* scores)
{
    if (!v || nq < 0 || nk < 0) return ORBGPU_ERR_ARG;
    if (nk == 0) return ORBGPU_OK;
    if (!kstart || !scores || (nq > 0 && (!qw || !qv))) return ORBGPU_ERR_ARG;
    if (v->scoring != 0) return ORBGPU_ERR_UNSUPPORTED;
    const int total = kstart[nk];
    if (total < 0 || (total > 0 && (!kw || !kv))) return ORBGPU_ERR_ARG;
    CUV(cudaSetDevice(v->device));
    const size_t o_qw = 0, o_qv = align256(o_qw + (size_t)nq * 4), o_kw = align256(o_qv + (size_t)nq * 8),
                 o_kv = align256(o_kw + (size_t)total * 4), o_ks = align256(o_kv + (size_t)total * 8),
                 o_out = align256(o_ks + ((size_t)nk + 1) * 4), bytes = align256(o_out + (size_t)nk * 8);
    if (bytes > v->score_cap) {
        if (v->d_score) { cudaFree(v->d_score); v->d_score = nullptr; }
        if (v->h_score) { cudaFreeHost(v->h_score); v->h_score = nullptr; }
        v->score_cap = 0;
        const size_t cap = bytes + bytes / 2;
        CUV(cudaMalloc(&v->d_score, cap));
        CUV(cudaHostAlloc(&v->h_score, cap, cudaHostAllocDefault));
        v->score_cap = cap;
    }
    if (nq) { memcpy(v->h_score + o_qw, qw, (size_t)nq * 4); memcpy(v->h_score + o_qv, qv, (size_t)nq * 8); }
    if (total) { memcpy(v->h_score + o_kw, kw, (size_t)total * 4); memcpy(v->h_score + o_kv, kv, (size_t)total * 8); }
    memcpy(v->h_score + o_ks, kstart, ((size_t)nk + 1) * 4);
    CUV(cudaMemcpyAsync(v->d_score, v->h_score, o_out, cudaMemcpyHostToDevice, v->st));
    k_bow_score<<<(nk + 7) / 8, 256, 0, v->st>>>(reinterpret_cast<const uint32_t*>(v->d_score + o_qw),
                                                  reinterpret_cast<const double*>(v->d_score + o_qv), nq,
                                                  reinterpret_cast<const uint32_t*>(v->d_score + o_kw),
                                                  reinterpret_cast<const double*>(v->d_score + o_kv),
                                                  reinterpret_cast<const int32_t*>(v->d_score + o_ks),
                                                  reinterpret_cast<const int32_t*>(v->d_score + o_ks) + 1, nk,
                                                  reinterpret_cast<double*>(v->d_score + o_out), nullptr, nullptr);
    CUV(cudaGetLastError());
    CUV(cudaMemcpyAsync(v->h_score + o_out, v->d_score + o_out, (size_t)nk * 8, cudaMemcpyDeviceToHost, v->st));
    CUV(cudaStreamSynchronize(v->st));
    memcpy(scores, v->h_score + o_out, (size_t)nk * 8);
    return ORBGPU_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------- KeyFrameDatabase: resident BowVectors
// src/KeyFrameDatabase.cc keeps an inverted file (word -> keyframes) and, per query, walks the lists of the query's words to
// count the words each keyframe shares (:86-105, :207-222) before scoring the ones above 0.8 * max (:107-139, :224-255).  Here
// the BowVectors of the keyframes stay in HBM and one launch of k_bow_score gives, for EVERY live keyframe, the number of shared
// words, the position of the first shared word in the query (the inverted-file walk meets keyframes in that order, then in
// insertion order) and the L1 score.
struct orbgpu_bowdb {
    int device = 0;
    cudaStream_t st = nullptr;
    size_t cap = 0, used = 0;                 // entries of the arena
    uint32_t* d_words = nullptr;
    double* d_values = nullptr;
    std::vector<int32_t> begin, end;          // per slot; end == begin for an erased slot
    bool table_dirty = true;
    int slot_cap = 0;
    int32_t* d_begin = nullptr; int32_t* d_end = nullptr;
    int32_t* d_common = nullptr; int32_t* d_first = nullptr; double* d_score = nullptr;
    size_t stage_cap = 0;                     // bytes
    uint8_t* h_stage = nullptr;               // pinned: uploads and results
    int q_cap = 0;
    uint32_t* d_qw = nullptr; double* d_qv = nullptr;
};

namespace {
int bowdb_stage(orbgpu_bowdb* db, size_t bytes)
{
    if (bytes <= db->stage_cap) return ORBGPU_OK;
    if (db->h_stage) { cudaFreeHost(db->h_stage); db->h_stage = nullptr; db->stage_cap = 0; }
    const size_t cap = align256(bytes + bytes / 2 + 4096);
    CUV(cudaHostAlloc(&db->h_stage, cap, cudaHostAllocDefault));
    db->stage_cap = cap;
    return ORBGPU_OK;
}
}  // namespace

extern "C" {

int orbgpu_bowdb_create(orbgpu_bowdb** out, int device)
{
    if (!out) return ORBGPU_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    CUV(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return ORBGPU_ERR_ARG;
    CUV(cudaSetDevice(device));
    orbgpu_bowdb* db = new (std::nothrow) orbgpu_bowdb();
    if (!db) return ORBGPU_ERR_NOMEM;
    db->device = device;
    const cudaError_t e = cudaStreamCreateWithFlags(&db->st, cudaStreamNonBlocking);
    if (e != cudaSuccess) { delete db; orbgpu_internal_set_error(cudaGetErrorString(e)); return ORBGPU_ERR_CUDA; }
    *out = db;
    return ORBGPU_OK;
}

void orbgpu_bowdb_destroy(orbgpu_bowdb* db)
{
    if (!db) return;
    cudaSetDevice(db->device);
    if (db->st) { cudaStreamSynchronize(db->st); cudaStreamDestroy(db->st); }
    cudaFree(db->d_words); cudaFree(db->d_values); cudaFree(db->d_begin); cudaFree(db->d_end);
    cudaFree(db->d_common); cudaFree(db->d_first); cudaFree(db->d_score); cudaFree(db->d_qw); cudaFree(db->d_qv);
    if (db->h_stage) cudaFreeHost(db->h_stage);
    delete db;
}

int orbgpu_bowdb_add(orbgpu_bowdb* db, const uint32_t* words, const double* values, int n, int* slot)
{
    if (!db || n < 0 || (n > 0 && (!words || !values))) return ORBGPU_ERR_ARG;
    CUV(cudaSetDevice(db->device));
    if (db->used + (size_t)n > db->cap) {                             // grow the arena (doubling), keeping what is there
        size_t cap = db->cap ? db->cap : ((size_t)1 << 20);
        while (cap < db->used + (size_t)n) cap <<= 1;
        if (cap > 0x7FFFFFFFull) return ORBGPU_ERR_CAPACITY;
        uint32_t* nw = nullptr; double* nv = nullptr;
        CUV(cudaMalloc(&nw, cap * 4));
        cudaError_t e = cudaMalloc(&nv, cap * 8);
        if (e != cudaSuccess) { cudaFree(nw); orbgpu_internal_set_error(cudaGetErrorString(e)); return ORBGPU_ERR_CUDA; }
        CUV(cudaStreamSynchronize(db->st));
        if (db->used) {
            CUV(cudaMemcpy(nw, db->d_words, db->used * 4, cudaMemcpyDeviceToDevice));
            CUV(cudaMemcpy(nv, db->d_values, db->used * 8, cudaMemcpyDeviceToDevice));
        }
        cudaFree(db->d_words); cudaFree(db->d_values);
        db->d_words = nw; db->d_values = nv; db->cap = cap;
    }
    if (n > 0) {
        CUV(cudaStreamSynchronize(db->st));                           // the staging block may still feed an earlier copy
        const int rc = bowdb_stage(db, (size_t)n * 12 + 256);
        if (rc != ORBGPU_OK) return rc;
        const size_t o_v = align256((size_t)n * 4);
        memcpy(db->h_stage, words, (size_t)n * 4);
        memcpy(db->h_stage + o_v, values, (size_t)n * 8);
        CUV(cudaMemcpyAsync(db->d_words + db->used, db->h_stage, (size_t)n * 4, cudaMemcpyHostToDevice, db->st));
        CUV(cudaMemcpyAsync(db->d_values + db->used, db->h_stage + o_v, (size_t)n * 8, cudaMemcpyHostToDevice, db->st));
    }
    db->begin.push_back((int32_t)db->used);
    db->end.push_back((int32_t)(db->used + (size_t)n));
    db->used += (size_t)n;
    db->table_dirty = true;
    if (slot) *slot = (int)db->begin.size() - 1;
    return ORBGPU_OK;
}

int orbgpu_bowdb_erase(orbgpu_bowdb* db, int slot)
{
    if (!db || slot < 0 || slot >= (int)db->begin.size()) return ORBGPU_ERR_ARG;
    db->end[(size_t)slot] = db->begin[(size_t)slot];
    db->table_dirty = true;
    return ORBGPU_OK;
}

int orbgpu_bowdb_clear(orbgpu_bowdb* db)
{
    if (!db) return ORBGPU_ERR_ARG;
    CUV(cudaSetDevice(db->device));
    CUV(cudaStreamSynchronize(db->st));
    db->begin.clear(); db->end.clear(); db->used = 0; db->table_dirty = true;
    return ORBGPU_OK;
}

int orbgpu_bowdb_size(const orbgpu_bowdb* db) { return db ? (int)db->begin.size() : ORBGPU_ERR_ARG; }

int orbgpu_bowdb_query(orbgpu_bowdb* db, const uint32_t* qw, const double* qv, int nq, int32_t* common_words,
                       int32_t* first_common, double* scores)
{
    if (!db || nq < 0 || (nq > 0 && (!qw || !qv))) return ORBGPU_ERR_ARG;
    const int ns = (int)db->begin.size();
    if (ns == 0) return ORBGPU_OK;
    if (!common_words || !first_common || !scores) return ORBGPU_ERR_ARG;
    CUV(cudaSetDevice(db->device));
    CUV(cudaStreamSynchronize(db->st));
    if (ns > db->slot_cap) {
        int cap = db->slot_cap ? db->slot_cap : 1024;
        while (cap < ns) cap <<= 1;
        cudaFree(db->d_begin); cudaFree(db->d_end); cudaFree(db->d_common); cudaFree(db->d_first); cudaFree(db->d_score);
        db->d_begin = db->d_end = db->d_common = db->d_first = nullptr; db->d_score = nullptr; db->slot_cap = 0;
        CUV(cudaMalloc(&db->d_begin, (size_t)cap * 4)); CUV(cudaMalloc(&db->d_end, (size_t)cap * 4));
        CUV(cudaMalloc(&db->d_common, (size_t)cap * 4)); CUV(cudaMalloc(&db->d_first, (size_t)cap * 4));
        CUV(cudaMalloc(&db->d_score, (size_t)cap * 8));
        db->slot_cap = cap; db->table_dirty = true;
    }
    if (nq > db->q_cap) {
        int cap = db->q_cap ? db->q_cap : 4096;
        while (cap < nq) cap <<= 1;
        cudaFree(db->d_qw); cudaFree(db->d_qv); db->d_qw = nullptr; db->d_qv = nullptr; db->q_cap = 0;
        CUV(cudaMalloc(&db->d_qw, (size_t)cap * 4)); CUV(cudaMalloc(&db->d_qv, (size_t)cap * 8));
        db->q_cap = cap;
    }
    // staging: [begin | end] (when dirty) | qw | qv ... results: common | first | score
    const size_t o_b = 0, o_e = align256((size_t)ns * 4), o_qw = align256(o_e + (size_t)ns * 4), o_qv = align256(o_qw + (size_t)nq * 4),
                 o_c = align256(o_qv + (size_t)nq * 8), o_f = align256(o_c + (size_t)ns * 4), o_s = align256(o_f + (size_t)ns * 4),
                 total = align256(o_s + (size_t)ns * 8);
    const int rc = bowdb_stage(db, total);
    if (rc != ORBGPU_OK) return rc;
    if (db->table_dirty) {
        memcpy(db->h_stage + o_b, db->begin.data(), (size_t)ns * 4);
        memcpy(db->h_stage + o_e, db->end.data(), (size_t)ns * 4);
        CUV(cudaMemcpyAsync(db->d_begin, db->h_stage + o_b, (size_t)ns * 4, cudaMemcpyHostToDevice, db->st));
        CUV(cudaMemcpyAsync(db->d_end, db->h_stage + o_e, (size_t)ns * 4, cudaMemcpyHostToDevice, db->st));
        db->table_dirty = false;
    }
    if (nq) {
        memcpy(db->h_stage + o_qw, qw, (size_t)nq * 4);
        memcpy(db->h_stage + o_qv, qv, (size_t)nq * 8);
        CUV(cudaMemcpyAsync(db->d_qw, db->h_stage + o_qw, (size_t)nq * 4, cudaMemcpyHostToDevice, db->st));
        CUV(cudaMemcpyAsync(db->d_qv, db->h_stage + o_qv, (size_t)nq * 8, cudaMemcpyHostToDevice, db->st));
    }
    k_bow_score<<<(ns + 7) / 8, 256, 0, db->st>>>(db->d_qw, db->d_qv, nq, db->d_words, db->d_values, db->d_begin, db->d_end, ns,
                                                  db->d_score, db->d_common, db->d_first);
    CUV(cudaGetLastError());
    CUV(cudaMemcpyAsync(db->h_stage + o_c, db->d_common, (size_t)ns * 4, cudaMemcpyDeviceToHost, db->st));
    CUV(cudaMemcpyAsync(db->h_stage + o_f, db->d_first, (size_t)ns * 4, cudaMemcpyDeviceToHost, db->st));
    CUV(cudaMemcpyAsync(db->h_stage + o_s, db->d_score, (size_t)ns * 8, cudaMemcpyDeviceToHost, db->st));
    CUV(cudaStreamSynchronize(db->st));
    memcpy(common_words, db->h_stage + o_c, (size_t)ns * 4);
    memcpy(first_common, db->h_stage + o_f, (size_t)ns * 4);
    memcpy(scores, db->h_stage + o_s, (size_t)ns * 8);
    return ORBGPU_OK;
}

}  // extern "C"
