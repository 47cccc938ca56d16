// orbgpu_extractor.cu -- host engine behind the orbgpu_* extractor entry points (include/orbgpu.h).
//
// One extractor handle = the reference's ORBextractor instance (include/ORBextractor.h:45-111):
// constructor tables, a per-shape plan (level sizes, FAST cell table, octree budgets, resize
// tables), and NSTREAM stream contexts that each own the device + pinned buffers of one chunk of
// frames, so upload, kernels and download of successive chunks overlap.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/orbgpu.h"
#include "orb_common.cuh"

static thread_local char g_cuda_err[256] = "";
extern "C" const char* orbgpu_last_cuda_error(void) { return g_cuda_err; }
void orbgpu_internal_set_error(const char* s) { snprintf(g_cuda_err, sizeof(g_cuda_err), "%s", s); }

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t _e = (call);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            snprintf(g_cuda_err, sizeof(g_cuda_err), "%s at %s:%d: %s", #call, __FILE__, __LINE__, \
                     cudaGetErrorString(_e));                                                      \
            return ORBGPU_ERR_CUDA;                                                                \
        }                                                                                          \
    } while (0)

namespace {

const int NSTREAM = 8;               // stream contexts per handle (queue depth); batch calls use `nstream` of them
enum { EV_BEGIN = 0, EV_UP, EV_PYR, EV_FAST, EV_OCT, EV_BLUR, EV_DESC, EV_DOWN, EV_COUNT };

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct Plan {
    int w = 0, h = 0;
    OrbPlanDev dev;                    // offsets + device table pointers; per-chunk pointers filled per launch
    std::vector<OrbCell> cells;
    OrbCell* d_cells = nullptr;
    short* d_rs_ofs = nullptr;
    short* d_rs_a = nullptr;
    int4* d_tiles = nullptr;
    int ntiles = 0;
    int fast_smem = 0, key_smem_cap = 0, node_smem_cap = 0;
    int kmax = 0;                      // max keypoints per frame
    size_t in_pitch = 0;               // row pitch of the upload buffer d_in: w rounded up to 16 B (aligned word loads)
    size_t in_frame_bytes = 0;         // in_pitch * h
    size_t bytes_per_frame = 0;
    int chunk_cap = 1;
    void free_dev()
    {
        cudaFree(d_cells); cudaFree(d_rs_ofs); cudaFree(d_rs_a); cudaFree(d_tiles);
        d_cells = nullptr; d_rs_ofs = nullptr; d_rs_a = nullptr; d_tiles = nullptr;
    }
};

struct StreamCtx {
    cudaStream_t st = nullptr;
    cudaStream_t st_side = nullptr;    // blur runs here, beside FAST + octree of the same chunk (both only need the pyramid)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaEvent_t ev[EV_COUNT] = {};
    cudaEvent_t done = nullptr;
    int cap = 0;                       // frames the buffers below can hold
    uint8_t* d_raw = nullptr; size_t raw_bytes = 0;   // dense copy of unaligned frames before k_repitch (lazily allocated)
    uint8_t *d_in = nullptr, *d_pyr = nullptr, *d_blur = nullptr, *d_scratch = nullptr, *d_desc = nullptr;
    uint32_t *d_cellcnt = nullptr, *d_slots = nullptr, *d_sel = nullptr;
    int *d_selcnt = nullptr, *d_candcnt = nullptr, *d_nout = nullptr;
    orbgpu_keypoint* d_kps = nullptr;
    uint8_t *h_in = nullptr, *h_desc = nullptr;     // pinned staging
    orbgpu_keypoint* h_kps = nullptr;
    int* h_nout = nullptr;
    // pending job
    bool busy = false, profiled = false;
    int job_frames = 0;
    bool staged_out = false;
    orbgpu_keypoint* user_kps = nullptr; uint8_t* user_desc = nullptr; int* user_nout = nullptr;
    int user_capacity = 0;
    uint64_t ticket = 0; bool ticket_live = false;
    int job_w = 0, job_h = 0;
    // the chunk's kernels (pyramid .. describe, with the blur fork / join on the side stream) captured once as a CUDA graph
    // per (buffers, frame count): one cudaGraphLaunch per chunk instead of 11 launches + 4 event calls from the host thread
    // (a few frame counts are kept: the last chunk of a batch is usually shorter than the others)
    static const int NGRAPH = 4;
    cudaGraphExec_t gexec[NGRAPH] = {};
    int g_frames[NGRAPH] = {}, g_launches[NGRAPH] = {};
    uint64_t g_epoch[NGRAPH] = {};     // plan generation each graph was captured for
    uint64_t g_used[NGRAPH] = {}, g_clock = 0;
};

}  // namespace

struct orbgpu_extractor {
    int device = 0;
    int nfeatures = 0, nlevels = 0, ini_th = 0, min_th = 0;
    double scale_factor = 0;           // the reference keeps the ctor's float in a double member
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> feats;
    int umax[16];
    Plan plan;
    StreamCtx ctx[NSTREAM];
    bool profiling = false;
    // how a batch call is cut up: [0] host-buffer calls (finer chunks pipeline PCIe against the kernels),
    // [1] device-buffer calls (fewer, larger launches); orbgpu_set_batching overrides both
    int nstream[2] = {8, 4};            // chunk streams a batch call deals its chunks to (1..NSTREAM)
    int max_chunk[2] = {64, 128};       // frames per chunk (upper bound; also capped by plan.chunk_cap); B200 sweep: profiles/r2_batching_sweep.txt
    double stage_ms[ORBGPU_NUM_STAGES] = {};
    uint64_t launches = 0;
    uint64_t next_ticket = 1;
    int in_channels = 1, in_rgb = 0;   // input pixel format of host/device frames (orbgpu_set_input_format)
    bool async_open = false;           // orbgpu_extract_batch_async batches outstanding (closed by orbgpu_wait)
    int next_ctx = 0;                  // next chunk stream of the host-batch round-robin
    cudaEvent_t ev_batch0 = nullptr, ev_batch1 = nullptr;   // bracket the most recent batch call on ctx[0].st
    double last_batch_ms = 0;
    uint8_t* d_aux = nullptr; size_t d_aux_cap = 0;   // stereo / undistort outputs of the last frame, grown on demand
    uint8_t* h_aux = nullptr; size_t h_aux_cap = 0;   // pinned mirror
    int last_ctx = -1;                 // context holding the most recent single frame (pyramid / debug views)
    bool use_graphs = true;            // host-buffer chunks replay a captured graph of their kernels (orbgpu_set_graphs)
    uint64_t epoch = 1;                // bumped whenever a plan or a context's buffers change: invalidates captured graphs
    // scratch for orbgpu_octree
    uint8_t* d_oct_scratch = nullptr; size_t oct_scratch_bytes = 0;
};

// -------------------------------------------------------------------------------------------------
// Constructor tables: same float/double mix as src/ORBextractor.cc:412-472.
// -------------------------------------------------------------------------------------------------
static void build_tables(orbgpu_extractor* e, float scale_factor)
{
    const int nl = e->nlevels;
    e->scale_factor = scale_factor;
    e->scale.assign(nl, 1.0f); e->sigma2.assign(nl, 1.0f);
    for (int i = 1; i < nl; ++i) {
        e->scale[i] = (float)(e->scale[i - 1] * e->scale_factor);
        e->sigma2[i] = e->scale[i] * e->scale[i];
    }
    e->inv_scale.resize(nl); e->inv_sigma2.resize(nl);
    for (int i = 0; i < nl; ++i) { e->inv_scale[i] = 1.0f / e->scale[i]; e->inv_sigma2[i] = 1.0f / e->sigma2[i]; }

    e->feats.assign(nl, 0);
    const float factor = (float)(1.0f / e->scale_factor);
    float per_scale = e->nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        e->feats[l] = (int)lrintf(per_scale);
        sum += e->feats[l];
        per_scale *= factor;
    }
    e->feats[nl - 1] = std::max(e->nfeatures - sum, 0);

    // row half-widths of the 31-px circular patch
    const int hp = ORB_HALF_PATCH;
    int* umax = e->umax;
    memset(umax, 0, sizeof(e->umax));
    const int vmax = (int)floor(hp * sqrtf(2.f) / 2 + 1), vmin = (int)ceil(hp * sqrtf(2.f) / 2);
    for (int v = 0; v <= vmax; ++v) umax[v] = (int)lrint(sqrt((double)hp * hp - v * v));
    for (int v = hp, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

// per-axis tables of cv::resize INTER_LINEAR 8U (see k_resize_level)
static void resize_axis(int ssize, int dsize, bool clamp_fraction, short* ofs, short* a)
{
    const double scale = 1.0 / ((double)dsize / ssize);
    for (int d = 0; d < dsize; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)floorf(f);
        f -= (float)s;
        if (clamp_fraction) {            // x axis: out-of-range taps collapse onto the edge pixel
            if (s < 0) { s = 0; f = 0.f; }
            if (s >= ssize - 1) { s = ssize - 1; f = 0.f; }
        }                                // y axis: both row indices are clipped in the kernel instead
        ofs[d] = (short)s;
        a[2 * d] = (short)lrintf((1.f - f) * 2048);
        a[2 * d + 1] = (short)lrintf(f * 2048);
    }
}

// -------------------------------------------------------------------------------------------------
// Per-shape plan: pyramid level sizes (src/ORBextractor.cc:1113-1114), FAST cell grid (:771-810),
// octree roots and budgets (:544-546), buffer offsets.
// -------------------------------------------------------------------------------------------------
static int build_plan(orbgpu_extractor* e, int w, int h)
{
    Plan& pl = e->plan;
    if (pl.w == w && pl.h == h) return ORBGPU_OK;
    for (int s = 0; s < NSTREAM; ++s)
        if (e->ctx[s].busy || e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;   // shape change with work in flight
    if (w < 1 || h < 1 || w > 32767 || h > 32767) return ORBGPU_ERR_UNSUPPORTED;

    Plan np;
    np.w = w; np.h = h;
    OrbPlanDev& D = np.dev;
    memset(&D, 0, sizeof(D));
    D.nlevels = e->nlevels; D.ini_th = e->ini_th; D.min_th = e->min_th;

    std::vector<short> rs_ofs, rs_a;
    std::vector<int4> tiles;
    size_t pyr = 0, blur = 0, scratch = 0;
    long long slot_words = 0;
    int sel_total = 0, max_cw = 7, max_ch = 7, max_node_cap = 0;

    for (int l = 0; l < e->nlevels; ++l) {
        OrbLevel& L = D.lv[l];
        const float s = e->inv_scale[l];
        L.w = (int)lrintf((float)w * s);
        L.h = (int)lrintf((float)h * s);
        if (L.w < 1 || L.h < 1) return ORBGPU_ERR_UNSUPPORTED;
        L.pitch = (int)align_up((size_t)L.w, 16);
        if (l > 0) { L.pyr_off = (long long)pyr; pyr += (size_t)L.pitch * L.h; }
        L.blur_off = (long long)blur; blur += (size_t)L.pitch * L.h;
        L.nfeat = e->feats[l];
        L.scale = e->scale[l];
        L.kp_size = (float)(int)(ORB_PATCH * e->scale[l]);
        if (l > 0) {
            const OrbLevel& S = D.lv[l - 1];
            if (S.w == 2 * L.w && S.h == 2 * L.h) return ORBGPU_ERR_UNSUPPORTED;   // cv::resize would switch to INTER_AREA
            L.xofs_off = (int)rs_ofs.size();
            rs_ofs.resize(rs_ofs.size() + L.w); rs_a.resize(rs_a.size() + 2 * (size_t)L.w);
            resize_axis(S.w, L.w, true, &rs_ofs[L.xofs_off], &rs_a[2 * (size_t)L.xofs_off]);
            L.yofs_off = (int)rs_ofs.size();
            rs_ofs.resize(rs_ofs.size() + L.h); rs_a.resize(rs_a.size() + 2 * (size_t)L.h);
            resize_axis(S.h, L.h, false, &rs_ofs[L.yofs_off], &rs_a[2 * (size_t)L.yofs_off]);
        }
        {   // blur work list: one entry per CTA of ORB_BLUR_THREADS tasks; task = (row strip, 4-pixel word column).
            // kind 0: interior columns on the aligned word path; kind 1: row-end columns on the byte path.  Level 0
            // is read from the caller's buffer, whose alignment is only known at launch: both kinds are listed for
            // the worst case and the kernel drops the entries that do not apply.
            const int nwc = (L.w + 3) / 4, nstrip = (L.h + ORB_BLUR_ROWS - 1) / ORB_BLUR_ROWS;
            const int nfast = L.w >= 8 ? (L.w - 4) / 4 : 0;
            for (int t0 = 0; t0 < nfast * nstrip; t0 += ORB_BLUR_THREADS) tiles.push_back(make_int4(l, t0, 0, 0));
            const int nslow_max = (l == 0) ? nwc : nwc - nfast;     // level 0 may turn out unaligned: all columns slow
            for (int t0 = 0; t0 < nslow_max * nstrip; t0 += ORB_BLUR_THREADS) tiles.push_back(make_int4(l, t0, 1, 0));
        }
        // FAST cell grid
        const int minB = ORB_BORDER, maxBX = L.w - ORB_BORDER, maxBY = L.h - ORB_BORDER;
        L.W = maxBX - minB; L.H = maxBY - minB;
        const float width = (float)L.W, height = (float)L.H;
        const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
        L.cell0 = (int)np.cells.size();
        L.slot_off = slot_words;
        int slot_total = 0;
        if (nCols >= 1 && nRows >= 1 && L.W > 0 && L.H > 0) {
            if (L.W > 4096 || L.H > 4096) return ORBGPU_ERR_UNSUPPORTED;     // 12-bit packed coordinates
            const int wCell = (int)ceilf(width / nCols), hCell = (int)ceilf(height / nRows);
            for (int i = 0; i < nRows; ++i) {
                const int iniY = minB + i * hCell;
                if (iniY >= maxBY - 3) continue;
                const int maxY = std::min(iniY + hCell + 6, maxBY);
                for (int j = 0; j < nCols; ++j) {
                    const int iniX = minB + j * wCell;
                    if (iniX >= maxBX - 6) continue;
                    const int maxX = std::min(iniX + wCell + 6, maxBX);
                    OrbCell c;
                    memset(&c, 0, sizeof(c));
                    c.level = (short)l;
                    c.cw = (short)(maxX - iniX); c.ch = (short)(maxY - iniY);
                    c.ini_x = (short)iniX; c.ini_y = (short)iniY;
                    c.off_x = (short)(j * wCell + 3); c.off_y = (short)(i * hCell + 3);
                    const int iw = std::max(c.cw - 6, 0), ih = std::max(c.ch - 6, 0);
                    c.cap = ((iw + 1) / 2) * ((ih + 1) / 2);
                    c.slot_off = slot_total;
                    slot_total += c.cap;
                    max_cw = std::max(max_cw, (int)c.cw); max_ch = std::max(max_ch, (int)c.ch);
                    np.cells.push_back(c);
                }
            }
            L.n_ini = (int)roundf((float)L.W / (float)L.H);
            if (L.n_ini < 1) return ORBGPU_ERR_UNSUPPORTED;   // the reference indexes an empty vector here (:571)
            L.hx = (float)L.W / L.n_ini;
        } else {
            L.n_ini = 0; L.hx = 1.f;                          // level too small for one 30-px cell: no keypoints
        }
        L.ncells = (int)np.cells.size() - L.cell0;
        L.slot_total = slot_total;
        if (slot_total >= (1 << 24)) return ORBGPU_ERR_UNSUPPORTED;
        slot_words += slot_total;
        L.node_cap = L.ncells ? std::max(L.nfeat, 4 * L.n_ini) + 4 : 0;
        if (L.node_cap > 65535) return ORBGPU_ERR_UNSUPPORTED;
        L.sel_cap = L.node_cap; L.sel_off = sel_total; sel_total += L.sel_cap;
        max_node_cap = std::max(max_node_cap, L.node_cap);
        L.key_scratch_off = (long long)scratch; scratch += align_up((size_t)7 * std::max(slot_total, 1), 16);
        L.node_scratch_off = (long long)scratch; scratch += align_up((size_t)64 * L.node_cap + 64, 16);
    }
    D.ncells = (int)np.cells.size();
    D.sel_total = sel_total;
    D.pyr_frame_bytes = (long long)align_up(pyr, 256);
    D.blur_frame_bytes = (long long)align_up(blur, 256);
    D.slot_frame_words = (long long)align_up((size_t)slot_words, 64);
    D.scratch_frame_bytes = (long long)align_up(scratch, 256);
    np.kmax = sel_total;
    np.ntiles = (int)tiles.size();
    np.fast_smem = orb_fast_smem_bytes(max_cw, max_ch);
    D.fast_rows = max_ch; D.fast_spw = orb_fast_spw(max_cw); D.fast_tp = orb_fast_tile_pitch(max_cw);
    if (np.fast_smem < 0 || np.fast_smem > 200 * 1024) return ORBGPU_ERR_UNSUPPORTED;
    np.key_smem_cap = std::min(std::max(8 * D.lv[0].nfeat, 1024), 6144);
    np.node_smem_cap = std::min(max_node_cap, 1024);
    {
        // k_octree runs 4 CTAs per SM (64 registers x 256 threads): keep 4 x (shared memory + 1 KB) inside the 196 KB carve-out
        // step, or every kernel co-resident with it loses its L1.  Node tables first (they are touched in
        // every pass), keys with what is left; whatever does not fit lives in the global scratch area as before.
        // B200, C2 (1241 x 376, 2000 features: 52 KB before): 89 k -> 123 k frames/s with 47 KB; 39 KB and 31 KB are no better.
        const int budget = 47 * 1024 - 64;
        if (64 * np.node_smem_cap > budget - 7 * 1024) np.node_smem_cap = 0;          // node tables too large: all in scratch
        const int left = budget - 64 * np.node_smem_cap;
        np.key_smem_cap = std::max(std::min(np.key_smem_cap, left / 7), 16);
    }
    np.in_pitch = align_up((size_t)w, 16);
    np.in_frame_bytes = np.in_pitch * h;
    np.bytes_per_frame = np.in_frame_bytes + (size_t)D.pyr_frame_bytes + (size_t)D.blur_frame_bytes +
                         (size_t)D.slot_frame_words * 4 + (size_t)D.scratch_frame_bytes + (size_t)D.ncells * 4 +
                         (size_t)sel_total * 4 + (size_t)np.kmax * 60 + 256;
    np.chunk_cap = (int)std::min<size_t>(128, std::max<size_t>(1, ((size_t)1536 << 20) / np.bytes_per_frame));

    struct Guard { Plan& p; bool keep = false; ~Guard() { if (!keep) p.free_dev(); } } guard{np};   // a failed upload frees what was allocated
    CU(cudaMalloc(&np.d_cells, sizeof(OrbCell) * std::max<size_t>(np.cells.size(), 1)));
    CU(cudaMalloc(&np.d_rs_ofs, sizeof(short) * std::max<size_t>(rs_ofs.size(), 1)));
    CU(cudaMalloc(&np.d_rs_a, sizeof(short) * std::max<size_t>(rs_a.size(), 1)));
    CU(cudaMalloc(&np.d_tiles, sizeof(int4) * std::max<size_t>(tiles.size(), 1)));
    if (!np.cells.empty()) CU(cudaMemcpy(np.d_cells, np.cells.data(), sizeof(OrbCell) * np.cells.size(), cudaMemcpyHostToDevice));
    if (!rs_ofs.empty()) CU(cudaMemcpy(np.d_rs_ofs, rs_ofs.data(), sizeof(short) * rs_ofs.size(), cudaMemcpyHostToDevice));
    if (!rs_a.empty()) CU(cudaMemcpy(np.d_rs_a, rs_a.data(), sizeof(short) * rs_a.size(), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(np.d_tiles, tiles.data(), sizeof(int4) * tiles.size(), cudaMemcpyHostToDevice));
    D.cells = np.d_cells; D.rs_ofs = np.d_rs_ofs; D.rs_a = np.d_rs_a;

    pl.free_dev();
    pl = np;
    guard.keep = true;
    e->last_ctx = -1;
    ++e->epoch;
    return ORBGPU_OK;
}

static void ctx_free_buffers(StreamCtx& c)
{
    for (int k = 0; k < StreamCtx::NGRAPH; ++k) {
        if (c.gexec[k]) { cudaGraphExecDestroy(c.gexec[k]); c.gexec[k] = nullptr; }
        c.g_frames[k] = 0;
    }
    cudaFree(c.d_raw); c.d_raw = nullptr; c.raw_bytes = 0;
    cudaFree(c.d_in); cudaFree(c.d_pyr); cudaFree(c.d_blur); cudaFree(c.d_scratch); cudaFree(c.d_desc);
    cudaFree(c.d_cellcnt); cudaFree(c.d_slots); cudaFree(c.d_sel); cudaFree(c.d_selcnt); cudaFree(c.d_candcnt);
    cudaFree(c.d_nout); cudaFree(c.d_kps);
    cudaFreeHost(c.h_in); cudaFreeHost(c.h_desc); cudaFreeHost(c.h_kps); cudaFreeHost(c.h_nout);
    c.d_in = c.d_pyr = c.d_blur = c.d_scratch = c.d_desc = nullptr;
    c.d_cellcnt = c.d_slots = c.d_sel = nullptr;
    c.d_selcnt = c.d_candcnt = c.d_nout = nullptr; c.d_kps = nullptr;
    c.h_in = c.h_desc = nullptr; c.h_kps = nullptr; c.h_nout = nullptr;
    c.cap = 0;
}

// (re)allocate one context's buffers for `frames` frames of the current plan
static int ctx_reserve(orbgpu_extractor* e, StreamCtx& c, int frames, int plan_w, int plan_h)
{
    const Plan& pl = e->plan;
    if (c.cap >= frames && c.job_w == plan_w && c.job_h == plan_h) return ORBGPU_OK;
    ctx_free_buffers(c);                                      // also drops the context's captured graph
    const OrbPlanDev& D = pl.dev;
    const size_t n = (size_t)frames;
    CU(cudaMalloc(&c.d_in, n * pl.in_frame_bytes + 64));
    CU(cudaMalloc(&c.d_pyr, n * (size_t)D.pyr_frame_bytes + 64));
    CU(cudaMalloc(&c.d_blur, n * (size_t)D.blur_frame_bytes + 64));
    CU(cudaMalloc(&c.d_scratch, n * (size_t)D.scratch_frame_bytes + 64));
    CU(cudaMalloc(&c.d_cellcnt, n * (size_t)std::max(D.ncells, 1) * 4));
    CU(cudaMalloc(&c.d_slots, n * (size_t)std::max<long long>(D.slot_frame_words, 1) * 4));
    CU(cudaMalloc(&c.d_sel, n * (size_t)std::max(D.sel_total, 1) * 4));
    CU(cudaMalloc(&c.d_selcnt, n * ORB_MAX_LEVELS * 4));
    CU(cudaMalloc(&c.d_candcnt, n * ORB_MAX_LEVELS * 4));
    CU(cudaMalloc(&c.d_nout, n * 4));
    CU(cudaMalloc(&c.d_kps, n * (size_t)std::max(pl.kmax, 1) * sizeof(orbgpu_keypoint)));
    CU(cudaMalloc(&c.d_desc, n * (size_t)std::max(pl.kmax, 1) * 32));
    CU(cudaHostAlloc(&c.h_in, n * (size_t)pl.w * pl.h * e->in_channels + 64, cudaHostAllocDefault));
    CU(cudaHostAlloc(&c.h_kps, n * (size_t)std::max(pl.kmax, 1) * sizeof(orbgpu_keypoint), cudaHostAllocDefault));
    CU(cudaHostAlloc(&c.h_desc, n * (size_t)std::max(pl.kmax, 1) * 32, cudaHostAllocDefault));
    CU(cudaHostAlloc(&c.h_nout, n * 4, cudaHostAllocDefault));
    c.cap = frames; c.job_w = plan_w; c.job_h = plan_h;
    return ORBGPU_OK;
}

// -------------------------------------------------------------------------------------------------
// One chunk of frames through the kernels, on the context's stream.
// -------------------------------------------------------------------------------------------------
static int enqueue_kernels(orbgpu_extractor* e, StreamCtx& c, const uint8_t* d_src, long long src_frame_stride,
                           int src_pitch, int nframes, orbgpu_keypoint* d_kps, uint8_t* d_desc, int* d_nout,
                           int kp_capacity)
{
    const Plan& pl = e->plan;
    OrbPlanDev P = pl.dev;
    P.src = d_src; P.src_frame_stride = src_frame_stride; P.src_pitch = src_pitch;
    P.pyr = c.d_pyr; P.blur = c.d_blur; P.cellcnt = c.d_cellcnt; P.slots = c.d_slots; P.sel = c.d_sel;
    P.selcnt = c.d_selcnt; P.candcnt = c.d_candcnt; P.scratch = c.d_scratch;
    P.kps = d_kps; P.desc = d_desc; P.nout = d_nout; P.kp_capacity = kp_capacity;
    const bool prof = e->profiling;
    int n, total = 0;
    if ((n = orb_launch_pyramid(P, nframes, c.st)) < 0) return ORBGPU_ERR_CUDA; total += n;
    if (prof) CU(cudaEventRecord(c.ev[EV_PYR], c.st));
    // blur only needs the pyramid: it runs on the side stream next to FAST + octree of the same chunk and joins before
    // describe (throughput of a busy GPU is unchanged, a single frame's latency drops by the blur time); the profiling
    // mode keeps everything on one stream so that the stage events bracket one kernel each
    const bool side = !prof && c.st_side;
    if (side) {
        CU(cudaEventRecord(c.ev_fork, c.st));
        CU(cudaStreamWaitEvent(c.st_side, c.ev_fork, 0));
        if ((n = orb_launch_blur(P, nframes, pl.d_tiles, pl.ntiles, c.st_side)) < 0) return ORBGPU_ERR_CUDA; total += n;
        CU(cudaEventRecord(c.ev_join, c.st_side));
    }
    if (P.ncells > 0) { if ((n = orb_launch_fast(P, nframes, pl.fast_smem, c.st)) < 0) return ORBGPU_ERR_CUDA; total += n; }
    if (prof) CU(cudaEventRecord(c.ev[EV_FAST], c.st));
    if ((n = orb_launch_octree(P, nframes, pl.key_smem_cap, pl.node_smem_cap, c.st)) < 0) return ORBGPU_ERR_CUDA; total += n;
    if (prof) CU(cudaEventRecord(c.ev[EV_OCT], c.st));
    if (side) CU(cudaStreamWaitEvent(c.st, c.ev_join, 0));
    else { if ((n = orb_launch_blur(P, nframes, pl.d_tiles, pl.ntiles, c.st)) < 0) return ORBGPU_ERR_CUDA; total += n; }
    if (prof) CU(cudaEventRecord(c.ev[EV_BLUR], c.st));
    if ((n = orb_launch_describe(P, nframes, c.st)) < 0) return ORBGPU_ERR_CUDA; total += n;
    if (prof) CU(cudaEventRecord(c.ev[EV_DESC], c.st));
    CU(cudaGetLastError());
    e->launches += (uint64_t)total;
    return ORBGPU_OK;
}

// The same chunk through a captured graph (host-buffer path: every pointer the kernels see belongs to the context, so the
// graph only depends on the frame count).  Captured on first use, replayed afterwards; profiling mode and handles with
// graphs switched off take the direct path.
static int enqueue_kernels_graph(orbgpu_extractor* e, StreamCtx& c, int nframes)
{
    const Plan& pl = e->plan;
    const uint8_t* src = c.d_in;
    const long long fs = (long long)(pl.in_pitch * pl.h);
    if (e->profiling || !e->use_graphs)
        return enqueue_kernels(e, c, src, fs, (int)pl.in_pitch, nframes, c.d_kps, c.d_desc, c.d_nout, pl.kmax);
    int slot = -1, victim = 0;
    for (int k = 0; k < StreamCtx::NGRAPH; ++k) {
        if (c.gexec[k] && c.g_frames[k] == nframes && c.g_epoch[k] == e->epoch) { slot = k; break; }
        if (!c.gexec[k] || c.g_epoch[k] != e->epoch) victim = k;               // prefer an empty or stale slot
        else if (c.gexec[victim] && c.g_epoch[victim] == e->epoch && c.g_used[k] < c.g_used[victim]) victim = k;
    }
    if (slot < 0) {
        slot = victim;
        if (c.gexec[slot]) { cudaGraphExecDestroy(c.gexec[slot]); c.gexec[slot] = nullptr; }
        const uint64_t l0 = e->launches;
        CU(cudaStreamBeginCapture(c.st, cudaStreamCaptureModeRelaxed));
        const int rc = enqueue_kernels(e, c, src, fs, (int)pl.in_pitch, nframes, c.d_kps, c.d_desc, c.d_nout, pl.kmax);
        cudaGraph_t g = nullptr;
        const cudaError_t ce = cudaStreamEndCapture(c.st, &g);
        c.g_launches[slot] = (int)(e->launches - l0);
        e->launches = l0;
        if (rc != ORBGPU_OK) { if (g) cudaGraphDestroy(g); return rc; }
        CU(ce);
        const cudaError_t ci = cudaGraphInstantiate(&c.gexec[slot], g, 0);
        cudaGraphDestroy(g);
        CU(ci);
        c.g_frames[slot] = nframes; c.g_epoch[slot] = e->epoch;
    }
    c.g_used[slot] = ++c.g_clock;
    CU(cudaGraphLaunch(c.gexec[slot], c.st));
    e->launches += (uint64_t)c.g_launches[slot];
    return ORBGPU_OK;
}

static bool is_pinned(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

// wait for the context's pending job and hand its results to the caller
static int ctx_finish(orbgpu_extractor* e, StreamCtx& c)
{
    if (!c.busy) return ORBGPU_OK;
    CU(cudaEventSynchronize(c.done));
    c.busy = false;
    int rc = ORBGPU_OK;
    const int kmax = e->plan.kmax;
    const int* nout = c.staged_out ? c.h_nout : c.user_nout;
    if (c.staged_out && c.user_nout) {
        for (int i = 0; i < c.job_frames; ++i) {
            const int n = c.h_nout[i];
            c.user_nout[i] = n;
            if (n > 0) {
                memcpy(c.user_kps + (size_t)i * c.user_capacity, c.h_kps + (size_t)i * kmax, sizeof(orbgpu_keypoint) * (size_t)n);
                memcpy(c.user_desc + (size_t)i * c.user_capacity * 32, c.h_desc + (size_t)i * kmax * 32, 32 * (size_t)n);
            }
        }
    }
    if (nout)
        for (int i = 0; i < c.job_frames; ++i) if (nout[i] < 0) rc = ORBGPU_ERR_CAPACITY;
    if (c.profiled) {
        for (int s = 0; s < ORBGPU_NUM_STAGES; ++s) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, c.ev[s], c.ev[s + 1]) == cudaSuccess) e->stage_ms[s] += ms;
            else cudaGetLastError();
        }
        c.profiled = false;
    }
    return rc;
}

static int use_device(orbgpu_extractor* e)
{
    CU(cudaSetDevice(e->device));
    return ORBGPU_OK;
}

static int reserve_raw(StreamCtx& c, size_t bytes)
{
    if (c.raw_bytes >= bytes) return ORBGPU_OK;
    cudaFree(c.d_raw); c.d_raw = nullptr; c.raw_bytes = 0;
    CU(cudaMalloc(&c.d_raw, bytes + 64));
    c.raw_bytes = bytes;
    return ORBGPU_OK;
}

// upload `m` frames starting at host pointer `src` into c.d_in (row pitch plan.in_pitch, frames in_pitch * h apart).
// PCIe copies are always plain 1-D copies of dense frames; when the device layout differs from the dense one
// (w not a multiple of 16, or 3/4-channel input) the frames land in d_raw and k_repitch / k_cvt_gray lays them out.
static int upload_frames(orbgpu_extractor* e, StreamCtx& c, const uint8_t* src, int m, int w, int h,
                         size_t row_stride, size_t frame_stride, bool pinned)
{
    const int bpp = e->in_channels;
    const size_t rb = (size_t)w * bpp, fb = rb * h, dp = e->plan.in_pitch;
    const bool relayout = dp != (size_t)w || bpp != 1;
    int rc;
    if (relayout && (rc = reserve_raw(c, fb * m)) != ORBGPU_OK) return rc;
    uint8_t* dense = relayout ? c.d_raw : c.d_in;
    if (pinned && row_stride == rb && frame_stride == fb) {
        CU(cudaMemcpyAsync(dense, src, fb * m, cudaMemcpyHostToDevice, c.st));
    } else if (pinned && row_stride == rb) {
        for (int i = 0; i < m; ++i)
            CU(cudaMemcpyAsync(dense + (size_t)i * fb, src + (size_t)i * frame_stride, fb, cudaMemcpyHostToDevice, c.st));
    } else {                                               // strided rows or pageable memory: gather into the pinned staging buffer
        for (int i = 0; i < m; ++i) {
            const uint8_t* f = src + (size_t)i * frame_stride;
            uint8_t* d = c.h_in + (size_t)i * fb;
            if (row_stride == rb) memcpy(d, f, fb);
            else for (int y = 0; y < h; ++y) memcpy(d + (size_t)y * rb, f + (size_t)y * row_stride, rb);
        }
        CU(cudaMemcpyAsync(dense, c.h_in, fb * m, cudaMemcpyHostToDevice, c.st));
    }
    if (relayout) {
        const int n = bpp == 1 ? orb_launch_repitch(c.d_raw, (long long)fb, w, c.d_in, (int)dp, w, h, m, c.st)
                               : orb_launch_cvt_gray(c.d_raw, (long long)fb, (int)rb, bpp, e->in_rgb, c.d_in, (int)dp, w, h, m, c.st);
        if (n < 0) return ORBGPU_ERR_CUDA;
        e->launches += 1;
    }
    return ORBGPU_OK;
}

static int download_results(orbgpu_extractor* e, StreamCtx& c, int m, orbgpu_keypoint* kps, uint8_t* desc, int* n_out,
                            int capacity, bool pinned_out)
{
    const size_t kmax = (size_t)e->plan.kmax;
    c.staged_out = !pinned_out;
    c.user_kps = kps; c.user_desc = desc; c.user_nout = n_out; c.user_capacity = capacity;
    if (pinned_out) {
        CU(cudaMemcpy2DAsync(kps, (size_t)capacity * sizeof(orbgpu_keypoint), c.d_kps, kmax * sizeof(orbgpu_keypoint),
                             kmax * sizeof(orbgpu_keypoint), (size_t)m, cudaMemcpyDeviceToHost, c.st));
        CU(cudaMemcpy2DAsync(desc, (size_t)capacity * 32, c.d_desc, kmax * 32, kmax * 32, (size_t)m, cudaMemcpyDeviceToHost, c.st));
        CU(cudaMemcpyAsync(n_out, c.d_nout, sizeof(int) * (size_t)m, cudaMemcpyDeviceToHost, c.st));
    } else {
        CU(cudaMemcpyAsync(c.h_kps, c.d_kps, kmax * sizeof(orbgpu_keypoint) * m, cudaMemcpyDeviceToHost, c.st));
        CU(cudaMemcpyAsync(c.h_desc, c.d_desc, kmax * 32 * m, cudaMemcpyDeviceToHost, c.st));
        CU(cudaMemcpyAsync(c.h_nout, c.d_nout, sizeof(int) * (size_t)m, cudaMemcpyDeviceToHost, c.st));
    }
    return ORBGPU_OK;
}

// one chunk, host buffers: upload -> kernels -> download, all asynchronous on the context's stream
struct UndistortReq { const float* K4; const float* dist; int ndist; };   // orbgpu_extract_undistorted: one more kernel + copy per frame
extern "C" { static int reserve_aux(orbgpu_extractor* e, size_t bytes); }   // defined with the frame-side entry points below

static int submit_chunk(orbgpu_extractor* e, StreamCtx& c, const uint8_t* src, int m, int w, int h, size_t row_stride,
                        size_t frame_stride, bool pinned_in, orbgpu_keypoint* kps, uint8_t* desc, int* n_out,
                        int capacity, bool pinned_out, const UndistortReq* und = nullptr)
{
    int rc;
    if ((rc = ctx_finish(e, c)) != ORBGPU_OK) return rc;
    if ((rc = ctx_reserve(e, c, m, w, h)) != ORBGPU_OK) return rc;
    const bool prof = e->profiling;
    if (prof) CU(cudaEventRecord(c.ev[EV_BEGIN], c.st));
    if ((rc = upload_frames(e, c, src, m, w, h, row_stride, frame_stride, pinned_in)) != ORBGPU_OK) return rc;
    if (prof) CU(cudaEventRecord(c.ev[EV_UP], c.st));
    if ((rc = enqueue_kernels_graph(e, c, m)) != ORBGPU_OK) return rc;
    if (und) {                                                  // mvKeysUn beside mvKeys, behind the same synchronisation (m == 1)
        orb_launch_undistort_keypoints(und->K4, und->dist, und->ndist, c.d_kps, e->d_aux, c.d_nout, e->plan.kmax, c.st);
        e->launches += 1;
    }
    if ((rc = download_results(e, c, m, kps, desc, n_out, capacity, pinned_out)) != ORBGPU_OK) return rc;
    if (und) CU(cudaMemcpyAsync(e->h_aux, e->d_aux, sizeof(orbgpu_keypoint) * (size_t)e->plan.kmax, cudaMemcpyDeviceToHost, c.st));
    if (prof) CU(cudaEventRecord(c.ev[EV_DOWN], c.st));
    CU(cudaEventRecord(c.done, c.st));
    c.busy = true; c.profiled = prof; c.job_frames = m;
    return ORBGPU_OK;
}

// Device-side bracket of one batch call: begin is recorded on ctx[0].st before any chunk is enqueued
// (all streams are idle between calls); end is recorded on ctx[0].st after it has been made to wait
// for the last chunk of every other stream.
static int batch_begin(orbgpu_extractor* e)
{
    CU(cudaEventRecord(e->ev_batch0, e->ctx[0].st));
    return ORBGPU_OK;
}
static int batch_end(orbgpu_extractor* e)
{
    for (int s = 1; s < NSTREAM; ++s)
        if (e->ctx[s].busy) CU(cudaStreamWaitEvent(e->ctx[0].st, e->ctx[s].done, 0));
    CU(cudaEventRecord(e->ev_batch1, e->ctx[0].st));
    return ORBGPU_OK;
}
static void batch_collect_ms(orbgpu_extractor* e)
{
    float ms = 0.f;
    if (cudaEventSynchronize(e->ev_batch1) == cudaSuccess && cudaEventElapsedTime(&ms, e->ev_batch0, e->ev_batch1) == cudaSuccess)
        e->last_batch_ms = ms;
    else { cudaGetLastError(); e->last_batch_ms = 0; }
}

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

int orbgpu_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* orbgpu_version(void) { return "orbgpu 0.1 (sm_100a)"; }

const char* orbgpu_strerror(int code)
{
    switch (code) {
        case ORBGPU_OK: return "ok";
        case ORBGPU_ERR_ARG: return "invalid argument";
        case ORBGPU_ERR_CAPACITY: return "output capacity too small";
        case ORBGPU_ERR_CUDA: return "CUDA error (see orbgpu_last_cuda_error)";
        case ORBGPU_ERR_NOMEM: return "out of memory";
        case ORBGPU_ERR_UNSUPPORTED: return "unsupported image shape or parameters";
        case ORBGPU_ERR_STATE: return "invalid state (ticket, queue or no frame extracted yet)";
        default: return "unknown error";
    }
}

int orbgpu_create(orbgpu_extractor** out, int device, int nfeatures, float scale_factor, int nlevels,
                  int ini_th_fast, int min_th_fast)
{
    if (!out) return ORBGPU_ERR_ARG;
    *out = nullptr;
    if (nfeatures < 1 || nlevels < 1 || nlevels > ORB_MAX_LEVELS || !(scale_factor > 1.0f)) return ORBGPU_ERR_ARG;
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return ORBGPU_ERR_ARG;
    CU(cudaSetDevice(device));
    orbgpu_extractor* e = new (std::nothrow) orbgpu_extractor();
    if (!e) return ORBGPU_ERR_NOMEM;
    e->device = device; e->nfeatures = nfeatures; e->nlevels = nlevels;
    e->ini_th = std::min(std::max(ini_th_fast, 0), 255);     // cv::FAST clamps its threshold
    e->min_th = std::min(std::max(min_th_fast, 0), 255);
    build_tables(e, scale_factor);
    if (orb_upload_constants(e->umax) != 0) { delete e; return ORBGPU_ERR_CUDA; }
    for (int s = 0; s < NSTREAM; ++s) {
        StreamCtx& c = e->ctx[s];
        if (cudaStreamCreateWithFlags(&c.st, cudaStreamNonBlocking) != cudaSuccess) { delete e; return ORBGPU_ERR_CUDA; }
        if (cudaStreamCreateWithFlags(&c.st_side, cudaStreamNonBlocking) != cudaSuccess) { delete e; return ORBGPU_ERR_CUDA; }
        cudaEventCreateWithFlags(&c.ev_fork, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&c.ev_join, cudaEventDisableTiming);
        for (int k = 0; k < EV_COUNT; ++k) cudaEventCreate(&c.ev[k]);
        cudaEventCreateWithFlags(&c.done, cudaEventDisableTiming);
    }
    cudaEventCreate(&e->ev_batch0); cudaEventCreate(&e->ev_batch1);
    *out = e;
    return ORBGPU_OK;
}

void orbgpu_destroy(orbgpu_extractor* e)
{
    if (!e) return;
    cudaSetDevice(e->device);
    for (int s = 0; s < NSTREAM; ++s) {
        StreamCtx& c = e->ctx[s];
        if (c.st) cudaStreamSynchronize(c.st);
        ctx_free_buffers(c);
        for (int k = 0; k < EV_COUNT; ++k) if (c.ev[k]) cudaEventDestroy(c.ev[k]);
        if (c.done) cudaEventDestroy(c.done);
        if (c.ev_fork) cudaEventDestroy(c.ev_fork);
        if (c.ev_join) cudaEventDestroy(c.ev_join);
        if (c.st_side) cudaStreamDestroy(c.st_side);
        if (c.st) cudaStreamDestroy(c.st);
    }
    if (e->ev_batch0) cudaEventDestroy(e->ev_batch0);
    if (e->ev_batch1) cudaEventDestroy(e->ev_batch1);
    e->plan.free_dev();
    cudaFree(e->d_oct_scratch);
    cudaFree(e->d_aux);
    cudaFreeHost(e->h_aux);
    delete e;
}

int orbgpu_get_tables(orbgpu_extractor* e, float* scale4, int* feats, int* umax)
{
    if (!e) return ORBGPU_ERR_ARG;
    const int nl = e->nlevels;
    if (scale4)
        for (int i = 0; i < nl; ++i) {
            scale4[i] = e->scale[i]; scale4[nl + i] = e->inv_scale[i];
            scale4[2 * nl + i] = e->sigma2[i]; scale4[3 * nl + i] = e->inv_sigma2[i];
        }
    if (feats) for (int i = 0; i < nl; ++i) feats[i] = e->feats[i];
    if (umax) for (int i = 0; i < 16; ++i) umax[i] = e->umax[i];
    return nl;
}

int orbgpu_max_keypoints(orbgpu_extractor* e, int w, int h)
{
    if (!e) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    if ((rc = build_plan(e, w, h)) != ORBGPU_OK) return rc;
    return e->plan.kmax;
}

// enqueue one host-buffer batch on the chunk streams; with `rotate` the chunk -> stream assignment continues
// where the previous batch stopped, so back-to-back asynchronous batches keep every stream fed
static int enqueue_host_batch(orbgpu_extractor* e, const uint8_t* frames, int n, int w, int h, size_t row_stride,
                              size_t frame_stride, orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out)
{
    int rc;
    const bool pin_in = is_pinned(frames);
    const bool pin_out = is_pinned(kps) && is_pinned(desc) && is_pinned(n_out);
    const int chunk = std::max(1, std::min(std::min(e->plan.chunk_cap, e->max_chunk[0]), (n + e->nstream[0] - 1) / e->nstream[0]));
    int result = ORBGPU_OK;
    for (int i0 = 0; i0 < n; i0 += chunk) {
        StreamCtx& c = e->ctx[e->next_ctx % e->nstream[0]];
        e->next_ctx = (e->next_ctx + 1) % e->nstream[0];
        const int m = std::min(chunk, n - i0);
        if ((rc = ctx_finish(e, c)) != ORBGPU_OK) result = rc;
        if ((rc = ctx_reserve(e, c, chunk, w, h)) != ORBGPU_OK) { result = rc; break; }
        rc = submit_chunk(e, c, frames + (size_t)i0 * frame_stride, m, w, h, row_stride, frame_stride, pin_in,
                          kps + (size_t)i0 * capacity, desc + (size_t)i0 * capacity * 32, n_out + i0, capacity, pin_out);
        if (rc != ORBGPU_OK) { result = rc; break; }
    }
    return result;
}

static int finish_all(orbgpu_extractor* e, int result)
{
    int rc;
    if (result == ORBGPU_OK) result = batch_end(e);
    for (int s = 0; s < NSTREAM; ++s)
        if ((rc = ctx_finish(e, e->ctx[s])) != ORBGPU_OK) result = rc;
    batch_collect_ms(e);
    e->async_open = false;
    return result;
}

int orbgpu_extract_batch(orbgpu_extractor* e, const uint8_t* frames, int n, int w, int h, size_t row_stride,
                         size_t frame_stride, orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out)
{
    if (!e || !frames || !kps || !desc || !n_out || n < 1 || w < 1 || h < 1 || row_stride < (size_t)w * e->in_channels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    if (e->async_open) return ORBGPU_ERR_STATE;                // orbgpu_wait first
    for (int s = 0; s < NSTREAM; ++s) if (e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;
    if ((rc = build_plan(e, w, h)) != ORBGPU_OK) return rc;
    if (capacity < e->plan.kmax) return ORBGPU_ERR_CAPACITY;
    if ((rc = batch_begin(e)) != ORBGPU_OK) return rc;
    e->next_ctx = 0;
    const int result = finish_all(e, enqueue_host_batch(e, frames, n, w, h, row_stride, frame_stride, kps, desc, capacity, n_out));
    e->last_ctx = (n == 1 && result == ORBGPU_OK) ? 0 : -1;
    return result;
}

int orbgpu_extract_batch_async(orbgpu_extractor* e, const uint8_t* frames, int n, int w, int h, size_t row_stride,
                               size_t frame_stride, orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out)
{
    if (!e || !frames || !kps || !desc || !n_out || n < 1 || w < 1 || h < 1 || row_stride < (size_t)w * e->in_channels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    for (int s = 0; s < NSTREAM; ++s) if (e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;
    if (e->async_open && (w != e->plan.w || h != e->plan.h)) return ORBGPU_ERR_STATE;   // one shape per open window
    if ((rc = build_plan(e, w, h)) != ORBGPU_OK) return rc;
    if (capacity < e->plan.kmax) return ORBGPU_ERR_CAPACITY;
    if (!e->async_open) {
        if ((rc = batch_begin(e)) != ORBGPU_OK) return rc;
        e->async_open = true;
        e->next_ctx = 0;
    }
    e->last_ctx = -1;
    return enqueue_host_batch(e, frames, n, w, h, row_stride, frame_stride, kps, desc, capacity, n_out);
}

// One process, several GPUs: handle r (its own device) takes the r-th contiguous block of frames on its own host thread --
// the reference's one-extractor-per-thread pattern (src/Frame.cc:78-81) with one GPU per thread.  No data-path exchange.
int orbgpu_extract_batch_multi(orbgpu_extractor** handles, int num_handles, const uint8_t* frames, int n, int w, int h,
                               size_t row_stride, size_t frame_stride, orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out)
{
    if (!handles || num_handles < 1 || !frames || !kps || !desc || !n_out || n < 1) return ORBGPU_ERR_ARG;
    for (int r = 0; r < num_handles; ++r) if (!handles[r]) return ORBGPU_ERR_ARG;
    if (num_handles == 1) return orbgpu_extract_batch(handles[0], frames, n, w, h, row_stride, frame_stride, kps, desc, capacity, n_out);
    std::vector<int> rc((size_t)num_handles, ORBGPU_OK);
    std::vector<std::string> err((size_t)num_handles);
    std::vector<std::thread> th;
    for (int r = 0; r < num_handles; ++r) {
        const int i0 = (int)((long long)n * r / num_handles), i1 = (int)((long long)n * (r + 1) / num_handles);
        if (i1 <= i0) continue;
        th.emplace_back([=, &rc, &err]() {
            rc[r] = orbgpu_extract_batch(handles[r], frames + (size_t)i0 * frame_stride, i1 - i0, w, h, row_stride, frame_stride,
                                         kps + (size_t)i0 * capacity, desc + (size_t)i0 * capacity * 32, capacity, n_out + i0);
            if (rc[r] != ORBGPU_OK) err[r] = g_cuda_err;          // the error text is thread-local: carry it to the caller
        });
    }
    for (auto& t : th) t.join();
    for (int r = 0; r < num_handles; ++r)
        if (rc[r] != ORBGPU_OK) { orbgpu_internal_set_error(err[r].c_str()); return rc[r]; }
    return ORBGPU_OK;
}

int orbgpu_wait(orbgpu_extractor* e)
{
    if (!e) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    if (!e->async_open) return ORBGPU_OK;
    return finish_all(e, ORBGPU_OK);
}

int orbgpu_extract(orbgpu_extractor* e, const uint8_t* img, int w, int h, size_t stride,
                   orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out)
{
    if (!e || !n_out) return ORBGPU_ERR_ARG;
    if (!img || w == 0 || h == 0) { *n_out = -1; return ORBGPU_OK; }   // empty image: outputs untouched
    return orbgpu_extract_batch(e, img, 1, w, h, stride, stride * (size_t)h, kps, desc, capacity, n_out);
}

int orbgpu_extract_batch_device(orbgpu_extractor* e, const uint8_t* d_frames, int n, int w, int h, size_t row_stride,
                                size_t frame_stride, orbgpu_keypoint* d_kps, uint8_t* d_desc, int capacity, int* d_n_out)
{
    if (!e || !d_frames || !d_kps || !d_desc || !d_n_out || n < 1 || w < 1 || h < 1 || row_stride < (size_t)w * e->in_channels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    if (e->async_open) return ORBGPU_ERR_STATE;
    for (int s = 0; s < NSTREAM; ++s) if (e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;
    if ((rc = build_plan(e, w, h)) != ORBGPU_OK) return rc;
    if (capacity < e->plan.kmax) return ORBGPU_ERR_CAPACITY;
    const int chunk = std::max(1, std::min(std::min(e->plan.chunk_cap, e->max_chunk[1]), (n + e->nstream[1] - 1) / e->nstream[1]));
    int result = ORBGPU_OK, k = 0;
    if ((rc = batch_begin(e)) != ORBGPU_OK) return rc;
    for (int i0 = 0; i0 < n; i0 += chunk, ++k) {
        StreamCtx& c = e->ctx[k % e->nstream[1]];
        const int m = std::min(chunk, n - i0);
        if ((rc = ctx_finish(e, c)) != ORBGPU_OK) result = rc;
        if ((rc = ctx_reserve(e, c, chunk, w, h)) != ORBGPU_OK) { result = rc; break; }
        const bool prof = e->profiling;
        if (prof) { cudaEventRecord(c.ev[EV_BEGIN], c.st); cudaEventRecord(c.ev[EV_UP], c.st); }
        const uint8_t* lvl0 = d_frames + (size_t)i0 * frame_stride;
        long long lvl0_frame = (long long)frame_stride;
        int lvl0_pitch = (int)row_stride;
        if (e->in_channels != 1 || (row_stride & 3) || (frame_stride & 3) || (reinterpret_cast<uintptr_t>(d_frames) & 3)) {
            // colour input, or rows not 4-byte aligned (e.g. 1241-px KITTI frames, dense): one pass into the pitched
            // gray buffer keeps every kernel on its aligned word path
            const size_t dp = e->plan.in_pitch;
            const int nl = e->in_channels == 1
                               ? orb_launch_repitch(lvl0, (long long)frame_stride, (int)row_stride, c.d_in, (int)dp, w, h, m, c.st)
                               : orb_launch_cvt_gray(lvl0, (long long)frame_stride, (int)row_stride, e->in_channels, e->in_rgb, c.d_in,
                                                     (int)dp, w, h, m, c.st);
            if (nl < 0) { result = ORBGPU_ERR_CUDA; break; }      // earlier chunks are in flight: fall through to the common drain
            e->launches += 1;
            lvl0 = c.d_in; lvl0_frame = (long long)(dp * h); lvl0_pitch = (int)dp;
        }
        rc = enqueue_kernels(e, c, lvl0, lvl0_frame, lvl0_pitch, m,
                             d_kps + (size_t)i0 * capacity, d_desc + (size_t)i0 * capacity * 32, d_n_out + i0, capacity);
        if (rc != ORBGPU_OK) { result = rc; break; }
        if (prof) cudaEventRecord(c.ev[EV_DOWN], c.st);
        if (cudaEventRecord(c.done, c.st) != cudaSuccess) { result = ORBGPU_ERR_CUDA; break; }
        c.busy = true; c.profiled = prof; c.job_frames = m; c.staged_out = false; c.user_nout = nullptr;
    }
    if (result == ORBGPU_OK) result = batch_end(e);
    for (int s = 0; s < NSTREAM; ++s)
        if ((rc = ctx_finish(e, e->ctx[s])) != ORBGPU_OK) result = rc;
    batch_collect_ms(e);
    e->last_ctx = -1;
    return result;
}

// ---- pinned frame queue ---------------------------------------------------------------------------
int orbgpu_queue_depth(orbgpu_extractor* e) { return e ? NSTREAM : ORBGPU_ERR_ARG; }

int orbgpu_submit(orbgpu_extractor* e, const uint8_t* img, int w, int h, size_t stride, uint64_t* ticket)
{
    if (!e || !img || !ticket || w < 1 || h < 1 || stride < (size_t)w * e->in_channels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    if (e->async_open) return ORBGPU_ERR_STATE;
    if ((rc = build_plan(e, w, h)) != ORBGPU_OK) return rc;
    const uint64_t t = e->next_ticket;
    StreamCtx& c = e->ctx[t % NSTREAM];
    if (c.ticket_live) return ORBGPU_ERR_STATE;          // ring slot not collected yet
    if ((rc = ctx_reserve(e, c, 1, w, h)) != ORBGPU_OK) return rc;
    // always staged: the caller's image may be reused as soon as submit returns
    rc = submit_chunk(e, c, img, 1, w, h, stride, stride * (size_t)h, false, nullptr, nullptr, nullptr, 0, false);
    if (rc != ORBGPU_OK) return rc;
    c.ticket = t; c.ticket_live = true;
    e->next_ticket = t + 1;
    *ticket = t;
    return ORBGPU_OK;
}

int orbgpu_collect(orbgpu_extractor* e, uint64_t ticket, orbgpu_keypoint* kps, uint8_t* desc, int capacity, int* n_out)
{
    if (!e || !kps || !desc || !n_out) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    StreamCtx& c = e->ctx[ticket % NSTREAM];
    if (!c.ticket_live || c.ticket != ticket) return ORBGPU_ERR_STATE;
    if (capacity < e->plan.kmax) return ORBGPU_ERR_CAPACITY;
    rc = ctx_finish(e, c);                                // staged, user pointers null: nothing copied yet
    c.ticket_live = false;
    if (rc != ORBGPU_OK) return rc;
    const int n = c.h_nout[0];
    if (n < 0) return ORBGPU_ERR_CAPACITY;
    *n_out = n;
    memcpy(kps, c.h_kps, sizeof(orbgpu_keypoint) * (size_t)n);
    memcpy(desc, c.h_desc, 32 * (size_t)n);
    e->last_ctx = (int)(ticket % NSTREAM);
    return ORBGPU_OK;
}

// ---- views of the most recent single frame ------------------------------------------------------
static StreamCtx* last_frame_ctx(orbgpu_extractor* e)
{
    if (!e || e->last_ctx < 0) return nullptr;
    StreamCtx& c = e->ctx[e->last_ctx];
    return c.busy ? nullptr : &c;
}

int orbgpu_pyramid_level(orbgpu_extractor* e, int level, uint8_t* dst, size_t dst_stride, int* w, int* h)
{
    StreamCtx* c = last_frame_ctx(e);
    if (!c) return ORBGPU_ERR_STATE;
    if (level < 0 || level >= e->nlevels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    const OrbLevel& L = e->plan.dev.lv[level];
    if (w) *w = L.w;
    if (h) *h = L.h;
    if (!dst) return ORBGPU_OK;
    if (dst_stride < (size_t)L.w) return ORBGPU_ERR_ARG;
    if (level == 0) CU(cudaMemcpy2D(dst, dst_stride, c->d_in, e->plan.in_pitch, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost));
    else CU(cudaMemcpy2D(dst, dst_stride, c->d_pyr + L.pyr_off, (size_t)L.pitch, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost));
    return ORBGPU_OK;
}

int orbgpu_debug_level_blurred(orbgpu_extractor* e, int level, uint8_t* dst, size_t dst_stride)
{
    StreamCtx* c = last_frame_ctx(e);
    if (!c) return ORBGPU_ERR_STATE;
    if (level < 0 || level >= e->nlevels || !dst) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    const OrbLevel& L = e->plan.dev.lv[level];
    CU(cudaMemcpy2D(dst, dst_stride, c->d_blur + L.blur_off, (size_t)L.pitch, (size_t)L.w, (size_t)L.h, cudaMemcpyDeviceToHost));
    return ORBGPU_OK;
}

int orbgpu_debug_level_candidates(orbgpu_extractor* e, int level, int* x, int* y, int* score, int cap)
{
    StreamCtx* c = last_frame_ctx(e);
    if (!c) return ORBGPU_ERR_STATE;
    if (level < 0 || level >= e->nlevels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    const Plan& pl = e->plan;
    const OrbLevel& L = pl.dev.lv[level];
    if (L.ncells == 0) return 0;
    int total = 0;
    CU(cudaMemcpy(&total, c->d_candcnt + level, sizeof(int), cudaMemcpyDeviceToHost));
    std::vector<int> base((size_t)L.ncells);          // exclusive scan left in place by the octree kernel
    CU(cudaMemcpy(base.data(), c->d_cellcnt + L.cell0, sizeof(int) * (size_t)L.ncells, cudaMemcpyDeviceToHost));
    std::vector<uint32_t> slots((size_t)L.slot_total);
    CU(cudaMemcpy(slots.data(), c->d_slots + L.slot_off, sizeof(uint32_t) * (size_t)L.slot_total, cudaMemcpyDeviceToHost));
    int o = 0;
    for (int ci = 0; ci < L.ncells; ++ci) {
        const int cnt = (ci + 1 < L.ncells ? base[ci + 1] : total) - base[ci];
        const uint32_t* s = &slots[(size_t)pl.cells[L.cell0 + ci].slot_off];
        for (int i = 0; i < cnt; ++i, ++o)
            if (o < cap) { x[o] = ORB_PX(s[i]); y[o] = ORB_PY(s[i]); score[o] = ORB_PS(s[i]); }
    }
    return o;
}

int orbgpu_debug_level_selected(orbgpu_extractor* e, int level, int* x, int* y, int* score, int cap)
{
    StreamCtx* c = last_frame_ctx(e);
    if (!c) return ORBGPU_ERR_STATE;
    if (level < 0 || level >= e->nlevels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    const OrbLevel& L = e->plan.dev.lv[level];
    int n = 0;
    CU(cudaMemcpy(&n, c->d_selcnt + level, sizeof(int), cudaMemcpyDeviceToHost));
    if (n <= 0) return n;
    std::vector<uint32_t> sel((size_t)n);
    CU(cudaMemcpy(sel.data(), c->d_sel + L.sel_off, sizeof(uint32_t) * (size_t)n, cudaMemcpyDeviceToHost));
    for (int i = 0; i < n && i < cap; ++i) { x[i] = ORB_PX(sel[i]); y[i] = ORB_PY(sel[i]); score[i] = ORB_PS(sel[i]); }
    return n;
}

int orbgpu_octree(orbgpu_extractor* e, const int* xs, const int* ys, const int* resp, int n, int W, int H, int N,
                  int* out_idx, int cap)
{
    if (!e || n < 0 || (n && (!xs || !ys || !resp)) || !out_idx || W < 1 || H < 1 || W > 4096 || H > 4096 || N < 1) return ORBGPU_ERR_ARG;
    if (n >= (1 << 24)) return ORBGPU_ERR_UNSUPPORTED;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    const int n_ini = (int)roundf((float)W / (float)H);
    if (n_ini < 1) return ORBGPU_ERR_UNSUPPORTED;
    const float hx = (float)W / n_ini;
    const int node_cap = std::max(N, 4 * n_ini) + 4;
    if (node_cap > 65535) return ORBGPU_ERR_UNSUPPORTED;
    std::vector<uint32_t> keys((size_t)std::max(n, 1));
    for (int i = 0; i < n; ++i) {
        if (xs[i] < 0 || xs[i] >= W || ys[i] < 0 || ys[i] > 4095 || resp[i] < 0 || resp[i] > 255) return ORBGPU_ERR_ARG;
        keys[i] = ORB_PACK(xs[i], ys[i], resp[i]);
    }
    const size_t key_bytes = align_up((size_t)7 * std::max(n, 1), 16), node_bytes = (size_t)64 * node_cap + 64;
    const size_t need = key_bytes + node_bytes + 4 * (size_t)std::max(n, 1) + 4 * (size_t)node_cap + 64;
    if (need > e->oct_scratch_bytes) {
        cudaFree(e->d_oct_scratch); e->d_oct_scratch = nullptr; e->oct_scratch_bytes = 0;
        CU(cudaMalloc(&e->d_oct_scratch, need));
        e->oct_scratch_bytes = need;
    }
    uint8_t* base = e->d_oct_scratch;
    uint32_t* d_keys = reinterpret_cast<uint32_t*>(base + key_bytes + node_bytes);
    uint32_t* d_out = d_keys + std::max(n, 1);
    int* d_count = reinterpret_cast<int*>(d_out + node_cap);
    cudaStream_t st = e->ctx[0].st;
    CU(cudaMemcpyAsync(d_keys, keys.data(), 4 * (size_t)std::max(n, 1), cudaMemcpyHostToDevice, st));
    const int key_smem_cap = std::min(std::max(8 * N, 1024), 6144), node_smem_cap = std::min(node_cap, 1024);
    if (orb_launch_octree_single(d_keys, n, W, H, N, n_ini, hx, node_cap, base, d_out, d_count, key_smem_cap,
                                 node_smem_cap, st) < 0) return ORBGPU_ERR_CUDA;
    e->launches += 1;
    int count = 0;
    CU(cudaMemcpyAsync(&count, d_count, sizeof(int), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (count < 0) return ORBGPU_ERR_CAPACITY;
    std::vector<uint32_t> out((size_t)std::max(count, 1));
    CU(cudaMemcpy(out.data(), d_out, 4 * (size_t)count, cudaMemcpyDeviceToHost));
    // the kernel returns packed keys; map them back to input indices (first occurrence = the key
    // that wins response ties, since equal packed values are the same candidate)
    std::vector<int> first_index;
    {
        std::vector<std::pair<uint32_t, int>> order((size_t)n);
        for (int i = 0; i < n; ++i) order[i] = std::make_pair(keys[i], i);
        std::sort(order.begin(), order.end());
        for (int i = 0; i < count; ++i) {
            auto it = std::lower_bound(order.begin(), order.end(), std::make_pair(out[i], -1));
            if (it == order.end() || it->first != out[i]) return ORBGPU_ERR_STATE;
            if (i < cap) out_idx[i] = it->second;
        }
    }
    return count;
}

// ---- Frame::ComputeStereoMatches (src/Frame.cc:466-640) --------------------------------------------------------
static int reserve_aux(orbgpu_extractor* e, size_t bytes)
{
    if (bytes > e->d_aux_cap) {
        cudaFree(e->d_aux); e->d_aux = nullptr; e->d_aux_cap = 0;
        CU(cudaMalloc(&e->d_aux, bytes));
        e->d_aux_cap = bytes;
    }
    if (bytes > e->h_aux_cap) {
        cudaFreeHost(e->h_aux); e->h_aux = nullptr; e->h_aux_cap = 0;
        CU(cudaMallocHost(&e->h_aux, bytes));
        e->h_aux_cap = bytes;
    }
    return ORBGPU_OK;
}

static int stereo_side(orbgpu_extractor* e, OrbStereoSide& S)
{
    StreamCtx* c = last_frame_ctx(e);
    if (!c) return ORBGPU_ERR_STATE;
    const Plan& pl = e->plan;
    for (int l = 0; l < e->nlevels; ++l) {
        const OrbLevel& L = pl.dev.lv[l];
        S.lvl[l] = l == 0 ? c->d_in : c->d_pyr + L.pyr_off;
        S.pitch[l] = l == 0 ? (int)pl.in_pitch : L.pitch;
        S.w[l] = L.w; S.h[l] = L.h;
    }
    S.kps = c->d_kps; S.desc = c->d_desc;
    S.n_ptr = c->d_nout;                                       // count of the frame this context extracted last
    return ORBGPU_OK;
}

int orbgpu_stereo_match(orbgpu_extractor* left, orbgpu_extractor* right, float bf, float b, float* u_right, float* depth,
                        int capacity, int* n_left)
{
    if (!left || !right || !u_right || !depth || !n_left || !(b > 0.f)) return ORBGPU_ERR_ARG;
    if (left->device != right->device || left->nlevels != right->nlevels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(left)) != ORBGPU_OK) return rc;
    OrbStereoArgs A;
    memset(&A, 0, sizeof(A));
    if ((rc = stereo_side(left, A.L)) != ORBGPU_OK || (rc = stereo_side(right, A.R)) != ORBGPU_OK) return rc;
    for (int l = 0; l < left->nlevels; ++l)
        if (A.L.w[l] != A.R.w[l] || A.L.h[l] != A.R.h[l] || left->scale[l] != right->scale[l]) return ORBGPU_ERR_ARG;
    const int kmax = left->plan.kmax;
    if (right->plan.kmax >= 65535) return ORBGPU_ERR_UNSUPPORTED;   // right index is packed into 16 bits
    A.nlevels = left->nlevels;
    for (int l = 0; l < left->nlevels; ++l) { A.scale[l] = left->scale[l]; A.inv_scale[l] = left->inv_scale[l]; }
    A.bf = bf; A.b = b;
    A.th_high = 100; A.th_orb_dist = (100 + 50) / 2;           // ORBmatcher::TH_HIGH, TH_LOW (src/ORBmatcher.cc:37-38)
    const size_t bytes = sizeof(float) * (3 * (size_t)kmax + 2);
    if ((rc = reserve_aux(left, bytes)) != ORBGPU_OK) return rc;
    float* d_out = reinterpret_cast<float*>(left->d_aux);
    A.u_right = d_out; A.depth = d_out + kmax; A.sad = reinterpret_cast<int*>(d_out + 2 * (size_t)kmax);
    A.counts = reinterpret_cast<int*>(d_out + 3 * (size_t)kmax);
    cudaStream_t st = left->ctx[0].st;
    // the right handle's work was synchronised by its own extract call; same device, so a plain launch suffices
    if (orb_launch_stereo(A, kmax, st) < 0) return ORBGPU_ERR_CUDA;
    left->launches += 1;
    CU(cudaMemcpyAsync(left->h_aux, left->d_aux, bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const float* h = reinterpret_cast<const float*>(left->h_aux);
    const int* sad = reinterpret_cast<const int*>(h + 2 * (size_t)kmax);
    const int N = reinterpret_cast<const int*>(h + 3 * (size_t)kmax)[0];
    if (N < 0) return ORBGPU_ERR_STATE;
    *n_left = N;
    if (capacity < N) return ORBGPU_ERR_CAPACITY;
    memcpy(u_right, h, sizeof(float) * (size_t)N);
    memcpy(depth, h + kmax, sizeof(float) * (size_t)N);
    // outlier rejection by the median SAD (:626-639): accepted matches sorted by (SAD, left index)
    std::vector<std::pair<int, int> > v;
    for (int i = 0; i < N; ++i) if (sad[i] >= 0) v.push_back(std::make_pair(sad[i], i));
    if (!v.empty()) {                                          // (the reference indexes an empty vector here when nothing matched)
        std::sort(v.begin(), v.end());
        const float median = (float)v[v.size() / 2].first;
        const float thDist = 1.5f * 1.4f * median;
        for (int i = (int)v.size() - 1; i >= 0; --i) {
            if ((float)v[i].first < thDist) break;
            u_right[v[i].second] = -1.f;
            depth[v[i].second] = -1.f;
        }
    }
    return ORBGPU_OK;
}

// ---- Frame::UndistortKeyPoints / ComputeImageBounds (src/Frame.cc:404-463) -------------------------
int orbgpu_undistort_keypoints(orbgpu_extractor* e, const float* K4, const float* dist, int ndist, orbgpu_keypoint* kps_un,
                               int capacity, int* n_out)
{
    if (!e || !K4 || !dist || !kps_un || !n_out || ndist < 4 || ndist > 12) return ORBGPU_ERR_ARG;
    StreamCtx* c = last_frame_ctx(e);
    if (!c) return ORBGPU_ERR_STATE;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    const int kmax = e->plan.kmax;
    const size_t rec = sizeof(orbgpu_keypoint) * (size_t)kmax, bytes = rec + 16;
    if ((rc = reserve_aux(e, bytes)) != ORBGPU_OK) return rc;
    cudaStream_t st = c->st;
    orb_launch_undistort_keypoints(K4, dist, ndist, c->d_kps, e->d_aux, c->d_nout, kmax, st);
    e->launches += 1;
    CU(cudaMemcpyAsync(e->d_aux + rec, c->d_nout, sizeof(int), cudaMemcpyDeviceToDevice, st));
    CU(cudaMemcpyAsync(e->h_aux, e->d_aux, bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const int n = *reinterpret_cast<const int*>(e->h_aux + rec);
    if (n < 0) return ORBGPU_ERR_STATE;
    *n_out = n;
    if (capacity < n) return ORBGPU_ERR_CAPACITY;
    memcpy(kps_un, e->h_aux, sizeof(orbgpu_keypoint) * (size_t)n);
    return ORBGPU_OK;
}

// orbgpu_extract + orbgpu_undistort_keypoints behind one synchronisation: the Frame constructor's ExtractORB + UndistortKeyPoints
// (src/Frame.cc:118-126, :404-435) cost one device round trip instead of two.
int orbgpu_extract_undistorted(orbgpu_extractor* e, const uint8_t* img, int w, int h, size_t stride, const float* K4, const float* dist,
                               int ndist, orbgpu_keypoint* kps, orbgpu_keypoint* kps_un, uint8_t* desc, int capacity, int* n_out)
{
    if (!e || !n_out || !K4 || !dist || ndist < 4 || ndist > 12) return ORBGPU_ERR_ARG;
    if (!img || w == 0 || h == 0) { *n_out = -1; return ORBGPU_OK; }       // empty image: outputs untouched, as orbgpu_extract
    if (!kps || !kps_un || !desc || w < 1 || h < 1 || stride < (size_t)w * e->in_channels) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    if (e->async_open) return ORBGPU_ERR_STATE;
    for (int s = 0; s < NSTREAM; ++s) if (e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;
    if ((rc = build_plan(e, w, h)) != ORBGPU_OK) return rc;
    if (capacity < e->plan.kmax) return ORBGPU_ERR_CAPACITY;
    for (int s = 0; s < NSTREAM; ++s) if ((rc = ctx_finish(e, e->ctx[s])) != ORBGPU_OK) return rc;   // d_aux may be regrown below
    if ((rc = reserve_aux(e, sizeof(orbgpu_keypoint) * (size_t)e->plan.kmax + 16)) != ORBGPU_OK) return rc;
    if ((rc = batch_begin(e)) != ORBGPU_OK) return rc;
    e->next_ctx = 0;
    const UndistortReq und = {K4, dist, ndist};
    StreamCtx& c = e->ctx[0];
    rc = submit_chunk(e, c, img, 1, w, h, stride, stride * (size_t)h, is_pinned(img), kps, desc, n_out, capacity,
                      is_pinned(kps) && is_pinned(desc) && is_pinned(n_out), &und);
    const int result = finish_all(e, rc);
    e->last_ctx = result == ORBGPU_OK ? 0 : -1;
    if (result == ORBGPU_OK && *n_out > 0) memcpy(kps_un, e->h_aux, sizeof(orbgpu_keypoint) * (size_t)*n_out);
    return result;
}

int orbgpu_undistort_points(int device, const float* xy, int n, const float* K4, const float* dist, int ndist, float* out_xy)
{
    if (!xy || !K4 || !dist || !out_xy || n < 0 || ndist < 4 || ndist > 12) return ORBGPU_ERR_ARG;
    if (n == 0) return ORBGPU_OK;
    if (cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return ORBGPU_ERR_CUDA; }
    float* d = nullptr;
    if (cudaMalloc(&d, 4 * sizeof(float) * (size_t)n) != cudaSuccess) { cudaGetLastError(); return ORBGPU_ERR_CUDA; }
    cudaError_t ce = cudaMemcpy(d, xy, 2 * sizeof(float) * (size_t)n, cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) { orb_launch_undistort_points(K4, dist, ndist, d, d + 2 * (size_t)n, n, 0); ce = cudaGetLastError(); }
    if (ce == cudaSuccess) ce = cudaMemcpy(out_xy, d + 2 * (size_t)n, 2 * sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost);
    cudaFree(d);
    return ce == cudaSuccess ? ORBGPU_OK : ORBGPU_ERR_CUDA;
}

// ---- float-step test hooks ------------------------------------------------------------------------
namespace {
struct DevBuf {                        // frees on every exit path of the hooks below
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
};
}  // namespace

int orbgpu_debug_atan2(orbgpu_extractor* e, const float* y, const float* x, int n, float* out_deg)
{
    if (!e || !y || !x || !out_deg || n < 1) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    DevBuf buf;
    CU(cudaMalloc(&buf.p, 3 * sizeof(float) * (size_t)n));
    float* d = static_cast<float*>(buf.p);
    CU(cudaMemcpy(d, y, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d + n, x, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
    orb_debug_atan2(d, d + n, n, d + 2 * (size_t)n, 0);
    e->launches += 1;
    CU(cudaMemcpy(out_deg, d + 2 * (size_t)n, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    return ORBGPU_OK;
}

int orbgpu_debug_sincosf(orbgpu_extractor* e, const float* rad, int n, float* out_sin, float* out_cos)
{
    if (!e || !rad || !out_sin || !out_cos || n < 1) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    DevBuf buf;
    CU(cudaMalloc(&buf.p, 3 * sizeof(float) * (size_t)n));
    float* d = static_cast<float*>(buf.p);
    CU(cudaMemcpy(d, rad, sizeof(float) * (size_t)n, cudaMemcpyHostToDevice));
    orb_debug_sincosf(d, n, d + n, d + 2 * (size_t)n, 0);
    e->launches += 1;
    CU(cudaMemcpy(out_sin, d + n, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(out_cos, d + 2 * (size_t)n, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    return ORBGPU_OK;
}

int orbgpu_debug_sincos_checksum(orbgpu_extractor* e, uint32_t first_bits, uint32_t count, uint32_t step, uint64_t* checksum)
{
    if (!e || !checksum) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    DevBuf buf;
    CU(cudaMalloc(&buf.p, sizeof(unsigned long long)));
    unsigned long long* d = static_cast<unsigned long long*>(buf.p);
    CU(cudaMemset(d, 0, sizeof(unsigned long long)));
    orb_debug_sincos_checksum(first_bits, count, step, d, 0);
    e->launches += 1;
    unsigned long long h = 0;
    CU(cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost));
    *checksum = h;
    return ORBGPU_OK;
}

int orbgpu_set_profiling(orbgpu_extractor* e, int on)
{
    if (!e) return ORBGPU_ERR_ARG;
    e->profiling = on != 0;
    return ORBGPU_OK;
}

int orbgpu_get_stage_ms(orbgpu_extractor* e, double* ms, int n, int reset)
{
    if (!e || !ms) return ORBGPU_ERR_ARG;
    for (int i = 0; i < n && i < ORBGPU_NUM_STAGES; ++i) ms[i] = e->stage_ms[i];
    if (reset) memset(e->stage_ms, 0, sizeof(e->stage_ms));
    return ORBGPU_OK;
}

uint64_t orbgpu_launch_count(orbgpu_extractor* e) { return e ? e->launches : 0; }

int orbgpu_set_input_format(orbgpu_extractor* e, int channels, int rgb_order)
{
    if (!e || (channels != 1 && channels != 3 && channels != 4)) return ORBGPU_ERR_ARG;
    if (e->async_open) return ORBGPU_ERR_STATE;
    for (int s = 0; s < NSTREAM; ++s)
        if (e->ctx[s].busy || e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;
    if (cudaSetDevice(e->device) != cudaSuccess) return ORBGPU_ERR_CUDA;
    if (channels != e->in_channels)
        for (int s = 0; s < NSTREAM; ++s) ctx_free_buffers(e->ctx[s]);     // the pinned staging buffers are sized per format
    e->in_channels = channels; e->in_rgb = rgb_order != 0;
    e->last_ctx = -1;
    return ORBGPU_OK;
}

int orbgpu_set_batching(orbgpu_extractor* e, int streams, int max_chunk_frames)
{
    if (!e || streams < 1 || streams > NSTREAM || max_chunk_frames < 1) return ORBGPU_ERR_ARG;
    for (int s = 0; s < NSTREAM; ++s)
        if (e->ctx[s].busy || e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;
    e->nstream[0] = e->nstream[1] = streams;
    e->max_chunk[0] = e->max_chunk[1] = max_chunk_frames;
    return ORBGPU_OK;
}

int orbgpu_host_alloc(void** out, size_t bytes, int write_combined)
{
    if (!out || bytes == 0) return ORBGPU_ERR_ARG;
    *out = nullptr;
    CU(cudaHostAlloc(out, bytes, write_combined ? (cudaHostAllocWriteCombined | cudaHostAllocPortable) : cudaHostAllocPortable));
    return ORBGPU_OK;
}

void orbgpu_host_free(void* p)
{
    if (p) cudaFreeHost(p);
}

int orbgpu_chunk_capacity(orbgpu_extractor* e, int w, int h)
{
    if (!e) return ORBGPU_ERR_ARG;
    int rc;
    if ((rc = use_device(e)) != ORBGPU_OK) return rc;
    if ((rc = build_plan(e, w, h)) != ORBGPU_OK) return rc;
    return e->plan.chunk_cap;
}

int orbgpu_set_graphs(orbgpu_extractor* e, int on)
{
    if (!e) return ORBGPU_ERR_ARG;
    for (int s = 0; s < NSTREAM; ++s)
        if (e->ctx[s].busy || e->ctx[s].ticket_live) return ORBGPU_ERR_STATE;
    e->use_graphs = on != 0;
    return ORBGPU_OK;
}

double orbgpu_last_batch_ms(orbgpu_extractor* e) { return e ? e->last_batch_ms : 0.0; }

}  // extern "C"
