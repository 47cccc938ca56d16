// orb_common.cuh -- device-side layout shared by the kernels and the host engine.
//
// Data layout in HBM (one "chunk" = up to chunk_cap same-shape frames processed per launch set):
//   pyr    [frame][level 1..L-1]  u8, level l at lv[l].pyr_off, row pitch lv[l].pitch (multiple of 16 B)
//          level 0 is read in place from the upload / caller buffer (its own row + frame strides)
//   blur   [frame][level 0..L-1]  u8, same pitches, lv[l].blur_off
//   cellcnt[frame][cell]          u32 survivors of each FAST cell
//   slots  [frame][cell][cap]     u32 packed candidates, cell c at cells[c].slot_off
//   sel    [frame][level][selcap] u32 packed octree survivors (list order), selcnt[frame][level]
//   kps / desc / nout             final records (28 B / 32 B per keypoint, capacity per frame)
// Packed candidate = x | y << 12 | score << 24, x/y relative to the (16,16) cell-grid origin
// (src/ORBextractor.cc:824-825), so level width/height - 32 must be <= 4096.
#pragma once
#include <stdint.h>

#define ORB_MAX_LEVELS 16
#define ORB_EDGE 19          // EDGE_THRESHOLD, src/ORBextractor.cc:76
#define ORB_BORDER 16        // EDGE_THRESHOLD - 3: origin of the FAST cell grid, :773
#define ORB_HALF_PATCH 15    // HALF_PATCH_SIZE, :75
#define ORB_PATCH 31         // PATCH_SIZE, :74
#ifndef ORB_BLUR_ROWS
#define ORB_BLUR_ROWS 28     // output rows per k_blur thread (multiple of 7)
#endif
#ifndef ORB_BLUR_THREADS
#define ORB_BLUR_THREADS 128 // tasks per k_blur CTA

// Index checks of our own (compute-sanitizer is closed on this pool): compiled in with -DORB_DEBUG_CHECKS, a failed check
// traps the kernel with file:line (device assert).  The release build compiles them out.
#ifdef ORB_DEBUG_CHECKS
#include <assert.h>
#define ORB_CHECK(cond) assert(cond)
#else
#define ORB_CHECK(cond) ((void)0)
#endif

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of (kernel, DEVICE): the opt-in is remembered per device
// ordinal, behind a mutex, so that handles on several GPUs of one process -- possibly on different host threads -- each
// raise the limit on their own device.  One instance per kernel (a function-local static at the launch site).
#include <mutex>
struct OrbSmemOptIn {
    std::mutex mu;
    int configured[64] = {};
    template <class K>
    cudaError_t ensure(K kernel, int bytes)
    {
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
        std::lock_guard<std::mutex> lk(mu);
        if (bytes > configured[dev]) {
            e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
            if (e != cudaSuccess) return e;
            configured[dev] = bytes;
        }
        return cudaSuccess;
    }
};

#endif

struct OrbLevel {
    int w, h, pitch;             // level size; pitch of pyr/blur rows for this level
    long long pyr_off;           // byte offset inside one frame's pyr block (levels >= 1)
    long long blur_off;          // byte offset inside one frame's blur block
    int cell0, ncells;           // this level's cells in the cell table
    int slot_total;              // sum of cell caps = worst-case candidate count
    long long slot_off;          // u32 offset of this level's slots inside one frame's slot block
    int nfeat;                   // mnFeaturesPerLevel[level]
    int n_ini;                   // octree roots, :544
    float hx;                    // root width, :546
    int W, H;                    // maxBorder - minBorder
    int sel_cap, sel_off;        // capacity / offset of this level inside one frame's sel block
    int node_cap;                // max list nodes the octree can hold for this level
    long long key_scratch_off;   // byte offset of this level's key spill area in one frame's scratch
    long long node_scratch_off;  // byte offset of this level's node spill area
    float scale;                 // mvScaleFactor[level]
    float kp_size;               // (float)(int)(31 * scale), :839
    int xofs_off, yofs_off;      // offsets into the resize tables (levels >= 1)
};

struct OrbCell {                 // one FAST cell (src/ORBextractor.cc:791-831), 24 B
    short level;
    short cw, ch;                // cell image size (interior = cw-6 x ch-6)
    short ini_x, ini_y;          // top-left of the cell image in level coordinates
    short off_x, off_y;          // j*wCell + 3, i*hCell + 3: added to interior coordinates
    short pad;
    int slot_off;                // u32 offset inside the level's slot area
    int cap;                     // ceil(iw/2)*ceil(ih/2): most strict 8-neighbour maxima a cell can hold
};

struct OrbPlanDev {              // passed by value to every kernel
    OrbLevel lv[ORB_MAX_LEVELS];
    int nlevels;
    int ini_th, min_th;
    int ncells;
    int sel_total;               // sum of sel_cap
    int kp_capacity;             // per-frame capacity of kps/desc
    int fast_rows, fast_spw, fast_tp;   // k_fast_cells: rows of a warp's cell tile (max cell height); pitch of its score map and of the tile
    long long pyr_frame_bytes, blur_frame_bytes, slot_frame_words, scratch_frame_bytes;
    // level 0 source
    const uint8_t* src;          // frame 0, row 0
    long long src_frame_stride;
    int src_pitch;
    // work buffers (device pointers, frame 0)
    uint8_t* pyr;
    uint8_t* blur;
    uint32_t* cellcnt;
    uint32_t* slots;
    uint32_t* sel;
    int* selcnt;                 // [frame][ORB_MAX_LEVELS]
    int* candcnt;                // [frame][ORB_MAX_LEVELS] FAST candidates per level (after the octree's cell scan)
    uint8_t* scratch;
    const OrbCell* cells;
    const short* rs_ofs;         // resize tables: source index per destination index
    const short* rs_a;           // interleaved (a0, a1) 11-bit weights
    // outputs
    void* kps;                   // orbgpu_keypoint[frame][kp_capacity]
    uint8_t* desc;               // [frame][kp_capacity][32]
    int* nout;                   // [frame]
};

__host__ __device__ inline const uint8_t* orb_level_ptr(const OrbPlanDev& P, int l, int frame)
{
    return l == 0 ? P.src + (long long)frame * P.src_frame_stride
                  : P.pyr + (long long)frame * P.pyr_frame_bytes + P.lv[l].pyr_off;
}
__host__ __device__ inline int orb_level_pitch(const OrbPlanDev& P, int l) { return l == 0 ? P.src_pitch : P.lv[l].pitch; }

#define ORB_PACK(x, y, s) ((uint32_t)(x) | ((uint32_t)(y) << 12) | ((uint32_t)(s) << 24))
#define ORB_PX(p) ((int)((p) & 0xFFFu))
#define ORB_PY(p) ((int)(((p) >> 12) & 0xFFFu))
#define ORB_PS(p) ((int)((p) >> 24))

// Frame::ComputeStereoMatches on the device (src/Frame.cc:466-640): the unblurred pyramids, keypoints and
// descriptors of the most recent left / right frame stay where the two extractors left them in HBM.
struct OrbStereoSide {
    const uint8_t* lvl[ORB_MAX_LEVELS];      // level l, row 0 (frame 0 of the context that holds the frame)
    int pitch[ORB_MAX_LEVELS], w[ORB_MAX_LEVELS], h[ORB_MAX_LEVELS];
    const void* kps;                         // orbgpu_keypoint[n]
    const uint8_t* desc;                     // [n][32]
    const int* n_ptr;                        // device: keypoint count of that frame
};
struct OrbStereoArgs {
    OrbStereoSide L, R;
    int nlevels;
    float scale[ORB_MAX_LEVELS], inv_scale[ORB_MAX_LEVELS];
    float bf, b;                             // mbf, mb
    int th_orb_dist, th_high;                // (TH_HIGH + TH_LOW) / 2, TH_HIGH  (src/Frame.cc:471, :522)
    float* u_right;                          // [L.n] mvuRight before the median filter (-1 = no match)
    float* depth;                            // [L.n] mvDepth
    int* sad;                                // [L.n] best SAD of the sub-pixel search (-1 = no match)
    int* counts;                             // [2] nL, nR as the kernel read them
};
int orb_launch_stereo(const OrbStereoArgs& A, int max_left, cudaStream_t st);
// orb_frame.cu: K4 = fx, fy, cx, cy; dist = k1, k2, p1, p2[, k3 ...] (ndist <= 12), both host pointers
int orb_launch_undistort_points(const float* K4, const float* dist, int ndist, const float* d_src, float* d_dst, int n, cudaStream_t st);
int orb_launch_undistort_keypoints(const float* K4, const float* dist, int ndist, const void* d_kps, void* d_kps_un, const int* d_n,
                                   int capacity, cudaStream_t st);

// kernel launchers (orb_kernels.cu); every one enqueues on `st` and returns the launch count
int orb_launch_pyramid(const OrbPlanDev& P, int nframes, cudaStream_t st);
int orb_launch_fast(const OrbPlanDev& P, int nframes, int smem_bytes, cudaStream_t st);
int orb_launch_octree(const OrbPlanDev& P, int nframes, int key_smem_cap, int node_smem_cap, cudaStream_t st);
int orb_launch_blur(const OrbPlanDev& P, int nframes, const int4* tiles, int ntiles, cudaStream_t st);
int orb_launch_describe(const OrbPlanDev& P, int nframes, cudaStream_t st);
// standalone octree on one candidate list already packed in d_keys (for orbgpu_octree)
int orb_launch_octree_single(const uint32_t* d_keys, int n, int W, int H, int N, int n_ini, float hx,
                             int node_cap, uint8_t* d_scratch, uint32_t* d_out, int* d_count,
                             int key_smem_cap, int node_smem_cap, cudaStream_t st);
int orb_launch_repitch(const uint8_t* src, long long src_frame_stride, int src_pitch, uint8_t* dst, int dst_pitch, int w, int h,
                       int nframes, cudaStream_t st);
int orb_launch_cvt_gray(const uint8_t* src, long long src_frame_stride, int src_pitch, int channels, int rgb_order, uint8_t* dst,
                        int dst_pitch, int w, int h, int nframes, cudaStream_t st);
int orb_upload_constants(const int* umax16);
int orb_fast_smem_bytes(int max_cw, int max_ch);
// pitch of the per-warp FAST score map: interior width + the 1-px rim the NMS reads, rounded up to 4
int orb_fast_tile_pitch(int max_cw);
inline int orb_fast_spw(int max_cw) { return ((max_cw - 6 + 2) + 3) & ~3; }
int orb_debug_atan2(const float* d_y, const float* d_x, int n, float* d_out, cudaStream_t st);
int orb_debug_sincosf(const float* d_rad, int n, float* d_s, float* d_c, cudaStream_t st);
int orb_debug_sincos_checksum(unsigned first_bits, unsigned count, unsigned step, unsigned long long* d_sum, cudaStream_t st);
