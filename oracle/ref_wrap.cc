// oracle/ref_wrap.cc -- TEST INFRASTRUCTURE, not product code.
//
// C entry points around the UNMODIFIED reference extractor (/root/reference/src/ORBextractor.cc,
// compiled from the mount by oracle/Makefile against oracle/cvshim) -> oracle/_ref/liborbref.so.
//
// Allocator pin.  DistributeOctTree sorts pair<int, ExtractorNode*> (src/ORBextractor.cc:686), so
// ties between equal-size nodes are broken by heap address, i.e. by malloc behaviour.  While an
// extraction runs, this library serves every operator new from a per-thread bump arena that hands
// out strictly ascending (mode 1) or strictly descending (mode 2) addresses and never reuses one:
// pointer order == creation order (or its reverse).  Mode 1 is the pinned semantics: among nodes
// of equal size the most recently created one is split first.  Mode 2 exists only to detect
// frames whose result depends on the tie-break.  Linked with -Bsymbolic-functions so these
// operators bind inside this .so only.
#include "ORBextractor.h"
#include <sys/mman.h>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

namespace {

const size_t ARENA_BYTES = (size_t)2 << 30;   // per thread, virtual (MAP_NORESERVE)
const int ARENA_SLOTS = 256;
char* g_region = 0;
std::atomic<int> g_next_slot{0};
std::atomic<int> g_region_state{0};           // 0 none, 1 being made, 2 ready

struct Arena {
    char* base = 0;
    char* lo = 0;      // ascending cursor
    char* hi = 0;      // descending cursor
    int mode = 0;      // 0 inactive, 1 ascending, 2 descending
    bool overflow = false;
    size_t peak = 0;
};
thread_local Arena t_arena;

bool in_region(void* p)
{
    return g_region && (char*)p >= g_region && (char*)p < g_region + ARENA_BYTES * ARENA_SLOTS;
}

void arena_begin(int mode)
{
    Arena& a = t_arena;
    if (!a.base) {
        int expected = 0;
        if (g_region_state.compare_exchange_strong(expected, 1)) {
            void* r = mmap(0, ARENA_BYTES * ARENA_SLOTS, PROT_READ | PROT_WRITE,
                           MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
            g_region = (r == MAP_FAILED) ? 0 : (char*)r;
            g_region_state.store(2);
        }
        while (g_region_state.load() != 2) std::this_thread::yield();
        int slot = g_next_slot.fetch_add(1);
        if (g_region && slot < ARENA_SLOTS) a.base = g_region + (size_t)slot * ARENA_BYTES;
    }
    a.lo = a.base;
    a.hi = a.base ? a.base + ARENA_BYTES : 0;
    a.overflow = (a.base == 0);
    a.mode = a.base ? mode : 0;
}

void arena_end()
{
    Arena& a = t_arena;
    if (!a.base) return;
    size_t used_lo = (size_t)(a.lo - a.base), used_hi = (size_t)(a.base + ARENA_BYTES - a.hi);
    if (used_lo + used_hi > a.peak) a.peak = used_lo + used_hi;
    const size_t keep = (size_t)32 << 20;     // give big frames' pages back to the OS
    if (used_lo > keep) madvise(a.base + keep, used_lo - keep, MADV_DONTNEED);
    if (used_hi > keep) madvise(a.hi, used_hi - keep, MADV_DONTNEED);
    a.mode = 0;
}

inline void* arena_alloc(size_t n)
{
    Arena& a = t_arena;
    n = (n + 15) & ~(size_t)15;
    if (n == 0) n = 16;
    if ((size_t)(a.hi - a.lo) < n) { a.overflow = true; return 0; }
    if (a.mode == 1) { void* p = a.lo; a.lo += n; return p; }
    a.hi -= n;
    return a.hi;
}

}  // namespace

void* operator new(size_t n)
{
    if (t_arena.mode) { void* p = arena_alloc(n); if (p) return p; }
    void* p = std::malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !in_region(p)) std::free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

namespace {

// protected members of the reference class are reached through a derived class; the reference
// source itself is untouched.
struct RefExtractor : public ORB_SLAM2::ORBextractor {
    RefExtractor(int nf, float sf, int nl, int ini, int mn) : ORB_SLAM2::ORBextractor(nf, sf, nl, ini, mn) {}
    void pyramid(const cv::Mat& im) { ComputePyramid(im); }
    std::vector<cv::KeyPoint> octree(const std::vector<cv::KeyPoint>& keys, int minX, int maxX, int minY,
                                     int maxY, int N, int level)
    {
        return DistributeOctTree(keys, minX, maxX, minY, maxY, N, level);
    }
    const std::vector<int>& features_per_level() const { return mnFeaturesPerLevel; }
    const std::vector<int>& umax_table() const { return umax; }
    int n_levels() const { return nlevels; }
};

cv::Mat wrap_image(const uint8_t* img, int w, int h, size_t stride)
{
    cv::Mat m(h, w, CV_8UC1);
    for (int y = 0; y < h; ++y) std::memcpy(m.ptr(y), img + (size_t)y * stride, (size_t)w);
    return m;
}

}  // namespace

extern "C" {

typedef struct { float x, y, size, angle, response; int32_t octave, class_id; } orbref_keypoint;

void* orbref_create(int nfeatures, float scale, int nlevels, int iniTh, int minTh)
{
    return new RefExtractor(nfeatures, scale, nlevels, iniTh, minTh);
}

void orbref_destroy(void* h) { delete (RefExtractor*)h; }

// tables: out_scale[4*nlevels] = scale, inv scale, sigma2, inv sigma2; feats[nlevels]; umax[16]
int orbref_tables(void* h, float* out_scale, int* feats, int* umax)
{
    RefExtractor* e = (RefExtractor*)h;
    int nl = e->GetLevels();
    std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(),
                       c = e->GetScaleSigmaSquares(), d = e->GetInverseScaleSigmaSquares();
    for (int i = 0; i < nl; ++i) {
        out_scale[i] = a[i]; out_scale[nl + i] = b[i]; out_scale[2 * nl + i] = c[i]; out_scale[3 * nl + i] = d[i];
        feats[i] = e->features_per_level()[i];
    }
    for (int i = 0; i < 16; ++i) umax[i] = e->umax_table()[i];
    return nl;
}

// Full operator().  arena_mode: 1 ascending (pinned), 2 descending (tie probe), 0 plain malloc.
// Returns 0, or -1 if cap is too small, -2 if the arena overflowed (tie-break pin lost).
int orbref_extract(void* h, const uint8_t* img, int w, int hgt, size_t stride, orbref_keypoint* kps,
                   uint8_t* desc, int cap, int* n_out, int arena_mode)
{
    RefExtractor* e = (RefExtractor*)h;
    cv::Mat image = wrap_image(img, w, hgt, stride);
    cv::Mat descriptors;
    int rc = 0;
    if (arena_mode) arena_begin(arena_mode);
    {
        std::vector<cv::KeyPoint> keypoints;
        (*e)(image, cv::Mat(), keypoints, descriptors);
        int n = (int)keypoints.size();
        *n_out = n;
        if (n > cap) rc = -1;
        else {
            static_assert(sizeof(cv::KeyPoint) == sizeof(orbref_keypoint), "KeyPoint must be 28 B");
            if (n) std::memcpy(kps, keypoints.data(), sizeof(orbref_keypoint) * (size_t)n);
            for (int i = 0; i < n; ++i) std::memcpy(desc + 32 * (size_t)i, descriptors.ptr(i), 32);
        }
    }
    if (arena_mode) { if (t_arena.overflow) rc = -2; arena_end(); }
    return rc;
}

// mvImagePyramid[level] after the last extract / pyramid call (interior, without the 19-px border)
int orbref_pyramid_level(void* h, int level, uint8_t* dst, size_t dstride, int* w, int* hgt)
{
    RefExtractor* e = (RefExtractor*)h;
    if (level < 0 || level >= e->GetLevels()) return -1;
    const cv::Mat& m = e->mvImagePyramid[level];
    *w = m.cols; *hgt = m.rows;
    if (dst)
        for (int y = 0; y < m.rows; ++y) std::memcpy(dst + (size_t)y * dstride, m.ptr(y), (size_t)m.cols);
    return 0;
}

// the bordered (w+38) x (h+38) buffer behind mvImagePyramid[level]
int orbref_pyramid_level_bordered(void* h, int level, uint8_t* dst, size_t dstride)
{
    RefExtractor* e = (RefExtractor*)h;
    const cv::Mat& m = e->mvImagePyramid[level];
    const int b = 19;
    for (int y = -b; y < m.rows + b; ++y)
        std::memcpy(dst + (size_t)(y + b) * dstride, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step.v - b, (size_t)m.cols + 2 * b);
    return 0;
}

int orbref_compute_pyramid(void* h, const uint8_t* img, int w, int hgt, size_t stride)
{
    RefExtractor* e = (RefExtractor*)h;
    e->pyramid(wrap_image(img, w, hgt, stride));
    return 0;
}

// DistributeOctTree on an arbitrary candidate list (x, y, response triples, coordinates relative
// to the (16,16) border origin as at src/ORBextractor.cc:824-825).  Returns the selected count.
int orbref_octree(void* h, const float* xs, const float* ys, const float* resp, int n, int minX, int maxX,
                  int minY, int maxY, int N, float* out_x, float* out_y, float* out_resp, int cap, int arena_mode)
{
    RefExtractor* e = (RefExtractor*)h;
    int cnt = 0, rc = 0;
    if (arena_mode) arena_begin(arena_mode);
    {
        std::vector<cv::KeyPoint> keys((size_t)n);
        for (int i = 0; i < n; ++i) keys[i] = cv::KeyPoint(xs[i], ys[i], 7.f, -1.f, resp[i]);
        std::vector<cv::KeyPoint> out = e->octree(keys, minX, maxX, minY, maxY, N, 0);
        cnt = (int)out.size();
        for (int i = 0; i < cnt && i < cap; ++i) { out_x[i] = out[i].pt.x; out_y[i] = out[i].pt.y; out_resp[i] = out[i].response; }
    }
    if (arena_mode) { if (t_arena.overflow) rc = -2; arena_end(); }
    return rc ? rc : cnt;
}

// CPU baseline: time `reps` passes over n frames (dense, frame_stride apart) on `threads` host
// threads, one extractor instance per thread over disjoint frames (the pattern of
// src/Frame.cc:78-81).  Returns wall seconds; total_kp receives the keypoint count (keeps the
// work observable).
double orbref_bench(const uint8_t* frames, int n, int w, int hgt, size_t frame_stride, int nfeatures,
                    float scale, int nlevels, int iniTh, int minTh, int threads, int reps, long* total_kp)
{
    if (threads < 1) threads = 1;
    std::vector<long> counts((size_t)threads, 0);
    std::vector<RefExtractor*> ex((size_t)threads);
    for (int t = 0; t < threads; ++t) ex[t] = new RefExtractor(nfeatures, scale, nlevels, iniTh, minTh);
    auto t0 = std::chrono::steady_clock::now();
    {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t)
            pool.emplace_back([&, t]() {
                long c = 0;
                for (int r = 0; r < reps; ++r)
                    for (int i = t; i < n; i += threads) {
                        cv::Mat image = wrap_image(frames + (size_t)i * frame_stride, w, hgt, (size_t)w);
                        cv::Mat descriptors;
                        arena_begin(1);
                        {
                            std::vector<cv::KeyPoint> kps;
                            (*ex[t])(image, cv::Mat(), kps, descriptors);
                            c += (long)kps.size();
                        }
                        arena_end();
                    }
                counts[t] = c;
            });
        for (auto& th : pool) th.join();
    }
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    long tot = 0;
    for (int t = 0; t < threads; ++t) { tot += counts[t]; delete ex[t]; }
    if (total_kp) *total_kp = tot;
    return sec;
}

size_t orbref_arena_peak(void) { return t_arena.peak; }

}  // extern "C"
