// oracle/orb_port.cc -- TEST INFRASTRUCTURE, not product code.  See orb_port.h.
// Compile with -ffp-contract=off (oracle/Makefile).
#include "orb_port.h"
#include "orb_prims.h"
#include "orb_pattern.h"   // the 512-point rBRIEF table (data constant, my-slam_b200/csrc)

#include <algorithm>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <list>
#include <thread>
#include <vector>

namespace {

const int kPatch = 31, kHalfPatch = 15, kEdge = 19;   // src/ORBextractor.cc:74-76

struct Cand { int x, y, score; };   // x, y relative to the (16,16) grid origin

struct Level {
    int w = 0, h = 0;
    std::vector<uint8_t> pix, blur;
    std::vector<Cand> cand, sel;
};

}  // namespace

struct orbport {
    int nfeatures, nlevels, ini_th, min_th;
    double scale_factor;                       // the reference keeps a double member (ORBextractor.h:98)
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> feats, umax;
    std::vector<Level> lv;
};

// ---- ORBextractor::ORBextractor, src/ORBextractor.cc:412-472 -----------------------------------
static void build_tables(orbport* p, float scale_factor_f)
{
    const int nl = p->nlevels;
    p->scale_factor = scale_factor_f;
    p->scale.assign(nl, 1.0f); p->sigma2.assign(nl, 1.0f);
    for (int i = 1; i < nl; ++i) {
        p->scale[i] = (float)(p->scale[i - 1] * p->scale_factor);    // float * double, narrowed
        p->sigma2[i] = p->scale[i] * p->scale[i];
    }
    p->inv_scale.resize(nl); p->inv_sigma2.resize(nl);
    for (int i = 0; i < nl; ++i) {
        p->inv_scale[i] = 1.0f / p->scale[i];
        p->inv_sigma2[i] = 1.0f / p->sigma2[i];
    }
    p->feats.assign(nl, 0);
    float factor = (float)(1.0f / p->scale_factor);
    float desired = p->nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        p->feats[l] = orbo_roundf(desired);
        sum += p->feats[l];
        desired *= factor;
    }
    p->feats[nl - 1] = std::max(p->nfeatures - sum, 0);

    // circular patch row extents, :456-471
    p->umax.assign(kHalfPatch + 1, 0);
    int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) p->umax[v] = orbo_round(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (p->umax[v0] == p->umax[v0 + 1]) ++v0;
        p->umax[v] = v0;
        ++v0;
    }
}

// ---- ORBextractor::ComputePyramid, src/ORBextractor.cc:1109-1137 ------------------------------
// (the 19-px REFLECT_101 frame around each level is never read by the extractor; not kept here)
static void compute_pyramid(orbport* p, const uint8_t* img, int w, int h, size_t stride)
{
    p->lv.assign(p->nlevels, Level());
    for (int l = 0; l < p->nlevels; ++l) {
        Level& L = p->lv[l];
        float s = p->inv_scale[l];
        L.w = orbo_roundf((float)w * s);
        L.h = orbo_roundf((float)h * s);
        L.pix.resize((size_t)L.w * L.h);
        if (l == 0) {
            for (int y = 0; y < h; ++y) std::memcpy(&L.pix[(size_t)y * w], img + (size_t)y * stride, (size_t)w);
        } else {
            const Level& S = p->lv[l - 1];     // chained: from the previous level, not the source
            orbo_resize_linear_u8(S.pix.data(), S.w, S.h, (size_t)S.w, L.pix.data(), L.w, L.h, (size_t)L.w);
        }
    }
}

// ---- cell grid + FAST with per-cell threshold fallback, src/ORBextractor.cc:767-831 ------------
static void level_candidates(const orbport* p, Level& L)
{
    L.cand.clear();
    const int minBX = kEdge - 3, minBY = minBX;
    const int maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
    if (nCols < 1 || nRows < 1) return;        // reference: empty loops (ceil(x/0) is never used)
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);

    std::vector<int> xs(4096), ys(4096), sc(4096);
    for (int i = 0; i < nRows; ++i) {
        const int iniY = minBY + i * hCell;
        int maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = maxBY;
        for (int j = 0; j < nCols; ++j) {
            const int iniX = minBX + j * wCell;
            int maxX = iniX + wCell + 6;
            if (iniX >= maxBX - 6) continue;
            if (maxX > maxBX) maxX = maxBX;
            const int cw = maxX - iniX, ch = maxY - iniY;
            const int cap = ((cw + 1) / 2) * ((ch + 1) / 2) + 1;
            if ((int)xs.size() < cap) { xs.resize(cap); ys.resize(cap); sc.resize(cap); }
            const uint8_t* cell = &L.pix[(size_t)iniY * L.w + iniX];
            int n = orbo_fast9_16(cell, cw, ch, (size_t)L.w, p->ini_th, 1, xs.data(), ys.data(), sc.data(), cap);
            if (n == 0)
                n = orbo_fast9_16(cell, cw, ch, (size_t)L.w, p->min_th, 1, xs.data(), ys.data(), sc.data(), cap);
            for (int k = 0; k < n; ++k)
                L.cand.push_back(Cand{xs[k] + j * wCell, ys[k] + i * hCell, sc[k]});
        }
    }
}

// ---- ExtractorNode::DivideNode + ORBextractor::DistributeOctTree, src/ORBextractor.cc:483-765 --
// Nodes carry key indices; `seq` (creation counter) stands in for the heap address that breaks
// ties in the reference's sort of pair<int, ExtractorNode*> (:686) under the ascending-arena pin.
namespace {

struct QNode {
    int x0, x1, y0, y1;          // UL.x, UR.x, UL.y, BL.y
    std::vector<int> keys;       // indices into the candidate array, input order preserved
    long seq = 0;
};
typedef std::list<QNode> QList;

void split_node(const QNode& n, const int* xs, const int* ys, QNode out[4])
{
    const int halfX = (int)std::ceil((float)(n.x1 - n.x0) / 2), halfY = (int)std::ceil((float)(n.y1 - n.y0) / 2);
    const int mx = n.x0 + halfX, my = n.y0 + halfY;
    out[0].x0 = n.x0; out[0].x1 = mx;   out[0].y0 = n.y0; out[0].y1 = my;      // UL
    out[1].x0 = mx;   out[1].x1 = n.x1; out[1].y0 = n.y0; out[1].y1 = my;      // UR
    out[2].x0 = n.x0; out[2].x1 = mx;   out[2].y0 = my;   out[2].y1 = n.y1;    // BL
    out[3].x0 = mx;   out[3].x1 = n.x1; out[3].y0 = my;   out[3].y1 = n.y1;    // BR
    for (int k : n.keys) {
        const bool left = (float)xs[k] < (float)mx, top = (float)ys[k] < (float)my;
        out[left ? (top ? 0 : 2) : (top ? 1 : 3)].keys.push_back(k);
    }
}

}  // namespace

extern "C" int orbport_octree(const int* xs, const int* ys, const int* resp, int n, int W, int H, int N,
                              int* out_idx, int cap)
{
    const int nIni = (int)std::round((float)W / (float)H);
    if (nIni < 1) return -3;
    const float hX = (float)W / nIni;

    QList nodes;
    long seq = 0;
    std::vector<QList::iterator> roots;
    for (int i = 0; i < nIni; ++i) {
        QNode r;
        r.x0 = (int)(hX * (float)i); r.x1 = (int)(hX * (float)(i + 1)); r.y0 = 0; r.y1 = H;
        r.seq = seq++;
        nodes.push_back(r);
        roots.push_back(std::prev(nodes.end()));
    }
    for (int k = 0; k < n; ++k) roots[(size_t)((float)xs[k] / hX)]->keys.push_back(k);
    for (QList::iterator it = nodes.begin(); it != nodes.end();)
        it = it->keys.empty() ? nodes.erase(it) : std::next(it);

    // nodes created by the latest pass that still hold more than one key
    std::vector<QList::iterator> fresh;
    // split one node: children that own keys go to the FRONT of the list in UL,UR,BL,BR order
    auto expand = [&](QList::iterator it, std::vector<QList::iterator>& made) {
        QNode ch[4];
        split_node(*it, xs, ys, ch);
        for (int c = 0; c < 4; ++c) {
            if (ch[c].keys.empty()) continue;
            ch[c].seq = seq++;
            nodes.push_front(ch[c]);
            if (ch[c].keys.size() > 1) made.push_back(nodes.begin());
        }
        nodes.erase(it);
    };

    bool done = false;
    while (!done) {
        // full sweep (:596-667): every multi-key node present at the start of the sweep is split
        const size_t before = nodes.size();
        fresh.clear();
        std::vector<QList::iterator> todo;
        for (QList::iterator it = nodes.begin(); it != nodes.end(); ++it)
            if (it->keys.size() > 1) todo.push_back(it);
        for (QList::iterator it : todo) expand(it, fresh);

        if ((int)nodes.size() >= N || nodes.size() == before) break;
        if ((int)nodes.size() + 3 * (int)fresh.size() <= N) continue;

        // budgeted passes (:675-739): biggest nodes first, newest first among equals, stop at N
        while (!done) {
            const size_t prev = nodes.size();
            std::vector<QList::iterator> order = fresh;
            fresh.clear();
            std::sort(order.begin(), order.end(), [](const QList::iterator& a, const QList::iterator& b) {
                if (a->keys.size() != b->keys.size()) return a->keys.size() > b->keys.size();
                return a->seq > b->seq;
            });
            for (QList::iterator it : order) {
                expand(it, fresh);
                if ((int)nodes.size() >= N) break;
            }
            if ((int)nodes.size() >= N || nodes.size() == prev) done = true;
        }
    }

    // best response per node, first key wins ties (:743-762)
    int cnt = 0;
    for (const QNode& nd : nodes) {
        int best = nd.keys[0];
        for (size_t k = 1; k < nd.keys.size(); ++k)
            if ((float)resp[nd.keys[k]] > (float)resp[best]) best = nd.keys[k];
        if (cnt < cap) out_idx[cnt] = best;
        ++cnt;
    }
    return cnt;
}

// ---- IC_Angle, src/ORBextractor.cc:79-106 -------------------------------------------------------
static float ic_angle(const uint8_t* img, int stride, int cx, int cy, const std::vector<int>& umax)
{
    const uint8_t* c = img + (size_t)cy * stride + cx;
    int m01 = 0, m10 = 0;
    for (int v = -kHalfPatch; v <= kHalfPatch; ++v) {
        const int d = umax[std::abs(v)];
        for (int u = -d; u <= d; ++u) {
            const int val = c[u + v * stride];
            m10 += u * val;
            m01 += v * val;
        }
    }
    return orbo_fast_atan2((float)m01, (float)m10);
}

// ---- computeOrbDescriptor, src/ORBextractor.cc:108-149 -----------------------------------------
static void brief_descriptor(const uint8_t* blur, int stride, int cx, int cy, float angle_deg, uint8_t* desc)
{
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    const float angle = angle_deg * factorPI;
    float a, b;
    orbo_sincosf(angle, &b, &a);            // a = cos, b = sin
    const uint8_t* c = blur + (size_t)cy * stride + cx;
    for (int i = 0; i < 32; ++i) {
        int byte = 0;
        for (int k = 0; k < 8; ++k) {
            int v[2];
            for (int e = 0; e < 2; ++e) {
                const int p = 16 * i + 2 * k + e;
                const float px = (float)ORB_PATTERN_XY[2 * p], py = (float)ORB_PATTERN_XY[2 * p + 1];
                const int row = orbo_roundf(px * b + py * a);
                const int col = orbo_roundf(px * a - py * b);
                v[e] = c[row * stride + col];
            }
            byte |= (v[0] < v[1]) << k;
        }
        desc[i] = (uint8_t)byte;
    }
}

extern "C" {

orbport* orbport_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th)
{
    orbport* p = new orbport();
    p->nfeatures = nfeatures; p->nlevels = nlevels; p->ini_th = ini_th; p->min_th = min_th;
    build_tables(p, scale_factor);
    return p;
}
void orbport_destroy(orbport* p) { delete p; }

int orbport_tables(orbport* p, float* s4, int* feats, int* umax)
{
    const int nl = p->nlevels;
    for (int i = 0; i < nl; ++i) {
        s4[i] = p->scale[i]; s4[nl + i] = p->inv_scale[i]; s4[2 * nl + i] = p->sigma2[i]; s4[3 * nl + i] = p->inv_sigma2[i];
        feats[i] = p->feats[i];
    }
    for (int i = 0; i < 16; ++i) umax[i] = p->umax[i];
    return nl;
}

int orbport_pyramid(orbport* p, const uint8_t* img, int w, int h, size_t stride)
{
    compute_pyramid(p, img, w, h, stride);
    return 0;
}

// ---- ORBextractor::operator(), src/ORBextractor.cc:1045-1107 -----------------------------------
int orbport_extract(orbport* p, const uint8_t* img, int w, int h, size_t stride,
                    orbport_keypoint* kps, uint8_t* desc, int cap, int* n_out)
{
    *n_out = 0;
    if (!img || w <= 0 || h <= 0) return 0;
    compute_pyramid(p, img, w, h, stride);
    int total = 0;
    for (int l = 0; l < p->nlevels; ++l) {
        Level& L = p->lv[l];
        level_candidates(p, L);
        L.sel.clear();
        const int W = (L.w - kEdge + 3) - (kEdge - 3), H = (L.h - kEdge + 3) - (kEdge - 3);
        const int n = (int)L.cand.size();
        std::vector<int> xs(n), ys(n), rs(n), idx(n + 8);
        for (int k = 0; k < n; ++k) { xs[k] = L.cand[k].x; ys[k] = L.cand[k].y; rs[k] = L.cand[k].score; }
        int m = 0;
        if (W > 0 && H > 0) {
            m = orbport_octree(xs.data(), ys.data(), rs.data(), n, W, H, p->feats[l], idx.data(), n + 8);
            if (m < 0) return m;
        }
        for (int k = 0; k < m; ++k) L.sel.push_back(L.cand[idx[k]]);
        total += m;
    }
    *n_out = total;
    if (total > cap) return -1;

    int o = 0;
    for (int l = 0; l < p->nlevels; ++l) {
        Level& L = p->lv[l];
        if (L.sel.empty()) { L.blur.clear(); continue; }
        L.blur.resize(L.pix.size());
        orbo_gauss7_u8(L.pix.data(), L.w, L.h, (size_t)L.w, L.blur.data(), (size_t)L.w);
        const int size = (int)(kPatch * p->scale[l]);
        for (const Cand& c : L.sel) {
            const int x = c.x + kEdge - 3, y = c.y + kEdge - 3;
            orbport_keypoint& kp = kps[o];
            kp.angle = ic_angle(L.pix.data(), L.w, x, y, p->umax);            // on the unblurred level
            brief_descriptor(L.blur.data(), L.w, x, y, kp.angle, desc + 32 * (size_t)o);
            kp.x = (float)x; kp.y = (float)y;
            if (l != 0) { kp.x *= p->scale[l]; kp.y *= p->scale[l]; }
            kp.size = (float)size; kp.response = (float)c.score; kp.octave = l; kp.class_id = -1;
            ++o;
        }
    }
    return 0;
}

int orbport_level_size(orbport* p, int l, int* w, int* h)
{
    if (l < 0 || l >= (int)p->lv.size()) return -1;
    *w = p->lv[l].w; *h = p->lv[l].h;
    return 0;
}
static int copy_plane(const std::vector<uint8_t>& src, int w, int h, uint8_t* dst, size_t ds)
{
    if (src.empty()) return -1;
    for (int y = 0; y < h; ++y) std::memcpy(dst + (size_t)y * ds, &src[(size_t)y * w], (size_t)w);
    return 0;
}
int orbport_level_pixels(orbport* p, int l, uint8_t* dst, size_t ds) { return copy_plane(p->lv[l].pix, p->lv[l].w, p->lv[l].h, dst, ds); }
int orbport_level_blurred(orbport* p, int l, uint8_t* dst, size_t ds) { return copy_plane(p->lv[l].blur, p->lv[l].w, p->lv[l].h, dst, ds); }
static int copy_cands(const std::vector<Cand>& v, int* x, int* y, int* s, int cap)
{
    for (int k = 0; k < (int)v.size() && k < cap; ++k) { x[k] = v[k].x; y[k] = v[k].y; s[k] = v[k].score; }
    return (int)v.size();
}
int orbport_level_candidates(orbport* p, int l, int* x, int* y, int* s, int cap) { return copy_cands(p->lv[l].cand, x, y, s, cap); }
int orbport_level_selected(orbport* p, int l, int* x, int* y, int* s, int cap) { return copy_cands(p->lv[l].sel, x, y, s, cap); }

// ---- ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1647-1663 -------------------------------
int orbport_hamming256(const uint8_t* a, const uint8_t* b)
{
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4); std::memcpy(&y, b + 4 * i, 4);
        uint32_t v = x ^ y;
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0x0F0F0F0Fu) * 0x01010101u) >> 24);
    }
    return dist;
}

// ---- best / second-best scan, canonical form src/ORBmatcher.cc:201-226 -------------------------
void orbport_match_bruteforce(const uint8_t* A, int nA, const uint8_t* B, int nB,
                              int32_t* best_idx, uint16_t* best, uint16_t* second)
{
    for (int i = 0; i < nA; ++i) {
        int b1 = 256, b2 = 256, bi = -1;
        for (int j = 0; j < nB; ++j) {
            const int d = orbport_hamming256(A + 32 * (size_t)i, B + 32 * (size_t)j);
            if (d < b1) { b2 = b1; b1 = d; bi = j; }
            else if (d < b2) b2 = d;
        }
        best_idx[i] = bi; best[i] = (uint16_t)b1; second[i] = (uint16_t)b2;
    }
}

int orbport_match_count(const uint8_t* A, int nA, const uint8_t* B, int nB, int th_low, float ratio)
{
    std::vector<int32_t> bi(nA); std::vector<uint16_t> b1(nA), b2(nA);
    orbport_match_bruteforce(A, nA, B, nB, bi.data(), b1.data(), b2.data());
    int c = 0;
    for (int i = 0; i < nA; ++i)
        if (b1[i] <= th_low && (float)b1[i] < ratio * (float)b2[i]) ++c;
    return c;
}

// All-pairs form of the scan (BASELINE config 5) for the timed CPU baseline of the Hamming half of the metric:
// pair p = (a, b) counts the rows of keyframe a whose best match in keyframe b passes the accept rule of
// src/ORBmatcher.cc:228-232.  popcnt = 0: DescriptorDistance exactly as the reference writes it (the SWAR formula of
// src/ORBmatcher.cc:1647-1663, orbport_hamming256); popcnt = 1: the same scan with the hardware popcount
// instruction (what a CPU maintainer would write today).  Pairs are dealt to `threads` host threads in contiguous
// slices.  Returns the wall time in seconds.
}  // extern "C"
static inline int hamming256_popcnt(const uint8_t* a, const uint8_t* b)
{
    uint64_t x[4], y[4];
    std::memcpy(x, a, 32); std::memcpy(y, b, 32);
    return __builtin_popcountll(x[0] ^ y[0]) + __builtin_popcountll(x[1] ^ y[1]) + __builtin_popcountll(x[2] ^ y[2]) +
           __builtin_popcountll(x[3] ^ y[3]);
}
template <bool POPCNT>
static void count_pairs_slice(const uint8_t* desc, int n, const int32_t* pairs, int p0, int p1, int th_low, float ratio, int32_t* counts)
{
    for (int p = p0; p < p1; ++p) {
        const uint8_t* A = desc + 32 * (size_t)n * pairs[2 * p];
        const uint8_t* B = desc + 32 * (size_t)n * pairs[2 * p + 1];
        int c = 0;
        for (int i = 0; i < n; ++i) {
            int b1 = 256, b2 = 256;
            for (int j = 0; j < n; ++j) {
                const int d = POPCNT ? hamming256_popcnt(A + 32 * (size_t)i, B + 32 * (size_t)j)
                                     : orbport_hamming256(A + 32 * (size_t)i, B + 32 * (size_t)j);
                if (d < b1) { b2 = b1; b1 = d; }
                else if (d < b2) b2 = d;
            }
            if (b1 <= th_low && (float)b1 < ratio * (float)b2) ++c;
        }
        counts[p] = c;
    }
}
extern "C" {
double orbport_count_pairs_bench(const uint8_t* desc, int n_per_kf, const int32_t* pairs, int num_pairs, int th_low, float ratio,
                                 int32_t* counts, int threads, int popcnt)
{
    if (threads < 1) threads = 1;
    if (threads > num_pairs) threads = num_pairs > 0 ? num_pairs : 1;
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < threads; ++t) {
        const int p0 = (int)((long long)num_pairs * t / threads), p1 = (int)((long long)num_pairs * (t + 1) / threads);
        if (popcnt) th.emplace_back(count_pairs_slice<true>, desc, n_per_kf, pairs, p0, p1, th_low, ratio, counts);
        else        th.emplace_back(count_pairs_slice<false>, desc, n_per_kf, pairs, p0, p1, th_low, ratio, counts);
    }
    for (auto& x : th) x.join();
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // extern "C"


// ---- Frame::ComputeStereoMatches, src/Frame.cc:466-640 (pinned by oracle/_ref/libframeref.so, see orb_port.h) --------
namespace {
inline int refl101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}
inline float lvl_px(const uint8_t* lvl, int w, int h, int x, int y) { return (float)lvl[(size_t)refl101(y, h) * w + refl101(x, w)]; }
}  // namespace

extern "C" void orbport_stereo_matches(const orbport_keypoint* kL, const uint8_t* dL, int nL, const orbport_keypoint* kR,
                                       const uint8_t* dR, int nR, const uint8_t* const* lvlL, const uint8_t* const* lvlR,
                                       const int* w, const int* h, const float* scale, const float* inv_scale, float bf, float b,
                                       float* u_right, float* depth)
{
    const int TH_HIGH = 100, TH_LOW = 50;                      // src/ORBmatcher.cc:37-38
    for (int i = 0; i < nL; ++i) { u_right[i] = -1.0f; depth[i] = -1.0f; }        // :468-469
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;              // :471
    const int nRows = h[0];                                    // :473
    std::vector<std::vector<size_t> > rows((size_t)nRows);     // :476-495 right keypoints per image row
    for (int iR = 0; iR < nR; ++iR) {
        const float kpY = kR[iR].y;
        const float r = 2.0f * scale[kR[iR].octave];
        const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; ++yi)
            if (yi >= 0 && yi < nRows) rows[(size_t)yi].push_back((size_t)iR);      // (the reference does not guard yi)
    }
    const float minZ = b, minD = 0, maxD = bf / minZ;          // :497-499
    std::vector<std::pair<int, int> > vDistIdx;
    for (int iL = 0; iL < nL; ++iL) {
        const orbport_keypoint& kpL = kL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const std::vector<size_t>& cand = rows[(size_t)vL];    // :512 float -> index
        if (cand.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        for (size_t iC = 0; iC < cand.size(); ++iC) {          // :528-549
            const size_t iR = cand[iC];
            const orbport_keypoint& kpR = kR[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = orbport_hamming256(dL + 32 * (size_t)iL, dR + 32 * iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {                            // :552 sub-pixel match by correlation
            const float uR0 = kR[bestIdxR].x;
            const float sf = inv_scale[levelL];
            const float scaleduL = std::round(kpL.x * sf), scaledvL = std::round(kpL.y * sf), scaleduR0 = std::round(uR0 * sf);
            const int wdw = 5, L = 5;
            const int cxL = (int)scaleduL, cy = (int)scaledvL, cxR = (int)scaleduR0;
            const uint8_t* PL = lvlL[levelL]; const uint8_t* PR = lvlR[levelL];
            const int lw = w[levelL], lh = h[levelL];
            int bestSad = INT_MAX, bestincR = 0;
            float vDists[2 * 5 + 1];
            const float iniu = scaleduR0 + L - wdw, endu = scaleduR0 + L + wdw + 1;
            if (iniu < 0 || endu >= lw) continue;              // :572-575
            const float cL = lvl_px(PL, lw, lh, cxL, cy);
            for (int incR = -L; incR <= L; ++incR) {           // :577-592
                const float cR = lvl_px(PR, lw, lh, cxR + incR, cy);
                double acc = 0;                                // cv::norm(NORM_L1) accumulates in double
                for (int dy = -wdw; dy <= wdw; ++dy)
                    for (int dx = -wdw; dx <= wdw; ++dx) {
                        const float a = lvl_px(PL, lw, lh, cxL + dx, cy + dy) - cL;
                        const float bb = lvl_px(PR, lw, lh, cxR + incR + dx, cy + dy) - cR;
                        acc += std::fabs(a - bb);
                    }
                const float dist = (float)acc;
                if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;     // :594
            const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));     // :602
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = scale[levelL] * ((float)scaleduR0 + (float)bestincR + deltaR);       // :608
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {       // :612-622
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depth[iL] = bf / disparity;
                u_right[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(bestSad, iL));
            }
        }
    }
    if (vDistIdx.empty()) return;                              // (the reference indexes the empty vector, :627)
    std::sort(vDistIdx.begin(), vDistIdx.end());               // :626-639
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)vDistIdx.size() - 1; i >= 0; --i) {
        if (vDistIdx[i].first < thDist) break;
        u_right[vDistIdx[i].second] = -1;
        depth[vDistIdx[i].second] = -1;
    }
}
