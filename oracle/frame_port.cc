// frame_port.cc -- TEST INFRASTRUCTURE (oracle); pinned against the reference's own code, see frame_port.h.
#include "frame_port.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <vector>

namespace {

const int GRID_ROWS = 48, GRID_COLS = 64;                     // include/Frame.h:37-38
const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;      // src/ORBmatcher.cc:37-39

struct Grid {                                                 // Frame::mGrid + the members GetFeaturesInArea reads
    std::vector<size_t> cell[GRID_COLS][GRID_ROWS];
    float minX, maxX, minY, maxY, invW, invH;
    const orbport_keypoint* keys;
    int N;
};

// src/Frame.cc:383-391 + :230-245; grid element sizes as src/Frame.cc:155-156
void build_grid(Grid& g, const orbport_keypoint* keys, int n, const float* b)
{
    g.minX = b[0]; g.maxX = b[1]; g.minY = b[2]; g.maxY = b[3];
    g.invW = static_cast<float>(GRID_COLS) / static_cast<float>(g.maxX - g.minX);
    g.invH = static_cast<float>(GRID_ROWS) / static_cast<float>(g.maxY - g.minY);
    g.keys = keys; g.N = n;
    for (int i = 0; i < n; ++i) {
        const int posX = std::round((keys[i].x - g.minX) * g.invW);
        const int posY = std::round((keys[i].y - g.minY) * g.invH);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        g.cell[posX][posY].push_back((size_t)i);
    }
}

// src/Frame.cc:327-380
std::vector<size_t> features_in_area(const Grid& g, float x, float y, float r, int minLevel, int maxLevel)
{
    std::vector<size_t> v;
    const int nMinCellX = std::max(0, (int)std::floor((x - g.minX - r) * g.invW));
    if (nMinCellX >= GRID_COLS) return v;
    const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - g.minX + r) * g.invW));
    if (nMaxCellX < 0) return v;
    const int nMinCellY = std::max(0, (int)std::floor((y - g.minY - r) * g.invH));
    if (nMinCellY >= GRID_ROWS) return v;
    const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - g.minY + r) * g.invH));
    if (nMaxCellY < 0) return v;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ++ix)
        for (int iy = nMinCellY; iy <= nMaxCellY; ++iy) {
            const std::vector<size_t>& c = g.cell[ix][iy];
            for (size_t j = 0; j < c.size(); ++j) {
                const orbport_keypoint& kp = g.keys[c[j]];
                if (bCheckLevels) {
                    if (kp.octave < minLevel) continue;
                    if (maxLevel >= 0 && kp.octave > maxLevel) continue;
                }
                const float distx = kp.x - x, disty = kp.y - y;
                if (std::fabs(distx) < r && std::fabs(disty) < r) v.push_back(c[j]);
            }
        }
    return v;
}

// src/ORBmatcher.cc:1601-1642
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3)
{
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; ++i) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

inline int rot_bin(float a1, float a2)                        // src/ORBmatcher.cc:474-480, :1432-1438
{
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

// cv::Mat "A*x + c" for 3x3 * 3x1 CV_32F = cv::gemm's small-matrix case: float products summed left to right,
// then (float)(t*alpha + c*beta) with alpha = beta = 1.0 as doubles
inline void mat3_mul_add(const float* A, int astep, const float* x, const float* c, float* d)
{
    for (int i = 0; i < 3; ++i) {
        const float* a = A + i * astep;
        const float t = a[0] * x[0] + a[1] * x[1] + a[2] * x[2];
        d[i] = (float)((double)t * 1.0 + (double)c[i] * 1.0);
    }
}

}  // namespace

extern "C" int frameport_features_in_area(const orbport_keypoint* keys_un, int n, const float* bounds4, float x, float y, float r,
                                          int min_level, int max_level, int* out, int cap)
{
    Grid* g = new Grid;
    build_grid(*g, keys_un, n, bounds4);
    const std::vector<size_t> v = features_in_area(*g, x, y, r, min_level, max_level);
    delete g;
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) out[i] = (int)v[i];
    return (int)v.size();
}

extern "C" int frameport_search_for_initialization(const orbport_keypoint* k1, const uint8_t* d1, int n1, const orbport_keypoint* k2,
                                                   const uint8_t* d2, int n2, const float* bounds4, float* prev_matched,
                                                   int window_size, float nnratio, int check_orientation, int* matches12)
{
    Grid* g = new Grid;
    build_grid(*g, k2, n2, bounds4);
    int nmatches = 0;
    for (int i = 0; i < n1; ++i) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance((size_t)n2, INT_MAX), vnMatches21((size_t)n2, -1);
    for (int i1 = 0; i1 < n1; ++i1) {
        const int level1 = k1[i1].octave;
        if (level1 > 0) continue;
        const std::vector<size_t> vIndices2 = features_in_area(*g, prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)window_size, level1, level1);
        if (vIndices2.empty()) continue;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (size_t c = 0; c < vIndices2.size(); ++c) {
            const size_t i2 = vIndices2[c];
            const int dist = orbport_hamming256(d1 + 32 * (size_t)i1, d2 + 32 * i2);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = (int)i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnratio) {
                if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                matches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (check_orientation) rotHist[rot_bin(k1[i1].angle, k2[bestIdx2].angle)].push_back(i1);
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); ++j) {
                const int idx1 = rotHist[i][j];
                if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
            }
        }
    }
    for (int i1 = 0; i1 < n1; ++i1)
        if (matches12[i1] >= 0) { prev_matched[2 * i1] = k2[matches12[i1]].x; prev_matched[2 * i1 + 1] = k2[matches12[i1]].y; }
    delete g;
    return nmatches;
}

extern "C" int frameport_search_by_projection(const frameport_projection_args* a, int* cur_match)
{
    Grid* g = new Grid;
    build_grid(*g, a->cur_keys_un, a->n_cur, a->bounds);
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float* Tc = a->Tcw_cur; const float* Tl = a->Tcw_last;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]}, tlw[3] = {Tl[3], Tl[7], Tl[11]};
    // twc = -Rcw.t() * tcw (:1341): cv::operator-(MatExpr) materialises the transpose (MatOp::subtract has no MatOp_T
    // override) and leaves "-1 * copy", so the product is cv::gemm(copy, tcw, alpha = -1, flags = 0): the small-matrix
    // case, float products summed left to right, then (float)(t * -1.0).  Not a GEMM_1_T call (which would accumulate in
    // double); pinned by tests/test_ref_frame.py against the reference's own code on a knife-edge pose.
    float twc[3];
    for (int i = 0; i < 3; ++i) {
        const float t = Tc[i] * tcw[0] + Tc[4 + i] * tcw[1] + Tc[8 + i] * tcw[2];
        twc[i] = (float)((double)t * -1.0);
    }
    float tlc[3];
    mat3_mul_add(Tl, 4, twc, tlw, tlc);                       // :1347
    const bool bForward = tlc[2] > a->b && !a->mono;
    const bool bBackward = -tlc[2] > a->b && !a->mono;
    std::vector<int> holder((size_t)a->n_cur, -1);            // index of the last-frame point now in mvpMapPoints[i2]
    std::vector<uint8_t> occupied(a->cur_occupied, a->cur_occupied + a->n_cur);
    for (int i = 0; i < a->n_cur; ++i) cur_match[i] = -1;
    for (int i = 0; i < a->n_last; ++i) {
        if (!a->mp_valid[i]) continue;
        float x3Dc[3];
        mat3_mul_add(Tc, 4, a->world_pos + 3 * (size_t)i, tcw, x3Dc);
        const float xc = x3Dc[0], yc = x3Dc[1];
        const float invzc = 1.0 / x3Dc[2];
        if (invzc < 0) continue;
        const float u = a->fx * xc * invzc + a->cx;
        const float v = a->fy * yc * invzc + a->cy;
        if (u < a->bounds[0] || u > a->bounds[1]) continue;
        if (v < a->bounds[2] || v > a->bounds[3]) continue;
        const int nLastOctave = a->last_octave[i];
        const float radius = a->th * a->scale_factors[nLastOctave];
        std::vector<size_t> vIndices2;
        if (bForward) vIndices2 = features_in_area(*g, u, v, radius, nLastOctave, -1);
        else if (bBackward) vIndices2 = features_in_area(*g, u, v, radius, 0, nLastOctave);
        else vIndices2 = features_in_area(*g, u, v, radius, nLastOctave - 1, nLastOctave + 1);
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (size_t c = 0; c < vIndices2.size(); ++c) {
            const size_t i2 = vIndices2[c];
            if (occupied[i2]) continue;
            if (a->cur_u_right[i2] > 0) {
                const float ur = u - a->bf * invzc;
                const float er = std::fabs(ur - a->cur_u_right[i2]);
                if (er > radius) continue;
            }
            const int dist = orbport_hamming256(a->mp_desc + 32 * (size_t)i, a->cur_desc + 32 * i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = (int)i2; }
        }
        if (bestDist <= TH_HIGH) {
            holder[bestIdx2] = i;
            occupied[bestIdx2] = a->mp_observed[i];
            cur_match[bestIdx2] = i;
            nmatches++;
            if (a->check_orientation) rotHist[rot_bin(a->last_angle[i], a->cur_keys_un[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (a->check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i)
            if (i != ind1 && i != ind2 && i != ind3)
                for (size_t j = 0; j < rotHist[i].size(); ++j) { cur_match[rotHist[i][j]] = -2; nmatches--; }
    }
    delete g;
    return nmatches;
}

// src/ORBmatcher.cc:127-133
static float radius_by_viewing_cos(const float& viewCos)
{
    if (viewCos > 0.998) return 2.5;
    else return 4.0;
}

// src/ORBmatcher.cc:44-125
extern "C" int frameport_search_local_points(const frameport_local_points_args* a, int* cur_match)
{
    Grid* g = new Grid;
    build_grid(*g, a->cur_keys_un, a->n_cur, a->bounds);
    int nmatches = 0;
    const bool bFactor = a->th != 1.0;
    std::vector<uint8_t> occupied(a->cur_occupied, a->cur_occupied + a->n_cur);
    for (int i = 0; i < a->n_cur; ++i) cur_match[i] = -1;
    for (int iMP = 0; iMP < a->n_points; ++iMP) {
        if (!a->in_view[iMP]) continue;                       // !mbTrackInView || isBad()
        const int nPredictedLevel = a->scale_level[iMP];
        float r = radius_by_viewing_cos(a->view_cos[iMP]);
        if (bFactor) r *= a->th;
        const std::vector<size_t> vIndices = features_in_area(*g, a->proj_x[iMP], a->proj_y[iMP], r * a->scale_factors[nPredictedLevel],
                                                              nPredictedLevel - 1, nPredictedLevel);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (size_t c = 0; c < vIndices.size(); ++c) {
            const size_t idx = vIndices[c];
            if (occupied[idx]) continue;
            if (a->cur_u_right[idx] > 0) {
                const float er = std::fabs(a->proj_xr[iMP] - a->cur_u_right[idx]);
                if (er > r * a->scale_factors[nPredictedLevel]) continue;
            }
            const int dist = orbport_hamming256(a->mp_desc + 32 * (size_t)iMP, a->cur_desc + 32 * idx);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist;
                bestLevel2 = bestLevel; bestLevel = a->cur_keys_un[idx].octave;
                bestIdx = (int)idx;
            } else if (dist < bestDist2) {
                bestLevel2 = a->cur_keys_un[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > a->nnratio * bestDist2) continue;
            cur_match[bestIdx] = iMP;
            occupied[bestIdx] = a->mp_observed[iMP];
            nmatches++;
        }
    }
    delete g;
    return nmatches;
}

// src/ORBmatcher.cc:159-288 and :522-655 (the two overloads differ in which side carries validity and in <= / < TH_LOW)
extern "C" int frameport_search_by_bow(const frameport_bow_side* s1, const frameport_bow_side* s2, float nnratio, int check_orientation,
                                       int strict_th_low, int* match12, int* match21)
{
    for (int i = 0; i < s1->n; ++i) match12[i] = -1;
    for (int i = 0; i < s2->n; ++i) match21[i] = -1;
    std::vector<bool> vbMatched2((size_t)s2->n, false);
    std::vector<int> rotHist[HISTO_LENGTH];                    // holds idx1; the pair's idx2 is match12[idx1]
    int nmatches = 0;
    int it1 = 0, it2 = 0;
    while (it1 != s1->n_nodes && it2 != s2->n_nodes) {
        if (s1->node_ids[it1] == s2->node_ids[it2]) {
            for (int a = s1->node_start[it1]; a < s1->node_start[it1 + 1]; ++a) {
                const unsigned idx1 = s1->indices[a];
                if (!s1->valid[idx1]) continue;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int c = s2->node_start[it2]; c < s2->node_start[it2 + 1]; ++c) {
                    const unsigned idx2 = s2->indices[c];
                    if (vbMatched2[idx2] || !s2->valid[idx2]) continue;
                    const int dist = orbport_hamming256(s1->desc + 32 * (size_t)idx1, s2->desc + 32 * (size_t)idx2);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = (int)idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (strict_th_low ? bestDist1 < TH_LOW : bestDist1 <= TH_LOW) {
                    if (static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {
                        match12[idx1] = bestIdx2;
                        match21[bestIdx2] = (int)idx1;
                        vbMatched2[bestIdx2] = true;
                        if (check_orientation) rotHist[rot_bin(s1->angle[idx1], s2->angle[bestIdx2])].push_back((int)idx1);
                        nmatches++;
                    }
                }
            }
            ++it1; ++it2;
        } else if (s1->node_ids[it1] < s2->node_ids[it2]) {
            while (it1 != s1->n_nodes && s1->node_ids[it1] < s2->node_ids[it2]) ++it1;       // lower_bound
        } else {
            while (it2 != s2->n_nodes && s2->node_ids[it2] < s1->node_ids[it1]) ++it2;
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); ++j) {
                const int idx1 = rotHist[i][j];
                match21[match12[idx1]] = -1;
                match12[idx1] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// src/ORBmatcher.cc:140-157
static bool check_dist_epipolar_line(const orbport_keypoint& kp1, const orbport_keypoint& kp2, const float* F12, const float* level_sigma2)
{
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * level_sigma2[kp2.octave];
}

// src/ORBmatcher.cc:657-820
extern "C" int frameport_search_for_triangulation(const frameport_triangulation_args* a, int* matches12)
{
    const frameport_triangulation_side &s1 = a->kf1, &s2 = a->kf2;
    // epipole in the second image (:664-670): R2w*Cw+t2w through cv::gemm's 3x3 float case
    const float t2w[3] = {a->T2w[3], a->T2w[7], a->T2w[11]};
    float C2[3];
    mat3_mul_add(a->T2w, 4, a->Cw1, t2w, C2);
    const float invz = 1.0f / C2[2];
    const float ex = a->fx * C2[0] * invz + a->cx;
    const float ey = a->fy * C2[1] * invz + a->cy;
    int nmatches = 0;
    std::vector<bool> vbMatched2((size_t)s2.n, false);         // never set in the reference either
    for (int i = 0; i < s1.n; ++i) matches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    int it1 = 0, it2 = 0;
    while (it1 != s1.n_nodes && it2 != s2.n_nodes) {
        if (s1.node_ids[it1] == s2.node_ids[it2]) {
            for (int c1 = s1.node_start[it1]; c1 < s1.node_start[it1 + 1]; ++c1) {
                const size_t idx1 = s1.indices[c1];
                if (s1.has_map_point[idx1]) continue;
                const bool bStereo1 = s1.u_right[idx1] >= 0;
                if (a->only_stereo) if (!bStereo1) continue;
                const orbport_keypoint& kp1 = s1.keys_un[idx1];
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (int c2 = s2.node_start[it2]; c2 < s2.node_start[it2 + 1]; ++c2) {
                    const size_t idx2 = s2.indices[c2];
                    if (vbMatched2[idx2] || s2.has_map_point[idx2]) continue;
                    const bool bStereo2 = s2.u_right[idx2] >= 0;
                    if (a->only_stereo) if (!bStereo2) continue;
                    const int dist = orbport_hamming256(s1.desc + 32 * idx1, s2.desc + 32 * idx2);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const orbport_keypoint& kp2 = s2.keys_un[idx2];
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kp2.x;
                        const float distey = ey - kp2.y;
                        if (distex * distex + distey * distey < 100 * a->scale_factors2[kp2.octave]) continue;
                    }
                    if (check_dist_epipolar_line(kp1, kp2, a->F12, a->level_sigma2)) { bestIdx2 = (int)idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    matches12[idx1] = bestIdx2;
                    nmatches++;
                    if (a->check_orientation) rotHist[rot_bin(kp1.angle, s2.keys_un[bestIdx2].angle)].push_back((int)idx1);
                }
            }
            ++it1; ++it2;
        } else if (s1.node_ids[it1] < s2.node_ids[it2]) {
            while (it1 != s1.n_nodes && s1.node_ids[it1] < s2.node_ids[it2]) ++it1;
        } else {
            while (it2 != s2.n_nodes && s2.node_ids[it2] < s1.node_ids[it1]) ++it2;
        }
    }
    if (a->check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); ++j) { matches12[rotHist[i][j]] = -1; nmatches--; }
        }
    }
    return nmatches;
}

extern "C" void frameport_logf_array(const float* x, int n, float* out)
{
    for (int i = 0; i < n; ++i) out[i] = logf(x[i]);
}

// src/Tracking.cc:1173-1199: isInFrustum (src/Frame.cc:269-325) for the candidate points, then the search when nToMatch > 0
extern "C" int frameport_search_local_map(const frameport_local_map_args* a, const frameport_track_fields* out, int* n_in_view, int* cur_match)
{
    const float* Tc = a->Tcw;
    const float mtcw[3] = {Tc[3], Tc[7], Tc[11]};
    float mOw[3];                                             // Frame::UpdatePoseMatrices: mOw = -mRcw.t()*mtcw (src/Frame.cc:266),
    for (int i = 0; i < 3; ++i) {                             // the cv::MatExpr of src/ORBmatcher.cc:1341 (see above)
        const float t = Tc[i] * mtcw[0] + Tc[4 + i] * mtcw[1] + Tc[8 + i] * mtcw[2];
        mOw[i] = (float)((double)t * -1.0);
    }
    int nToMatch = 0;
    for (int i = 0; i < a->n_points; ++i) {
        out->in_view[i] = 0; out->proj_x[i] = out->proj_y[i] = out->proj_xr[i] = out->view_cos[i] = 0.f; out->scale_level[i] = 0;
        if (!a->candidate[i]) continue;
        const float* P = a->world_pos + 3 * (size_t)i;
        float Pc[3];
        mat3_mul_add(Tc, 4, P, mtcw, Pc);
        const float PcX = Pc[0], PcY = Pc[1], PcZ = Pc[2];
        if (PcZ < 0.0f) continue;
        const float invz = 1.0f / PcZ;
        const float u = a->fx * PcX * invz + a->cx;
        const float v = a->fy * PcY * invz + a->cy;
        if (u < a->bounds[0] || u > a->bounds[1]) continue;
        if (v < a->bounds[2] || v > a->bounds[3]) continue;
        const float maxDistance = 1.2f * a->max_distance[i];
        const float minDistance = 0.8f * a->min_distance[i];
        const float PO[3] = {P[0] - mOw[0], P[1] - mOw[1], P[2] - mOw[2]};
        double s2 = 0;                                        // cv::norm(PO): double accumulation, sqrt in double
        for (int k = 0; k < 3; ++k) s2 += (double)PO[k] * (double)PO[k];
        const float dist = std::sqrt(s2);
        if (dist < minDistance || dist > maxDistance) continue;
        const float* Pn = a->normal + 3 * (size_t)i;
        double dot = 0;                                       // cv::Mat::dot: double accumulation
        for (int k = 0; k < 3; ++k) dot += (double)PO[k] * (double)Pn[k];
        const float viewCos = dot / dist;
        if (viewCos < a->viewing_cos_limit) continue;
        // MapPoint::PredictScale (src/MapPoint.cc:402-417); the reference object calls logf
        const float ratio = a->max_distance[i] / dist;
        int nScale = ceil(logf(ratio) / a->log_scale_factor);
        if (nScale < 0) nScale = 0;
        else if (nScale >= a->nlevels) nScale = a->nlevels - 1;
        out->in_view[i] = 1;                                  // "Data used by the tracking" (src/Frame.cc:316-322)
        out->proj_x[i] = u; out->proj_y[i] = v; out->view_cos[i] = viewCos;
        out->proj_xr[i] = u - a->bf * invz;
        out->scale_level[i] = nScale;
        nToMatch++;
    }
    *n_in_view = nToMatch;
    for (int i = 0; i < a->n_cur; ++i) cur_match[i] = -1;
    if (a->n_cur == 0) return 0;
    frameport_local_points_args s;
    s.n_points = a->n_points; s.in_view = out->in_view; s.proj_x = out->proj_x; s.proj_y = out->proj_y; s.proj_xr = out->proj_xr;
    s.scale_level = out->scale_level; s.view_cos = out->view_cos; s.mp_desc = a->mp_desc; s.mp_observed = a->mp_observed;
    s.n_cur = a->n_cur; s.cur_keys_un = a->cur_keys_un; s.cur_desc = a->cur_desc; s.cur_u_right = a->cur_u_right; s.cur_occupied = a->cur_occupied;
    for (int k = 0; k < 4; ++k) s.bounds[k] = a->bounds[k];
    s.scale_factors = a->scale_factors; s.nlevels = a->nlevels; s.th = a->th; s.nnratio = a->nnratio;
    return frameport_search_local_points(&s, cur_match);
}

// glibc 2.39 logf (sysdeps/ieee754/flt-32/e_logf.c, data e_logf_data.c) restated operation by operation -- the form
// k_frustum evaluates on the device (separate multiplies and adds).  frameport_logf_port_mismatches counts the float
// bit patterns first, first + step, ... (count of them) on which it differs from the host's logf.
static const double kLogfTab[32] = {
    0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2, 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2, 0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2,
    0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3, 0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3, 0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3,
    0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4, 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4, 0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5,
    0x1p+0, 0x0p+0, 0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5, 0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4,
    0x1.b2036576afce6p-1, 0x1.526e57720db08p-3, 0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3, 0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2,
    0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2};

static float logf_port(float x)
{
    uint32_t ix; memcpy(&ix, &x, 4);
    if (ix == 0x3f800000u) return 0.f;
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
        if (ix * 2 == 0) return -INFINITY;
        if (ix == 0x7f800000u) return x;
        if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return NAN;
        const float xs = x * 0x1p23f;
        memcpy(&ix, &xs, 4);
        ix -= 23u << 23;
    }
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> 19) & 15u);
    const int k = (int32_t)tmp >> 23;
    const uint32_t iz = ix - (tmp & (0x1ffu << 23));
    float zf; memcpy(&zf, &iz, 4);
    const double invc = kLogfTab[2 * i], logc = kLogfTab[2 * i + 1], z = zf;
    volatile double m;                                         // volatile: no contraction, as __dmul_rn / __dadd_rn on the device
    m = z * invc; const double r = m + -1.0;
    m = (double)k * 0x1.62e42fefa39efp-1; const double y0 = logc + m;
    m = r * r; const double r2 = m;
    m = 0x1.5575b0be00b6ap-2 * r; double y = m + -0x1.ffffef20a4123p-2;
    m = -0x1.00ea348b88334p-2 * r2; y = m + y;
    m = y * r2; y = m + (y0 + r);
    return (float)y;
}

extern "C" long long frameport_logf_port_mismatches(uint32_t first, uint32_t count, uint32_t step)
{
    long long bad = 0;
    for (uint32_t n = 0; n < count; ++n) {
        const uint32_t b = first + n * step;
        float x; memcpy(&x, &b, 4);
        const float a = logf(x), p = logf_port(x);
        if (std::isnan(a) ? !std::isnan(p) : memcmp(&a, &p, 4) != 0) ++bad;
    }
    return bad;
}

// the loop body of ORBmatcher::Fuse up to the decision (src/ORBmatcher.cc:845-943) with KeyFrame::GetFeaturesInArea
// (src/KeyFrame.cc:569-608: the grid was filled by the Frame with its float bounds, the window subtracts the keyframe's
// int-truncated copies) and KeyFrame::IsInImage (:610-613)
extern "C" void frameport_fuse_search(const frameport_fuse_args* a, int* best_idx)
{
    Grid* g = new Grid;
    build_grid(*g, a->keys_un, a->n_kf, a->bounds);
    const int mnMinX = (int)a->bounds[0], mnMaxX = (int)a->bounds[1], mnMinY = (int)a->bounds[2], mnMaxY = (int)a->bounds[3];
    const float* Tc = a->Tcw;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]};
    for (int i = 0; i < a->n_points; ++i) {
        best_idx[i] = -1;
        if (!a->present[i]) continue;
        const float* p3Dw = a->world_pos + 3 * (size_t)i;
        float p3Dc[3];
        mat3_mul_add(Tc, 4, p3Dw, tcw, p3Dc);
        if (p3Dc[2] < 0.0f) continue;
        const float invz = 1 / p3Dc[2];
        const float x = p3Dc[0] * invz;
        const float y = p3Dc[1] * invz;
        const float u = a->fx * x + a->cx;
        const float v = a->fy * y + a->cy;
        if (!(u >= mnMinX && u < mnMaxX && v >= mnMinY && v < mnMaxY)) continue;
        const float ur = u - a->bf * invz;
        const float maxDistance = 1.2f * a->max_distance[i];
        const float minDistance = 0.8f * a->min_distance[i];
        const float PO[3] = {p3Dw[0] - a->Ow[0], p3Dw[1] - a->Ow[1], p3Dw[2] - a->Ow[2]};
        double s2 = 0;
        for (int k = 0; k < 3; ++k) s2 += (double)PO[k] * (double)PO[k];
        const float dist3D = std::sqrt(s2);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const float* Pn = a->normal + 3 * (size_t)i;
        double dot = 0;
        for (int k = 0; k < 3; ++k) dot += (double)PO[k] * (double)Pn[k];
        if (dot < 0.5 * dist3D) continue;
        const float ratio = a->max_distance[i] / dist3D;       // PredictScale(dist3D, pKF)
        int nPredictedLevel = ceil(logf(ratio) / a->log_scale_factor);
        if (nPredictedLevel < 0) nPredictedLevel = 0;
        else if (nPredictedLevel >= a->nlevels) nPredictedLevel = a->nlevels - 1;
        const float radius = a->th * a->scale_factors[nPredictedLevel];
        // KeyFrame::GetFeaturesInArea(u, v, radius)
        std::vector<size_t> vIndices;
        {
            const float r = radius;
            const int nMinCellX = std::max(0, (int)floor((u - mnMinX - r) * g->invW));
            const int nMaxCellX = std::min(GRID_COLS - 1, (int)ceil((u - mnMinX + r) * g->invW));
            const int nMinCellY = std::max(0, (int)floor((v - mnMinY - r) * g->invH));
            const int nMaxCellY = std::min(GRID_ROWS - 1, (int)ceil((v - mnMinY + r) * g->invH));
            if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0))
                for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                        const std::vector<size_t>& vCell = g->cell[ix][iy];
                        for (size_t j = 0; j < vCell.size(); j++) {
                            const orbport_keypoint& kpUn = a->keys_un[vCell[j]];
                            const float distx = kpUn.x - u, disty = kpUn.y - v;
                            if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices.push_back(vCell[j]);
                        }
                    }
        }
        if (vIndices.empty()) continue;
        int bestDist = 256, bestIdx = -1;
        for (size_t c = 0; c < vIndices.size(); ++c) {
            const size_t idx = vIndices[c];
            const orbport_keypoint& kp = a->keys_un[idx];
            const int kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (a->u_right[idx] >= 0) {
                const float ex = u - kp.x, ey = v - kp.y, er = ur - a->u_right[idx];
                const float e2 = ex * ex + ey * ey + er * er;
                if (e2 * a->inv_level_sigma2[kpLevel] > 7.8) continue;
            } else {
                const float ex = u - kp.x, ey = v - kp.y;
                const float e2 = ex * ex + ey * ey;
                if (e2 * a->inv_level_sigma2[kpLevel] > 5.99) continue;
            }
            const int dist = orbport_hamming256(a->mp_desc + 32 * (size_t)i, a->desc + 32 * idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_LOW) best_idx[i] = bestIdx;
    }
    delete g;
}

// src/ORBmatcher.cc:1472-1597
extern "C" int frameport_search_by_projection_keyframe(const frameport_reloc_projection_args* a, int* cur_match)
{
    Grid* g = new Grid;
    build_grid(*g, a->cur_keys_un, a->n_cur, a->bounds);
    int nmatches = 0;
    const float* Tc = a->Tcw_cur;
    const float tcw[3] = {Tc[3], Tc[7], Tc[11]};
    float Ow[3];                                              // -Rcw.t()*tcw (:1478), lowered as at :1341
    for (int i = 0; i < 3; ++i) {
        const float t = Tc[i] * tcw[0] + Tc[4 + i] * tcw[1] + Tc[8 + i] * tcw[2];
        Ow[i] = (float)((double)t * -1.0);
    }
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<uint8_t> occupied(a->cur_occupied, a->cur_occupied + a->n_cur);
    for (int i = 0; i < a->n_cur; ++i) cur_match[i] = -1;
    for (int i = 0; i < a->n_kf; ++i) {
        if (!a->mp_valid[i]) continue;
        const float* x3Dw = a->world_pos + 3 * (size_t)i;
        float x3Dc[3];
        mat3_mul_add(Tc, 4, x3Dw, tcw, x3Dc);
        const float xc = x3Dc[0], yc = x3Dc[1];
        const float invzc = 1.0 / x3Dc[2];
        const float u = a->fx * xc * invzc + a->cx;
        const float v = a->fy * yc * invzc + a->cy;
        if (u < a->bounds[0] || u > a->bounds[1]) continue;
        if (v < a->bounds[2] || v > a->bounds[3]) continue;
        const float PO[3] = {x3Dw[0] - Ow[0], x3Dw[1] - Ow[1], x3Dw[2] - Ow[2]};
        double s2 = 0;
        for (int k = 0; k < 3; ++k) s2 += (double)PO[k] * (double)PO[k];
        float dist3D = std::sqrt(s2);
        const float maxDistance = 1.2f * a->max_distance[i];
        const float minDistance = 0.8f * a->min_distance[i];
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const float ratio = a->max_distance[i] / dist3D;
        int nPredictedLevel = ceil(logf(ratio) / a->log_scale_factor);
        if (nPredictedLevel < 0) nPredictedLevel = 0;
        else if (nPredictedLevel >= a->nlevels) nPredictedLevel = a->nlevels - 1;
        const float radius = a->th * a->scale_factors[nPredictedLevel];
        const std::vector<size_t> vIndices2 = features_in_area(*g, u, v, radius, nPredictedLevel - 1, nPredictedLevel + 1);
        if (vIndices2.empty()) continue;
        int bestDist = 256, bestIdx2 = -1;
        for (size_t c = 0; c < vIndices2.size(); ++c) {
            const size_t i2 = vIndices2[c];
            if (occupied[i2]) continue;
            const int dist = orbport_hamming256(a->mp_desc + 32 * (size_t)i, a->cur_desc + 32 * i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = (int)i2; }
        }
        if (bestDist <= a->orb_dist) {
            occupied[bestIdx2] = 1;
            cur_match[bestIdx2] = i;
            nmatches++;
            if (a->check_orientation) rotHist[rot_bin(a->kf_angle[i], a->cur_keys_un[bestIdx2].angle)].push_back(bestIdx2);
        }
    }
    if (a->check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i)
            if (i != ind1 && i != ind2 && i != ind3)
                for (size_t j = 0; j < rotHist[i].size(); ++j) { cur_match[rotHist[i][j]] = -2; nmatches--; }
    }
    delete g;
    return nmatches;
}

// src/MapPoint.cc:272-303
extern "C" void frameport_distinctive_descriptors(const uint8_t* desc, const int32_t* group_start, int n_groups, int* best_idx)
{
    for (int g = 0; g < n_groups; ++g) {
        const uint8_t* d = desc + 32 * (size_t)group_start[g];
        const size_t N = (size_t)(group_start[g + 1] - group_start[g]);
        best_idx[g] = -1;
        if (N == 0) continue;
        std::vector<float> Distances(N * N);
        for (size_t i = 0; i < N; i++) {
            Distances[i * N + i] = 0;
            for (size_t j = i + 1; j < N; j++) {
                const int distij = orbport_hamming256(d + 32 * i, d + 32 * j);
                Distances[i * N + j] = distij;
                Distances[j * N + i] = distij;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + (i + 1) * N);
            std::sort(vDists.begin(), vDists.end());
            const int median = vDists[0.5 * (N - 1)];
            if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
        }
        best_idx[g] = BestIdx;
    }
}

namespace {

// KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:569-608) on the grid the Frame built; the window subtracts the keyframe's
// int-truncated bounds
std::vector<size_t> kf_features_in_area(const Grid& g, const orbport_keypoint* keys, int mnMinX, int mnMinY, float x, float y, float r)
{
    std::vector<size_t> v;
    const int nMinCellX = std::max(0, (int)floor((x - mnMinX - r) * g.invW));
    if (nMinCellX >= GRID_COLS) return v;
    const int nMaxCellX = std::min(GRID_COLS - 1, (int)ceil((x - mnMinX + r) * g.invW));
    if (nMaxCellX < 0) return v;
    const int nMinCellY = std::max(0, (int)floor((y - mnMinY - r) * g.invH));
    if (nMinCellY >= GRID_ROWS) return v;
    const int nMaxCellY = std::min(GRID_ROWS - 1, (int)ceil((y - mnMinY + r) * g.invH));
    if (nMaxCellY < 0) return v;
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const std::vector<size_t>& vCell = g.cell[ix][iy];
            for (size_t j = 0; j < vCell.size(); j++) {
                const orbport_keypoint& kpUn = keys[vCell[j]];
                const float distx = kpUn.x - x, disty = kpUn.y - y;
                if (std::fabs(distx) < r && std::fabs(disty) < r) v.push_back(vCell[j]);
            }
        }
    return v;
}

// "Decompose Scw" (src/ORBmatcher.cc:299-303, :979-984)
void decompose_sim3(const float* S, float* Rcw9, float* tcw, float* Ow)
{
    double d = 0;                                             // sRcw.row(0).dot(sRcw.row(0)): double accumulation
    for (int c = 0; c < 3; ++c) d += (double)S[c] * (double)S[c];
    const float scw = sqrt(d);
    const float inv = (float)(1.0 / scw);                     // "M / scw" = convertTo with scale 1/scw: element * (float)scale, in float
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) Rcw9[3 * r + c] = S[4 * r + c] * inv;
        tcw[r] = S[4 * r + 3] * inv;
    }
    for (int i = 0; i < 3; ++i) {                             // -Rcw.t()*tcw, lowered as at :1341
        const float t = Rcw9[i] * tcw[0] + Rcw9[3 + i] * tcw[1] + Rcw9[6 + i] * tcw[2];
        Ow[i] = (float)((double)t * -1.0);
    }
}

int predict_scale(float max_distance, float dist, float log_scale, int nlevels)      // src/MapPoint.cc:385-400
{
    const float ratio = max_distance / dist;
    int nScale = ceil(logf(ratio) / log_scale);
    if (nScale < 0) nScale = 0;
    else if (nScale >= nlevels) nScale = nlevels - 1;
    return nScale;
}

}  // namespace

// src/ORBmatcher.cc:969-1097 up to the decision
extern "C" void frameport_fuse_search_sim3(const frameport_fuse_args* a, const float* Scw, int* best_idx)
{
    Grid* g = new Grid;
    build_grid(*g, a->keys_un, a->n_kf, a->bounds);
    const int mnMinX = (int)a->bounds[0], mnMaxX = (int)a->bounds[1], mnMinY = (int)a->bounds[2], mnMaxY = (int)a->bounds[3];
    float Rcw[9], tcw[3], Ow[3];
    decompose_sim3(Scw, Rcw, tcw, Ow);
    for (int iMP = 0; iMP < a->n_points; ++iMP) {
        best_idx[iMP] = -1;
        if (!a->present[iMP]) continue;
        const float* p3Dw = a->world_pos + 3 * (size_t)iMP;
        float p3Dc[3];
        mat3_mul_add(Rcw, 3, p3Dw, tcw, p3Dc);
        if (p3Dc[2] < 0.0f) continue;
        const float invz = 1.0 / p3Dc[2];
        const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
        const float u = a->fx * x + a->cx, v = a->fy * y + a->cy;
        if (!(u >= mnMinX && u < mnMaxX && v >= mnMinY && v < mnMaxY)) continue;
        const float maxDistance = 1.2f * a->max_distance[iMP], minDistance = 0.8f * a->min_distance[iMP];
        const float PO[3] = {p3Dw[0] - Ow[0], p3Dw[1] - Ow[1], p3Dw[2] - Ow[2]};
        double s2 = 0;
        for (int k = 0; k < 3; ++k) s2 += (double)PO[k] * (double)PO[k];
        const float dist3D = std::sqrt(s2);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const float* Pn = a->normal + 3 * (size_t)iMP;
        double dot = 0;
        for (int k = 0; k < 3; ++k) dot += (double)PO[k] * (double)Pn[k];
        if (dot < 0.5 * dist3D) continue;
        const int nPredictedLevel = predict_scale(a->max_distance[iMP], dist3D, a->log_scale_factor, a->nlevels);
        const float radius = a->th * a->scale_factors[nPredictedLevel];
        const std::vector<size_t> vIndices = kf_features_in_area(*g, a->keys_un, mnMinX, mnMinY, u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = INT_MAX, bestIdx = -1;
        for (size_t c = 0; c < vIndices.size(); ++c) {
            const size_t idx = vIndices[c];
            const int kpLevel = a->keys_un[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int dist = orbport_hamming256(a->mp_desc + 32 * (size_t)iMP, a->desc + 32 * idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_LOW) best_idx[iMP] = bestIdx;
    }
    delete g;
}

// src/ORBmatcher.cc:290-403
extern "C" int frameport_search_by_projection_sim3(const frameport_fuse_args* a, const float* Scw, const uint8_t* matched_in, int* matched)
{
    Grid* g = new Grid;
    build_grid(*g, a->keys_un, a->n_kf, a->bounds);
    const int mnMinX = (int)a->bounds[0], mnMaxX = (int)a->bounds[1], mnMinY = (int)a->bounds[2], mnMaxY = (int)a->bounds[3];
    float Rcw[9], tcw[3], Ow[3];
    decompose_sim3(Scw, Rcw, tcw, Ow);
    std::vector<uint8_t> vpMatched(matched_in, matched_in + a->n_kf);
    for (int i = 0; i < a->n_kf; ++i) matched[i] = -1;
    int nmatches = 0;
    for (int iMP = 0; iMP < a->n_points; ++iMP) {
        if (!a->present[iMP]) continue;
        const float* p3Dw = a->world_pos + 3 * (size_t)iMP;
        float p3Dc[3];
        mat3_mul_add(Rcw, 3, p3Dw, tcw, p3Dc);
        if (p3Dc[2] < 0.0) continue;
        const float invz = 1 / p3Dc[2];
        const float x = p3Dc[0] * invz, y = p3Dc[1] * invz;
        const float u = a->fx * x + a->cx, v = a->fy * y + a->cy;
        if (!(u >= mnMinX && u < mnMaxX && v >= mnMinY && v < mnMaxY)) continue;
        const float maxDistance = 1.2f * a->max_distance[iMP], minDistance = 0.8f * a->min_distance[iMP];
        const float PO[3] = {p3Dw[0] - Ow[0], p3Dw[1] - Ow[1], p3Dw[2] - Ow[2]};
        double s2 = 0;
        for (int k = 0; k < 3; ++k) s2 += (double)PO[k] * (double)PO[k];
        const float dist = std::sqrt(s2);
        if (dist < minDistance || dist > maxDistance) continue;
        const float* Pn = a->normal + 3 * (size_t)iMP;
        double dot = 0;
        for (int k = 0; k < 3; ++k) dot += (double)PO[k] * (double)Pn[k];
        if (dot < 0.5 * dist) continue;
        const int nPredictedLevel = predict_scale(a->max_distance[iMP], dist, a->log_scale_factor, a->nlevels);
        const float radius = (int)a->th * a->scale_factors[nPredictedLevel];
        const std::vector<size_t> vIndices = kf_features_in_area(*g, a->keys_un, mnMinX, mnMinY, u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = 256, bestIdx = -1;
        for (size_t c = 0; c < vIndices.size(); ++c) {
            const size_t idx = vIndices[c];
            if (vpMatched[idx]) continue;
            const int kpLevel = a->keys_un[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int d = orbport_hamming256(a->mp_desc + 32 * (size_t)iMP, a->desc + 32 * idx);
            if (d < bestDist) { bestDist = d; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_LOW) { vpMatched[bestIdx] = 1; matched[bestIdx] = iMP; nmatches++; }
    }
    delete g;
    return nmatches;
}

namespace {
// one direction of SearchBySim3 (src/ORBmatcher.cc:1147-1225 / :1227-1305)
void sim3_pass(const frameport_fuse_args* a, const float* sR, const float* tt, std::vector<int>& vnMatch)
{
    Grid* g = new Grid;
    build_grid(*g, a->keys_un, a->n_kf, a->bounds);
    const int mnMinX = (int)a->bounds[0], mnMaxX = (int)a->bounds[1], mnMinY = (int)a->bounds[2], mnMaxY = (int)a->bounds[3];
    const float* T = a->Tcw;
    const float tw[3] = {T[3], T[7], T[11]};
    vnMatch.assign((size_t)a->n_points, -1);
    for (int i = 0; i < a->n_points; ++i) {
        if (!a->present[i]) continue;
        float pA[3], pB[3];
        mat3_mul_add(T, 4, a->world_pos + 3 * (size_t)i, tw, pA);
        mat3_mul_add(sR, 3, pA, tt, pB);
        if (pB[2] < 0.0) continue;
        const float invz = 1.0 / pB[2];
        const float x = pB[0] * invz, y = pB[1] * invz;
        const float u = a->fx * x + a->cx, v = a->fy * y + a->cy;
        if (!(u >= mnMinX && u < mnMaxX && v >= mnMinY && v < mnMaxY)) continue;
        const float maxDistance = 1.2f * a->max_distance[i], minDistance = 0.8f * a->min_distance[i];
        double s2 = 0;
        for (int k = 0; k < 3; ++k) s2 += (double)pB[k] * (double)pB[k];
        const float dist3D = std::sqrt(s2);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int nPredictedLevel = predict_scale(a->max_distance[i], dist3D, a->log_scale_factor, a->nlevels);
        const float radius = a->th * a->scale_factors[nPredictedLevel];
        const std::vector<size_t> vIndices = kf_features_in_area(*g, a->keys_un, mnMinX, mnMinY, u, v, radius);
        if (vIndices.empty()) continue;
        int bestDist = INT_MAX, bestIdx = -1;
        for (size_t c = 0; c < vIndices.size(); ++c) {
            const size_t idx = vIndices[c];
            const int oct = a->keys_un[idx].octave;
            if (oct < nPredictedLevel - 1 || oct > nPredictedLevel) continue;
            const int dist = orbport_hamming256(a->mp_desc + 32 * (size_t)i, a->desc + 32 * idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = (int)idx; }
        }
        if (bestDist <= TH_HIGH) vnMatch[i] = bestIdx;
    }
    delete g;
}
}  // namespace

// src/ORBmatcher.cc:1102-1326
extern "C" int frameport_search_by_sim3(const frameport_fuse_args* a12, const frameport_fuse_args* a21, float s12, const float* R12,
                                        const float* t12, int* match12)
{
    float sR12[9], sR21[9], t21[3];
    const float inv = (float)(1.0 / s12);                     // (1.0/s12)*R12.t(): transpose, then convertTo with that scale
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) { sR12[3 * r + c] = R12[3 * r + c] * s12; sR21[3 * r + c] = R12[3 * c + r] * inv; }
    for (int i = 0; i < 3; ++i) {                             // -sR21*t12: gemm with alpha = -1
        const float t = sR21[3 * i] * t12[0] + sR21[3 * i + 1] * t12[1] + sR21[3 * i + 2] * t12[2];
        t21[i] = (float)((double)t * -1.0);
    }
    std::vector<int> vnMatch1, vnMatch2;
    sim3_pass(a12, sR21, t21, vnMatch1);
    sim3_pass(a21, sR12, t12, vnMatch2);
    int nFound = 0;
    for (int i1 = 0; i1 < a12->n_points; ++i1) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) { match12[i1] = idx2; nFound++; }
        }
    }
    return nFound;
}

// MapPoint::UpdateNormalAndDepth (src/MapPoint.cc:330-371) in OpenCV's arithmetic: "mWorldPos - Owi" in float; cv::norm
// accumulates squares in double; "normal + normali / norm" is MatOp_AddEx -> cv::scaleAdd(normali, 1.0 / norm, normal) =
// fmaf(v, (float)(1.0 / norm), acc) (pinned: tests/golden/scale_add_golden.npz); "normal / n" is v * (float)(1.0 / n).
extern "C" void frameport_update_normal_and_depth(const float* pos, const int32_t* obs_start, const float* centers,
                                                  const float* ref_center, const float* ref_scale, const float* last_scale,
                                                  int n_points, float* normal, float* max_dist, float* min_dist)
{
    for (int g = 0; g < n_points; ++g) {
        const int b = obs_start[g], e = obs_start[g + 1];
        if (e <= b) continue;                                                     // observations.empty(): return (:345-346)
        const float* P = pos + 3 * (size_t)g;
        float acc[3] = {0.f, 0.f, 0.f};
        for (int j = b; j < e; ++j) {
            const float* O = centers + 3 * (size_t)j;
            const float v[3] = {P[0] - O[0], P[1] - O[1], P[2] - O[2]};
            const double nrm = std::sqrt((double)v[0] * (double)v[0] + (double)v[1] * (double)v[1] + (double)v[2] * (double)v[2]);
            const float a = (float)(1.0 / nrm);
            for (int k = 0; k < 3; ++k) acc[k] = std::fmaf(v[k], a, acc[k]);
        }
        const float* R = ref_center + 3 * (size_t)g;
        const float pc[3] = {P[0] - R[0], P[1] - R[1], P[2] - R[2]};
        const float dist = (float)std::sqrt((double)pc[0] * (double)pc[0] + (double)pc[1] * (double)pc[1] + (double)pc[2] * (double)pc[2]);
        max_dist[g] = dist * ref_scale[g];
        min_dist[g] = max_dist[g] / last_scale[g];
        const float inv_n = (float)(1.0 / (double)(e - b));
        for (int k = 0; k < 3; ++k) normal[3 * (size_t)g + k] = acc[k] * inv_n;
    }
}
