// orb_stereo.cu -- Frame::ComputeStereoMatches (src/Frame.cc:466-640) as one kernel: one warp per LEFT keypoint.
//
//   1. row-band descriptor search (:502-549): right keypoints whose row band [floor(y - r), ceil(y + r)],
//      r = 2 * scale[octave], covers the left keypoint's row, within one octave, with uR in [uL - bf/b, uL];
//      best Hamming distance below TH_HIGH, earliest right index on ties (the reference visits the row's
//      candidate list in right-keypoint order with a strict '<');
//   2. if best < (TH_HIGH + TH_LOW) / 2: 11 x 11 SAD between the centre-subtracted patches of the unblurred
//      pyramid level of the left keypoint, right patch slid over -5..+5 px (:551-592), parabola fit (:597-605);
//   3. disparity / depth (:607-622).  The median-based outlier rejection (:626-639) needs a global sort of the
//      accepted matches and is finished on the host by orbgpu_stereo_match.
// Float steps use explicit round-to-nearest operations in the reference's order (no FMA contraction).
#include <cuda_runtime.h>
#include <limits.h>
#include <math.h>
#include <stdint.h>
#include "orb_common.cuh"

struct StereoKp { float x, y, size, angle, response; int octave, class_id; };

__device__ __forceinline__ int st_reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) i = i < 0 ? -i : 2 * (n - 1) - i;
    return i;
}

// pixel (x, y) of a level as the reference's bordered cv::Mat would return it (REFLECT_101 frame)
__device__ __forceinline__ float st_pix(const OrbStereoSide& S, int l, int x, int y)
{
    return (float)__ldg(S.lvl[l] + (long long)st_reflect101(y, S.h[l]) * S.pitch[l] + st_reflect101(x, S.w[l]));
}

__device__ __forceinline__ int st_hamming(const uint4& a0, const uint4& a1, const uint4* __restrict__ b)
{
    const uint4 b0 = __ldg(b), b1 = __ldg(b + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(256) k_stereo_match(OrbStereoArgs A)
{
    const int lane = threadIdx.x & 31;
    const int iL = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int nL = *A.L.n_ptr, nR = *A.R.n_ptr;
    if (blockIdx.x == 0 && threadIdx.x == 0) { A.counts[0] = nL; A.counts[1] = nR; }
    if (iL >= nL) return;
    const StereoKp* kL = reinterpret_cast<const StereoKp*>(A.L.kps);
    const StereoKp* kR = reinterpret_cast<const StereoKp*>(A.R.kps);
    const float uL = kL[iL].x, vL = kL[iL].y;
    const int levelL = kL[iL].octave;
    float out_u = -1.0f, out_d = -1.0f;
    int out_sad = -1;

    const float maxD = __fdiv_rn(A.bf, A.b);                  // mbf / minZ, minZ = mb (:497-499)
    const float minU = __fsub_rn(uL, maxD), maxU = uL;        // minD = 0
    const int rowL = (int)vL;                                  // vRowIndices[vL]: float -> index truncation (:512)
    unsigned best = ((unsigned)A.th_high << 16) | 0xFFFFu;     // (distance, right index) packed; lowest index wins ties
    if (!(maxU < 0.f)) {
        const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(A.L.desc) + 2 * (size_t)iL);
        const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(A.L.desc) + 2 * (size_t)iL + 1);
        for (int iR = lane; iR < nR; iR += 32) {
            const StereoKp k = kR[iR];
            const float r = __fmul_rn(2.0f, A.scale[k.octave]);
            const int maxr = (int)ceilf(__fadd_rn(k.y, r)), minr = (int)floorf(__fsub_rn(k.y, r));
            if (rowL < minr || rowL > maxr) continue;
            if (k.octave < levelL - 1 || k.octave > levelL + 1) continue;
            if (!(k.x >= minU && k.x <= maxU)) continue;
            const unsigned d = (unsigned)st_hamming(a0, a1, reinterpret_cast<const uint4*>(A.R.desc) + 2 * (size_t)iR);
            if (d < (unsigned)A.th_high) best = min(best, (d << 16) | (unsigned)iR);
        }
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
    const int bestDist = (int)(best >> 16);

    if (bestDist < A.th_orb_dist) {
        const int bestIdxR = (int)(best & 0xFFFFu);
        const float uR0 = kR[bestIdxR].x;
        const float sf = A.inv_scale[levelL];
        const int su = (int)roundf(__fmul_rn(uL, sf)), sv = (int)roundf(__fmul_rn(vL, sf));
        const int su0 = (int)roundf(__fmul_rn(uR0, sf));
        const int w = 5, Lw = 5;
        // iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1 (:572-575)
        if (!(su0 + Lw - w < 0 || su0 + Lw + w + 1 >= A.R.w[levelL])) {
            // lane k < 11 evaluates shift incR = k - 5: SAD of the two centre-subtracted 11 x 11 patches
            float dist = 0.f;
            if (lane < 2 * Lw + 1) {
                const int inc = lane - Lw;
                const float cL = st_pix(A.L, levelL, su, sv), cR = st_pix(A.R, levelL, su0 + inc, sv);
                for (int dy = -w; dy <= w; ++dy)
                    for (int dx = -w; dx <= w; ++dx) {
                        const float a = __fsub_rn(st_pix(A.L, levelL, su + dx, sv + dy), cL);
                        const float b = __fsub_rn(st_pix(A.R, levelL, su0 + inc + dx, sv + dy), cR);
                        dist = __fadd_rn(dist, fabsf(__fsub_rn(a, b)));
                    }
            }
            // minimum over the 11 shifts, first shift wins ties (strict '<' while incR ascends)
            unsigned key = lane < 2 * Lw + 1 ? (((unsigned)(int)dist) << 8) | (unsigned)lane : 0xFFFFFFFFu;
#pragma unroll
            for (int s = 16; s > 0; s >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, s));
            const int bi = (int)(key & 0xFFu);                 // index 0..10 of the best shift
            const int sadBest = (int)(key >> 8);
            const float d1 = __shfl_sync(0xffffffffu, dist, max(bi - 1, 0));
            const float d2 = __shfl_sync(0xffffffffu, dist, bi);
            const float d3 = __shfl_sync(0xffffffffu, dist, min(bi + 1, 2 * Lw));
            if (bi != 0 && bi != 2 * Lw) {
                const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2)));
                const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), den);
                if (!(deltaR < -1.f || deltaR > 1.f)) {
                    float bestuR = __fmul_rn(A.scale[levelL], __fadd_rn(__fadd_rn((float)su0, (float)(bi - Lw)), deltaR));
                    float disparity = __fsub_rn(uL, bestuR);
                    if (disparity >= 0.f && disparity < maxD) {
                        if (disparity <= 0.f) {
                            disparity = 0.01f;
                            bestuR = (float)((double)uL - 0.01);
                        }
                        out_d = __fdiv_rn(A.bf, disparity);
                        out_u = bestuR;
                        out_sad = sadBest;
                    }
                }
            }
        }
    }
    if (lane == 0) { A.u_right[iL] = out_u; A.depth[iL] = out_d; A.sad[iL] = out_sad; }
}

int orb_launch_stereo(const OrbStereoArgs& A, int max_left, cudaStream_t st)
{
    if (max_left < 1) return 0;
    k_stereo_match<<<(max_left + 7) / 8, 256, 0, st>>>(A);
    return 1;
}
