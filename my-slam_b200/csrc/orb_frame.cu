// orb_frame.cu -- Frame-side steps that sit between the extractor and the matcher.
//
//   k_undistort_points / k_undistort_keypoints   cv::undistortPoints(src, dst, K, D, noArray(), K) of OpenCV 4.13 as
//       Frame::UndistortKeyPoints and Frame::ComputeImageBounds call it (src/Frame.cc:404-463): per point, 5 fixed
//       iterations of the inverse distortion model in double, re-projection with P = K, result narrowed to float.
//       Every double step is an explicit round-to-nearest operation in OpenCV's expression order (no FMA contraction).
#include <cuda_runtime.h>
#include <stdint.h>
#include "orb_common.cuh"

struct UndistortCam { double fx, fy, cx, cy, ifx, ify, k[12]; };

#define DM(a, b) __dmul_rn((a), (b))
#define DA(a, b) __dadd_rn((a), (b))
#define DS(a, b) __dsub_rn((a), (b))

__device__ __forceinline__ float2 undistort_one(const UndistortCam& C, float uf, float vf)
{
    const double u = (double)uf, v = (double)vf;
    const double* k = C.k;
    double x = DM(DS(u, C.cx), C.ifx), y = DM(DS(v, C.cy), C.ify);
    const double x0 = x, y0 = y;
#pragma unroll 1
    for (int j = 0; j < 5; ++j) {
        const double r2 = DA(DM(x, x), DM(y, y));
        const double num = DA(1.0, DM(DA(DM(DA(DM(k[7], r2), k[6]), r2), k[5]), r2));
        const double den = DA(1.0, DM(DA(DM(DA(DM(k[4], r2), k[1]), r2), k[0]), r2));
        const double icdist = __ddiv_rn(num, den);
        if (icdist < 0) { x = DM(DS(u, C.cx), C.ifx); y = DM(DS(v, C.cy), C.ify); break; }
        // deltaX = 2*k2*x*y + k3*(r2 + 2*x*x) + k8*r2 + k9*r2*r2, summed left to right
        const double dX = DA(DA(DA(DM(DM(DM(2.0, k[2]), x), y), DM(k[3], DA(r2, DM(DM(2.0, x), x)))), DM(k[8], r2)), DM(DM(k[9], r2), r2));
        const double dY = DA(DA(DA(DM(k[2], DA(r2, DM(DM(2.0, y), y))), DM(DM(DM(2.0, k[3]), x), y)), DM(k[10], r2)), DM(DM(k[11], r2), r2));
        x = DM(DS(x0, dX), icdist);
        y = DM(DS(y0, dY), icdist);
    }
    // RR = P * I with P = K: rows (fx, 0, cx), (0, fy, cy), (0, 0, 1)
    const double xx = DA(DA(DM(C.fx, x), DM(0.0, y)), C.cx);
    const double yy = DA(DA(DM(0.0, x), DM(C.fy, y)), C.cy);
    const double ww = __ddiv_rn(1.0, DA(DA(DM(0.0, x), DM(0.0, y)), 1.0));
    return make_float2((float)DM(xx, ww), (float)DM(yy, ww));
}

__global__ void k_undistort_points(UndistortCam C, const float2* __restrict__ src, float2* __restrict__ dst, int n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const float2 p = src[i]; dst[i] = undistort_one(C, p.x, p.y); }
}

// 28-byte keypoint records: copy the record, replace pt (src/Frame.cc:425-432)
__global__ void k_undistort_keypoints(UndistortCam C, const float* __restrict__ src, float* __restrict__ dst, const int* __restrict__ n_ptr, int identity)
{
    const int n = *n_ptr;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float* s = src + 7 * (size_t)i;
    float* d = dst + 7 * (size_t)i;
    float2 p = make_float2(s[0], s[1]);
    if (!identity) p = undistort_one(C, p.x, p.y);
    d[0] = p.x; d[1] = p.y;
#pragma unroll
    for (int f = 2; f < 7; ++f) d[f] = s[f];
}

static UndistortCam make_cam(const float* K4, const float* dist, int ndist)
{
    UndistortCam C;
    C.fx = K4[0]; C.fy = K4[1]; C.cx = K4[2]; C.cy = K4[3];
    C.ifx = 1. / C.fx; C.ify = 1. / C.fy;
    for (int i = 0; i < 12; ++i) C.k[i] = i < ndist ? (double)dist[i] : 0.0;
    return C;
}

int orb_launch_undistort_points(const float* K4, const float* dist, int ndist, const float* d_src, float* d_dst, int n, cudaStream_t st)
{
    if (n < 1) return 0;
    k_undistort_points<<<(n + 127) / 128, 128, 0, st>>>(make_cam(K4, dist, ndist), reinterpret_cast<const float2*>(d_src),
                                                          reinterpret_cast<float2*>(d_dst), n);
    return 1;
}

int orb_launch_undistort_keypoints(const float* K4, const float* dist, int ndist, const void* d_kps, void* d_kps_un, const int* d_n,
                                   int capacity, cudaStream_t st)
{
    if (capacity < 1) return 0;
    const int identity = dist[0] == 0.0f;                      // mDistCoef.at<float>(0) == 0.0: mvKeysUn = mvKeys (src/Frame.cc:406-410)
    k_undistort_keypoints<<<(capacity + 127) / 128, 128, 0, st>>>(make_cam(K4, dist, ndist), reinterpret_cast<const float*>(d_kps),
                                                                  reinterpret_cast<float*>(d_kps_un), d_n, identity);
    return 1;
}
