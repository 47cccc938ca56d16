// ORBmatcher.cc -- host side of the Hamming drop-in (see ORBmatcher.h); kernels are in liborbgpu.so.
#include "ORBmatcher.h"

#include <cstring>

#include "../../include/orbgpu.h"

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;      // src/ORBmatcher.cc:37
const int ORBmatcher::TH_LOW = 50;        // :38
const int ORBmatcher::HISTO_LENGTH = 30;  // :39

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri), mpMatcher(0) {}

ORBmatcher::~ORBmatcher() { orbgpu_matcher_destroy(mpMatcher); }

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b)
{
    return orbgpu_hamming256(a.ptr(0), b.ptr(0));          // 256-bit XOR + hardware popcount
}

// dense nRows x 32 copy of a descriptor matrix (rows may be strided)
static const unsigned char* dense_rows(const cv::Mat& m, std::vector<unsigned char>& tmp)
{
    if (m.isContinuous()) return m.ptr(0);
    tmp.resize((size_t)m.rows * 32);
    for (int r = 0; r < m.rows; ++r) std::memcpy(&tmp[(size_t)32 * r], m.ptr(r), 32);
    return tmp.data();
}

bool ORBmatcher::BestTwo(const cv::Mat& A, const cv::Mat& B, std::vector<int>& bestIdx, std::vector<int>& bestDist,
                         std::vector<int>& secondDist, int cudaDevice)
{
    const int nA = A.rows, nB = B.rows;
    bestIdx.assign((size_t)nA, -1); bestDist.assign((size_t)nA, 256); secondDist.assign((size_t)nA, 256);
    if (nA == 0) return true;
    if (A.cols != 32 || (nB > 0 && B.cols != 32)) return false;
    std::vector<unsigned char> ta, tb;
    std::vector<int32_t> idx((size_t)nA);
    std::vector<uint16_t> b1((size_t)nA), b2((size_t)nA);
    static const unsigned char none[32] = {0};
    const int rc = orbgpu_match_bruteforce(cudaDevice, dense_rows(A, ta), nA, nB > 0 ? dense_rows(B, tb) : none, nB,
                                           idx.data(), b1.data(), b2.data());
    if (rc != ORBGPU_OK) return false;
    for (int i = 0; i < nA; ++i) { bestIdx[i] = idx[i]; bestDist[i] = b1[i]; secondDist[i] = b2[i]; }
    return true;
}

int ORBmatcher::SearchBruteForce(const cv::Mat& A, const cv::Mat& B, std::vector<int>& vnMatches12, int cudaDevice)
{
    std::vector<int> idx, d1, d2;
    if (!BestTwo(A, B, idx, d1, d2, cudaDevice)) return -1;
    vnMatches12.assign(idx.size(), -1);
    int nmatches = 0;
    for (size_t i = 0; i < idx.size(); ++i)
        if (d1[i] <= TH_LOW && (float)d1[i] < mfNNratio * (float)d2[i]) { vnMatches12[i] = idx[i]; ++nmatches; }
    return nmatches;
}

bool ORBmatcher::SetKeyFrames(const std::vector<cv::Mat>& descriptors, int cudaDevice)
{
    if (descriptors.empty()) return false;
    const int n = descriptors[0].rows;
    if (n < 1) return false;
    std::vector<unsigned char> all((size_t)descriptors.size() * n * 32);
    for (size_t k = 0; k < descriptors.size(); ++k) {
        if (descriptors[k].rows != n || descriptors[k].cols != 32) return false;
        for (int r = 0; r < n; ++r) std::memcpy(&all[(k * n + r) * 32], descriptors[k].ptr(r), 32);
    }
    if (!mpMatcher && orbgpu_matcher_create(&mpMatcher, cudaDevice) != ORBGPU_OK) return false;
    return orbgpu_matcher_set_keyframes(mpMatcher, all.data(), (int)descriptors.size(), n) == ORBGPU_OK;
}

bool ORBmatcher::CountPairs(const std::vector<int>& a, const std::vector<int>& b, std::vector<int>& counts)
{
    if (!mpMatcher || a.size() != b.size()) return false;
    counts.assign(a.size(), 0);
    if (a.empty()) return true;
    std::vector<int32_t> out(a.size());
    const int rc = orbgpu_matcher_count_pairs(mpMatcher, a.data(), b.data(), (int)a.size(), TH_LOW, mfNNratio, out.data());
    if (rc != ORBGPU_OK) return false;
    for (size_t i = 0; i < a.size(); ++i) counts[i] = out[i];
    return true;
}

// Three fullest histogram bins; the first bin reaching a size wins ties (strict '>' while scanning
// upwards); the 2nd / 3rd are dropped when they hold less than a tenth of the fullest.
void ORBmatcher::ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3)
{
    int top[3] = {0, 0, 0};
    int* ind[3] = {&ind1, &ind2, &ind3};
    for (int i = 0; i < L; ++i) {
        const int s = (int)histo[i].size();
        int rank = 0;
        while (rank < 3 && s <= top[rank]) ++rank;
        if (rank == 3) continue;
        for (int k = 2; k > rank; --k) { top[k] = top[k - 1]; *ind[k] = *ind[k - 1]; }
        top[rank] = s; *ind[rank] = i;
    }
    if ((float)top[1] < 0.1f * (float)top[0]) { ind2 = -1; ind3 = -1; }
    else if ((float)top[2] < 0.1f * (float)top[0]) { ind3 = -1; }
}

}  // namespace ORB_SLAM2
