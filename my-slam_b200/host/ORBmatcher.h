// ORBmatcher.h -- the Hamming part of the reference's include/ORBmatcher.h:37-102 as a drop-in.
//
// What is replaced: the static DescriptorDistance (src/ORBmatcher.cc:1647-1663, called from 15 sites in
// ORBmatcher.cc plus src/MapPoint.cc:281 and src/Frame.cc:541), the constants TH_LOW / TH_HIGH /
// HISTO_LENGTH (:37-39), the constructor, ComputeThreeMaxima (:1601-1642, host, as written) and the
// best / second-best scan (:201-226) in batched form on the GPU.  The ten Search* / Fuse members take
// Frame / KeyFrame / MapPoint pointers whose getters lock mutexes; their bodies stay in the
// reference's src/ORBmatcher.cc and keep calling DescriptorDistance (INTEGRATION.md shows the patch).
#ifndef ORBGPU_HOST_ORBMATCHER_H
#define ORBGPU_HOST_ORBMATCHER_H

#include <stdint.h>
#include <vector>
#include "cv_compat.h"

struct orbgpu_matcher;

namespace ORB_SLAM2 {

class ORBmatcher {
public:
    ORBmatcher(float nnratio = 0.6, bool checkOri = true);
    ~ORBmatcher();

    // Computes the Hamming distance between two ORB descriptors
    static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);

    // ---- batched forms of the scan every Search* routine ends in (src/ORBmatcher.cc:201-226) ----
    // every row of A (nA x 32, CV_8U) against every row of B in index order; strict '<' updates, so ties
    // keep the lowest index; best/second start at 256, index at -1.  Returns false on a CUDA failure.
    bool BestTwo(const cv::Mat& A, const cv::Mat& B, std::vector<int>& bestIdx, std::vector<int>& bestDist,
                 std::vector<int>& secondDist, int cudaDevice = 0);
    // accept rule of :228-232 on top of BestTwo: best <= TH_LOW && float(best) < mfNNratio * float(second);
    // vnMatches12[i] = index in B or -1.  Returns the number of accepted matches (-1 on failure).
    int SearchBruteForce(const cv::Mat& A, const cv::Mat& B, std::vector<int>& vnMatches12, int cudaDevice = 0);
    // all-pairs keyframe matching (BASELINE config 5): K keyframes x n descriptors resident on the GPU;
    // counts[p] = accepted matches of keyframe a[p] against keyframe b[p]
    bool SetKeyFrames(const std::vector<cv::Mat>& descriptors, int cudaDevice = 0);
    bool CountPairs(const std::vector<int>& a, const std::vector<int>& b, std::vector<int>& counts);

    // rotation-histogram maxima, host, same results as :1601-1642 (indices must arrive as -1)
    void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);

public:
    static const int TH_LOW;
    static const int TH_HIGH;
    static const int HISTO_LENGTH;

protected:
    float mfNNratio;
    bool mbCheckOrientation;
    orbgpu_matcher* mpMatcher;

private:
    ORBmatcher(const ORBmatcher&);
    ORBmatcher& operator=(const ORBmatcher&);
};

}  // namespace ORB_SLAM2

#endif
