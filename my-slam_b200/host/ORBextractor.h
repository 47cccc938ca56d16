// ORBextractor.h -- drop-in for the reference's include/ORBextractor.h:45-111.
//
// Same namespace, class name, constructor, operator(), getters and public mvImagePyramid, so
// src/Frame.cc:69-75,247-253 and src/Tracking.cc:121-127 compile against it unchanged.  All pixel
// work runs in liborbgpu.so (CUDA, sm_100a) through the C ABI of include/orbgpu.h; there is no CPU
// implementation behind this class -- without a CUDA device the constructor throws.
//
// Additions (not in the reference):
//   Submit / Collect      the pinned-memory frame queue: Tracking::GrabImage* submits frame t+1 while
//                         frame t is being tracked, Frame::ExtractORB collects
//   SetPyramidDownload    mvImagePyramid is filled only when asked (stereo: src/Frame.cc:473,563-580
//                         read it; mono / RGB-D never do), saving a 0.95 MB D2H copy per frame
#ifndef ORBGPU_HOST_ORBEXTRACTOR_H
#define ORBGPU_HOST_ORBEXTRACTOR_H

#include <stdint.h>
#include <vector>
#include "cv_compat.h"

struct orbgpu_extractor;

namespace ORB_SLAM2 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int cudaDevice);
    ~ORBextractor();

    // Compute the ORB features and descriptors on an image (mask is ignored, as in the reference).
    void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints,
                    cv::OutputArray descriptors);

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    std::vector<cv::Mat> mvImagePyramid;

    // ---- additions ----
    // mvImagePyramid is filled (device -> host, 19-px REFLECT_101 frame rebuilt) after every frame when this is on.  Default:
    // off while the process holds ONE extractor (monocular: nothing reads the member), on for every extractor from the
    // moment a second one exists -- the stereo Frame constructor's left / right pair, whose ComputeStereoMatches reads
    // mvImagePyramid of both (src/Frame.cc:473, :563-580).  An explicit call here overrides the default for good.
    void SetPyramidDownload(bool on) { mbDownloadPyramid = on; mbDownloadExplicit = true; }
    // Accept the camera image as it arrives (3 or 4 interleaved 8-bit channels, R,G,B or B,G,R order) and do the
    // cv::cvtColor(..., CV_*2GRAY) of Tracking::GrabImage* (src/Tracking.cc:244-257) on the device; channels = 1 restores
    // the reference contract (CV_8UC1).  GrayImage() returns what mImGray would hold after the host conversion.
    bool SetColorInput(int channels, bool rgbOrder);
    cv::Mat GrayImage();
    // asynchronous pair: returns a ticket (0 = failed / queue full); Collect blocks on that frame only
    uint64_t Submit(cv::InputArray image);
    bool Collect(uint64_t ticket, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors);
    int QueueDepth() const;
    // same-shape batch (frames[i] all w x h): one upload / launch set / download per chunk
    bool ExtractBatch(const std::vector<cv::Mat>& frames, std::vector<std::vector<cv::KeyPoint> >& keypoints,
                      std::vector<cv::Mat>& descriptors);
    // Frame::ComputeStereoMatches (src/Frame.cc:466-640) between the frames the two extractors processed last: fills
    // mvuRight / mvDepth (size = left keypoint count, -1 where unmatched) without downloading either pyramid.
    static bool ComputeStereoMatches(ORBextractor& left, ORBextractor& right, float bf, float b,
                                     std::vector<float>& mvuRight, std::vector<float>& mvDepth);
    const char* LastError() const { return mLastError; }
    orbgpu_extractor* Handle() { return mpHandle; }

protected:
    void FillPyramid();
    void Finish(int n, std::vector<cv::KeyPoint>& keypoints, cv::OutputArray descriptors);

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;

    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

    orbgpu_extractor* mpHandle;
    bool mbDownloadPyramid;
    bool mbDownloadExplicit;
    const char* mLastError;
    std::vector<unsigned char> mKpScratch, mDescScratch;   // staging for the C ABI (caller-owned buffers)
    int mLastW, mLastH;
    int mChannels;

private:
    ORBextractor(const ORBextractor&);
    ORBextractor& operator=(const ORBextractor&);
};

}  // namespace ORB_SLAM2

#endif
