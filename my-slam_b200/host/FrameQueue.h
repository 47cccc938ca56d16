// FrameQueue.h -- the pinned-memory frame queue placed in Frame construction (north_star).
//
// Tracking::GrabImageMonocular (src/Tracking.cc:240-267) receives frame t+1 from the camera loop while
// frame t is still being tracked.  Push() copies the image into one slot of the library's pinned ring
// and enqueues upload + kernels + download on that slot's CUDA stream without blocking; Pop() -- called
// where Frame::ExtractORB (src/Frame.cc:247-253) used to call the extractor -- blocks only until THAT
// frame's results have landed.  Frames leave in the order they entered.  Header-only; one queue per
// extractor; not thread-safe (same contract as the extractor itself, include/ORBextractor.h:45).
#ifndef ORBGPU_HOST_FRAMEQUEUE_H
#define ORBGPU_HOST_FRAMEQUEUE_H

#include <deque>
#include "ORBextractor.h"

namespace ORB_SLAM2 {

class FrameQueue {
public:
    explicit FrameQueue(ORBextractor* extractor) : mpExtractor(extractor) {}
    int Capacity() const { return mpExtractor->QueueDepth(); }
    size_t Pending() const { return mTickets.size(); }
    bool Full() const { return (int)mTickets.size() >= Capacity(); }

    // false when the ring is full (Pop first) or the image is empty
    bool Push(const cv::Mat& imGray, double timestamp)
    {
        if (Full()) return false;
        const uint64_t t = mpExtractor->Submit(imGray);
        if (!t) return false;
        mTickets.push_back(Entry(t, timestamp));
        return true;
    }

    // oldest pending frame; keypoints / descriptors as operator() would have produced them
    bool Pop(double& timestamp, std::vector<cv::KeyPoint>& keypoints, cv::Mat& descriptors)
    {
        if (mTickets.empty()) return false;
        const Entry e = mTickets.front();
        mTickets.pop_front();
        timestamp = e.second;
        return mpExtractor->Collect(e.first, keypoints, descriptors);
    }

private:
    typedef std::pair<uint64_t, double> Entry;
    ORBextractor* mpExtractor;
    std::deque<Entry> mTickets;
};

}  // namespace ORB_SLAM2
#endif
