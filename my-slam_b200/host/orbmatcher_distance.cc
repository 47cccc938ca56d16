// orbmatcher_distance.cc -- for builds INSIDE the reference tree only: compiled with the
// reference's own include/ORBmatcher.h, it supplies ORBmatcher::DescriptorDistance in place of the body at
// src/ORBmatcher.cc:1647-1663 (delete that body there).  Not part of liborbslam_frontend.so, whose
// host/ORBmatcher.cc already defines the same symbol against our own class declaration.
#include "ORBmatcher.h"
#include "orbgpu.h"

int ORB_SLAM2::ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b)
{
    return orbgpu_hamming256(a.ptr(), b.ptr());
}
