// oracle/cvshim2/Converter.h -- TEST INFRASTRUCTURE.  Stands in for the reference's include/Converter.h when
// src/Frame.cc is compiled for the oracle: the real header needs Eigen and g2o, which this image does not have, and
// Frame.cc only uses toDescriptorVector (in Frame::ComputeBoW, which the oracle never runs).
#ifndef ORBO_CVSHIM2_CONVERTER_H
#define ORBO_CVSHIM2_CONVERTER_H
#include <opencv2/core/core.hpp>
#include <vector>
namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};
}
#endif
