// oracle/shim/Converter.h -- TEST INFRASTRUCTURE.
// Stands in for the reference's include/Converter.h, which pulls in Eigen and g2o (absent here).  src/Frame.cc and
// src/KeyFrame.cc use exactly one member of it, Converter::toDescriptorVector (src/Frame.cc:576, src/KeyFrame.cc:70); it
// is declared here and defined in oracle/matchref_wrap.cc (one Mat row per descriptor, src/Converter.cc:29-40).
#pragma once
#include <vector>
#include <opencv2/core/core.hpp>

namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};
}  // namespace ORB_SLAM2
