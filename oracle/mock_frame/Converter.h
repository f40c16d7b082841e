// oracle/mock_frame/Converter.h -- TEST INFRASTRUCTURE: the one Converter function src/Frame.cc calls
// (Frame::ComputeBoW; the real header pulls Eigen and g2o).
#pragma once
#include <vector>

#include <opencv2/core/core.hpp>

namespace ORB_SLAM2 {

class Converter {
 public:
  static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors) {
    std::vector<cv::Mat> v;
    v.reserve(Descriptors.rows);
    for (int j = 0; j < Descriptors.rows; j++) v.push_back(Descriptors.row(j));
    return v;
  }
};

}  // namespace ORB_SLAM2
