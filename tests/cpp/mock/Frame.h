#pragma once
#include "KeyFrame.h"
namespace ORB_SLAM2 {
class Frame {
 public:
  int N = 0;
  std::vector<cv::KeyPoint> mvKeys;
  cv::Mat mDescriptors;
  DBoW2::FeatureVector mFeatVec;
};
}  // namespace ORB_SLAM2
