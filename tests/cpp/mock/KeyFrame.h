#pragma once
#include <map>
#include <vector>
#include <opencv2/core/core.hpp>
#include "MapPoint.h"
namespace DBoW2 {
typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;  // Thirdparty/DBoW2/DBoW2/FeatureVector.h
}
namespace ORB_SLAM2 {
class KeyFrame {
 public:
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  std::vector<cv::KeyPoint> mvKeysUn;
  cv::Mat mDescriptors;
  DBoW2::FeatureVector mFeatVec;
  std::vector<MapPoint*> mvpMapPoints;
};
}  // namespace ORB_SLAM2
