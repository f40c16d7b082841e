// oracle/mock_frame/KeyFrame.h -- TEST INFRASTRUCTURE.
// The KeyFrame that the reference's own src/Frame.cc and src/MapPoint.cc see when they are compiled
// unmodified into oracle/_ref/libframeref.so: only the members those two files touch.  (The real
// KeyFrame pulls KeyFrameDatabase, g2o and Eigen, which this image does not have.)
#pragma once
#include <cstddef>
#include <list>
#include <map>
#include <mutex>
#include <set>
#include <vector>

#include <opencv2/core/core.hpp>

namespace ORB_SLAM2 {

using namespace std;  // the reference's own KeyFrame.h leaks std into ORB_SLAM2, and its MapPoint.h / Frame.h rely on it

class MapPoint;

class KeyFrame {
 public:
  long unsigned int mnId = 0, mnFrameId = 0;
  std::vector<cv::KeyPoint> mvKeysUn;
  std::vector<float> mvuRight;
  cv::Mat mDescriptors;
  int mnScaleLevels = 0;
  float mfLogScaleFactor = 0;
  std::vector<float> mvScaleFactors;
  cv::Mat Ow;  // 3x1 f32
  bool mbBad = false;
  std::vector<MapPoint*> mvpMapPoints;

  bool isBad() { return mbBad; }
  cv::Mat GetCameraCenter() { return Ow.clone(); }
  void EraseMapPointMatch(const size_t& idx) { mvpMapPoints[idx] = nullptr; }
  void ReplaceMapPointMatch(const size_t& idx, MapPoint* pMP) { mvpMapPoints[idx] = pMP; }
};

}  // namespace ORB_SLAM2
