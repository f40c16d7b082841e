// tests/cpp/mock -- minimal stand-ins for the reference's MapPoint/KeyFrame/Frame (the real ones
// need Eigen, g2o, DBoW2 and full OpenCV, absent here).  Only the members that
// ORBmatcher::SearchByBoW touches (src/ORBmatcher.cc:360-489, 723-856) exist.
#pragma once
namespace ORB_SLAM2 {
class Map;
class MapPoint {
 public:
  explicit MapPoint(bool bad = false) : mbBad(bad) {}
  bool isBad() const { return mbBad; }
 private:
  bool mbBad;
};
}  // namespace ORB_SLAM2
