// oracle/mock_frame/Map.h -- TEST INFRASTRUCTURE: what src/MapPoint.cc needs of Map.
#pragma once
#include <mutex>
#include <set>

namespace ORB_SLAM2 {

class MapPoint;

class Map {
 public:
  std::mutex mMutexPointCreation;
  std::set<MapPoint*> erased;
  void EraseMapPoint(MapPoint* pMP) { erased.insert(pMP); }
};

}  // namespace ORB_SLAM2
