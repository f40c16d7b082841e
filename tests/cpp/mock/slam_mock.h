// tests/cpp/mock/slam_mock.h -- TEST INFRASTRUCTURE.
//
// Stand-ins for the reference's MapPoint / Map / KeyFrame / Frame (the real ones pull Eigen, g2o,
// DBoW2's vocabulary and full OpenCV, none of which exist in this image).  They carry exactly the
// members src/ORBmatcher.cc touches, with the small accessor bodies restated from the reference
// (file:line at each), so that
//   * the reference's OWN src/ORBmatcher.cc compiles unmodified against them (oracle/_ref/libmatchref.so)
//   * the product's include/ORBmatcher.h compiles against the very same types (tests/cpp/_build/libdropin.so)
// and the two can be run on identical scenes.
#pragma once
#include <cmath>
#include <cstddef>
#include <map>
#include <set>
#include <vector>

#include <opencv2/core/core.hpp>

#ifdef SLAM_MOCK_REAL_DBOW2  // reference build: the vendored header (and FeatureVector.cpp) from /root/reference
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"
#else
namespace DBoW2 {
typedef std::map<unsigned int, std::vector<unsigned int> > FeatureVector;  // Thirdparty/DBoW2/DBoW2/FeatureVector.h:21-22
}
#endif

#define FRAME_GRID_ROWS 48  // include/Frame.h:37-38
#define FRAME_GRID_COLS 64

namespace ORB_SLAM2 {

using namespace std;  // the reference's own headers leak std into ORB_SLAM2, and its ORBmatcher.h relies on it

class KeyFrame;
class Frame;
class Map;

class MapPoint {
 public:
  // scene state
  long unsigned int mnId = 0;
  cv::Mat mWorldPos;      // 3x1 f32
  cv::Mat mNormalVector;  // 3x1 f32
  cv::Mat mDescriptor;    // 1x32 u8
  float mfMinDistance = 0, mfMaxDistance = 0;
  int nObs = 1;
  bool mbBad = false;
  std::map<KeyFrame*, size_t> mObservations;
  MapPoint* mpReplaced = nullptr;  // set by Replace()

  // tracking scratch, include/MapPoint.h:88-107
  float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
  bool mbTrackInView = false;
  int mnTrackScaleLevel = 0;
  float mTrackViewCos = 0;
  float mRelocProjX = 0, mRelocProjY = 0, mRelocViewCos = 0;
  int mnRelocScaleLevel = 0;
  long unsigned int mnFuseCandidateForKF = 0;

  explicit MapPoint(bool bad = false) : mbBad(bad) {}

  bool isBad() const { return mbBad; }
  int Observations() const { return nObs; }
  cv::Mat GetWorldPos() const { return mWorldPos.clone(); }    // MapPoint.cc:76-80
  cv::Mat GetNormal() const { return mNormalVector.clone(); }  // :82-86
  cv::Mat GetDescriptor() const { return mDescriptor.clone(); }  // :371-375
  float GetMinDistanceInvariance() const { return 0.8f * mfMinDistance; }  // :436-440
  float GetMaxDistanceInvariance() const { return 1.2f * mfMaxDistance; }  // :442-446
  inline int PredictScale(const float& currentDist, KeyFrame* pKF);        // :448-463
  inline int PredictScale(const float& currentDist, Frame* pF);            // :465-480
  bool IsInKeyFrame(KeyFrame* pKF) const { return mObservations.count(pKF) != 0; }  // :383-387
  int GetIndexInKeyFrame(KeyFrame* pKF) const {                                      // :377-381 (returns -1 if absent)
    std::map<KeyFrame*, size_t>::const_iterator it = mObservations.find(pKF);
    return it == mObservations.end() ? -1 : (int)it->second;
  }
  void AddObservation(KeyFrame* pKF, size_t idx) {  // :94-106 (count only)
    if (mObservations.count(pKF)) return;
    mObservations[pKF] = idx;
    ++nObs;
  }
  void Replace(MapPoint* pMP) {  // :196-240: this point is dropped in favour of pMP
    if (pMP == this) return;
    mbBad = true;
    mpReplaced = pMP;
  }
};

class Map {
 public:
  std::vector<MapPoint*> mvpAll;
  std::vector<MapPoint*> GetAllMapPoints() { return mvpAll; }  // Map.cc:91-95
};

// Members shared by Frame and KeyFrame in the reference (calibration, scale pyramid, keypoints, grid).
struct FrameCommon {
  int N = 0;
  float fx = 0, fy = 0, cx = 0, cy = 0, invfx = 0, invfy = 0, mbf = 0, mb = 0;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
  std::vector<float> mvuRight, mvDepth;
  cv::Mat mDescriptors;
  DBoW2::FeatureVector mFeatVec;
  int mnScaleLevels = 0;
  float mfScaleFactor = 0, mfLogScaleFactor = 0;
  std::vector<float> mvScaleFactors, mvInvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
};

class Frame : public FrameCommon {
 public:
  long unsigned int mnId = 0;
  cv::Mat mTcw;  // 4x4 f32
  std::vector<MapPoint*> mvpMapPoints, mvpMapPointsInFrame;
  std::vector<float> fvMapPointSearchRadious;
  std::vector<bool> mvbOutlier;
  static float mnMinX, mnMaxX, mnMinY, mnMaxY;  // Frame.cc:32-33
  static float mfGridElementWidthInv, mfGridElementHeightInv;
  std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];

  // Frame.cc:397-410
  bool PosInGrid(const cv::KeyPoint& kp, int& posX, int& posY) {
    posX = round((kp.pt.x - mnMinX) * mfGridElementWidthInv);
    posY = round((kp.pt.y - mnMinY) * mfGridElementHeightInv);
    if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) return false;
    return true;
  }
  // Frame.cc:245-260
  void AssignFeaturesToGrid() {
    for (int i = 0; i < N; i++) {
      const cv::KeyPoint& kp = mvKeysUn[i];
      int nGridPosX, nGridPosY;
      if (PosInGrid(kp, nGridPosX, nGridPosY)) mGrid[nGridPosX][nGridPosY].push_back(i);
    }
  }
  // Frame.cc:342-395
  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r, const int minLevel = -1,
                                        const int maxLevel = -1) const {
    std::vector<size_t> vIndices;
    vIndices.reserve(N);
    const int nMinCellX = std::max(0, (int)floor((x - mnMinX - r) * mfGridElementWidthInv));
    if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
    const int nMaxCellX = std::min((int)FRAME_GRID_COLS - 1, (int)ceil((x - mnMinX + r) * mfGridElementWidthInv));
    if (nMaxCellX < 0) return vIndices;
    const int nMinCellY = std::max(0, (int)floor((y - mnMinY - r) * mfGridElementHeightInv));
    if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
    const int nMaxCellY = std::min((int)FRAME_GRID_ROWS - 1, (int)ceil((y - mnMinY + r) * mfGridElementHeightInv));
    if (nMaxCellY < 0) return vIndices;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
      for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
        const std::vector<size_t>& vCell = mGrid[ix][iy];
        for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
          const cv::KeyPoint& kpUn = mvKeysUn[vCell[j]];
          if (bCheckLevels) {
            if (kpUn.octave < minLevel) continue;
            if (maxLevel >= 0)
              if (kpUn.octave > maxLevel) continue;
          }
          const float distx = kpUn.pt.x - x;
          const float disty = kpUn.pt.y - y;
          if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(vCell[j]);
        }
      }
    }
    return vIndices;
  }
  // Frame.cc:481-524 (the live branch: the cached bounds)
  std::vector<int> GetImageBounds() {
    std::vector<int> bounds;
    bounds.push_back(mnMinX);
    bounds.push_back(mnMaxX);
    bounds.push_back(mnMinY);
    bounds.push_back(mnMaxY);
    return bounds;
  }
};

class KeyFrame : public FrameCommon {
 public:
  long unsigned int mnId = 0;
  int mnGridCols = FRAME_GRID_COLS, mnGridRows = FRAME_GRID_ROWS;  // KeyFrame.cc:35
  float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
  int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;              // include/KeyFrame.h: const int
  std::vector<std::vector<std::vector<size_t> > > mGrid;
  std::vector<MapPoint*> mvpMapPoints;
  cv::Mat Tcw;  // 4x4 f32
  bool mbBad = false;
  bool isBad() { return mbBad; }  // include/KeyFrame.h:102 (under mMutexConnections in the reference)

  KeyFrame() {}
  // KeyFrame.cc:30-65: takes the frame's keypoints, descriptors, grid and bounds
  explicit KeyFrame(const Frame& F) : FrameCommon(F), mnId(F.mnId) {
    mfGridElementWidthInv = Frame::mfGridElementWidthInv;
    mfGridElementHeightInv = Frame::mfGridElementHeightInv;
    mnMinX = Frame::mnMinX; mnMinY = Frame::mnMinY; mnMaxX = Frame::mnMaxX; mnMaxY = Frame::mnMaxY;
    mGrid.resize(mnGridCols);
    for (int i = 0; i < mnGridCols; i++) {
      mGrid[i].resize(mnGridRows);
      for (int j = 0; j < mnGridRows; j++) mGrid[i][j] = F.mGrid[i][j];
    }
    mvpMapPoints = F.mvpMapPoints;
    if (!F.mTcw.empty()) Tcw = F.mTcw.clone();
  }

  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }  // KeyFrame.cc:277-281
  MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }  // :283-287
  std::set<MapPoint*> GetMapPoints() {                                    // :233-246
    std::set<MapPoint*> s;
    for (size_t i = 0; i < mvpMapPoints.size(); i++) {
      if (!mvpMapPoints[i]) continue;
      MapPoint* pMP = mvpMapPoints[i];
      if (!pMP->isBad()) s.insert(pMP);
    }
    return s;
  }
  void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }  // :206-210
  cv::Mat GetRotation() { return Tcw.rowRange(0, 3).colRange(0, 3).clone(); }      // :117-121
  cv::Mat GetTranslation() { return Tcw.rowRange(0, 3).col(3).clone(); }           // :123-127
  cv::Mat GetCameraCenter() {                                                      // :98-102 (Ow = -Rwc... = -Rcw^T tcw)
    cv::Mat Rcw = Tcw.rowRange(0, 3).colRange(0, 3);
    cv::Mat tcw = Tcw.rowRange(0, 3).col(3);
    return -Rcw.t() * tcw;
  }
  bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }  // :654-657
  // KeyFrame.cc:613-652
  std::vector<size_t> GetFeaturesInArea(const float& x, const float& y, const float& r) const {
    std::vector<size_t> vIndices;
    vIndices.reserve(N);
    const int nMinCellX = std::max(0, (int)floor((x - mnMinX - r) * mfGridElementWidthInv));
    if (nMinCellX >= mnGridCols) return vIndices;
    const int nMaxCellX = std::min((int)mnGridCols - 1, (int)ceil((x - mnMinX + r) * mfGridElementWidthInv));
    if (nMaxCellX < 0) return vIndices;
    const int nMinCellY = std::max(0, (int)floor((y - mnMinY - r) * mfGridElementHeightInv));
    if (nMinCellY >= mnGridRows) return vIndices;
    const int nMaxCellY = std::min((int)mnGridRows - 1, (int)ceil((y - mnMinY + r) * mfGridElementHeightInv));
    if (nMaxCellY < 0) return vIndices;
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
      for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
        const std::vector<size_t>& vCell = mGrid[ix][iy];
        for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
          const cv::KeyPoint& kpUn = mvKeysUn[vCell[j]];
          const float distx = kpUn.pt.x - x;
          const float disty = kpUn.pt.y - y;
          if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(vCell[j]);
        }
      }
    }
    return vIndices;
  }
};

inline int MapPoint::PredictScale(const float& currentDist, KeyFrame* pKF) {
  float ratio = mfMaxDistance / currentDist;
  int nScale = ceil(log(ratio) / pKF->mfLogScaleFactor);
  if (nScale < 0) nScale = 0;
  else if (nScale >= pKF->mnScaleLevels) nScale = pKF->mnScaleLevels - 1;
  return nScale;
}
inline int MapPoint::PredictScale(const float& currentDist, Frame* pF) {
  float ratio = mfMaxDistance / currentDist;
  int nScale = ceil(log(ratio) / pF->mfLogScaleFactor);
  if (nScale < 0) nScale = 0;
  else if (nScale >= pF->mnScaleLevels) nScale = pF->mnScaleLevels - 1;
  return nScale;
}

}  // namespace ORB_SLAM2
