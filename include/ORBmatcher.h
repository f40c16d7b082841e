/* ORBmatcher.h -- source-compatible ORB_SLAM2::ORBmatcher whose descriptor searches run on the
 * B200 through the C ABI (orbfront.h).
 *
 * Drop-in for the reference's include/ORBmatcher.h:37-111.  The class declaration is the
 * reference's (every public method, constant and protected member), so Tracking.cc,
 * LocalMapping.cc, LoopClosing.cc, Frame.cc, MapPoint.cc and PnPsolver.cc compile against it
 * unchanged.  Defined HERE, inline, over the GPU library:
 *     ORBmatcher(float, bool)                         src/ORBmatcher.cc:42-44
 *     DescriptorDistance(const cv::Mat&, const cv::Mat&)  src/ORBmatcher.cc:1848-1864
 *     SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&)  src/ORBmatcher.cc:360-489
 *     SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&) src/ORBmatcher.cc:723-856
 * plus the extension StereoSearch() for Frame::ComputeStereoMatches (src/Frame.cc:527-610).
 * The windowed projection searches (SearchByProjection x5, SearchBySim3, Fuse x2,
 * SearchForTriangulation, SearchForInitialization) stay declared and keep their definitions in
 * the reference's ORBmatcher.cc (next); INTEGRATION.md shows the four
 * definitions to delete from that file.
 *
 * Like the reference header, this one includes MapPoint.h / KeyFrame.h / Frame.h from the
 * includer's path (the reference's own in its build; tests/cpp/mock in this repo's tests).
 */
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <cstdlib>
#include <cstring>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>

#include "MapPoint.h"
#include "KeyFrame.h"
#include "Frame.h"

#include "orbfront.h"

namespace ORB_SLAM2 {

using std::pair;
using std::vector;

class ORBmatcher {
 public:
  ORBmatcher(float nnratio = 0.6, bool checkOri = true) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

  // Computes the Hamming distance between two ORB descriptors
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b) { return orbm_hamming(a.data, b.data); }

  int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);
  int SearchByProjection(Frame& F, Map* pMap, double mCamRcw[3][3], double mCamtcw[3], vector<MapPoint*>& vMatchedMPs,
                         vector<cv::KeyPoint>& vMatchedKPs, vector<bool>& vbMatched, const float th);
  int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);
  int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                         const int ORBdist);
  int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                         std::vector<MapPoint*>& vpMatched, int th);

  // Brute force constrained to ORB that belong to the same vocabulary node -- on the GPU.
  int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
  int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);

  int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                              int windowSize = 10);
  int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12, std::vector<pair<size_t, size_t> >& vMatchedPairs,
                             const bool bOnlyStereo);
  int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12, const cv::Mat& R12,
                   const cv::Mat& t12, const float th);
  int Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th = 3.0);
  int Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th, vector<MapPoint*>& vpReplacePoint);

  /* Extension: descriptor part of Frame::ComputeStereoMatches (src/Frame.cc:537-610).  Fills
   * bestIdxR / bestDist per left keypoint (bestDist = -1 where the reference `continue`s before
   * the scan); the SAD refinement (:613-684) stays with the caller. */
  static void StereoSearch(const std::vector<cv::KeyPoint>& keysLeft, const cv::Mat& descLeft,
                           const std::vector<cv::KeyPoint>& keysRight, const cv::Mat& descRight,
                           const std::vector<float>& scaleFactors, int nRows, float minD, float maxD,
                           std::vector<int>& bestIdxR, std::vector<int>& bestDist);

 public:
  static const int TH_LOW = ORBM_TH_LOW;
  static const int TH_HIGH = ORBM_TH_HIGH;
  static const int TH_RELOC = 60;
  static const int HISTO_LENGTH = ORBM_HISTO_LENGTH;

 protected:
  bool CheckDistEpipolarLine(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& F12, const KeyFrame* pKF);
  double ComputeDistance(double R[3][3], double t[3], double pt[3], double po[3]);
  bool isInFrustum(MapPoint* pMP, Frame mFr, double mRcw[3][3], double mtcw[3], float viewingCosLimit);
  float RadiusByViewingCos(const float& viewCos);
  void ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3);

  float mfNNratio;
  bool mbCheckOrientation;

  // ---- helpers of the inline definitions (not in the reference)
  static int Device() {
    const char* e = getenv("ORBFRONT_DEVICE");
    return e ? atoi(e) : 0;
  }
  static void Check(int rc, const char* what) {
    if (rc != ORB_OK) throw std::runtime_error(std::string(what) + ": " + orb_last_error());
  }
  struct FlatFeatVec {  // DBoW2::FeatureVector (ordered map) -> CSR, ascending node ids
    std::vector<uint32_t> node, idx;
    std::vector<int32_t> off;
    orbm_featvec c;
    template <typename FV>
    explicit FlatFeatVec(const FV& fv) {
      off.push_back(0);
      for (typename FV::const_iterator it = fv.begin(); it != fv.end(); ++it) {
        node.push_back((uint32_t)it->first);
        for (size_t k = 0; k < it->second.size(); ++k) idx.push_back((uint32_t)it->second[k]);
        off.push_back((int32_t)idx.size());
      }
      c.node = node.data(); c.off = off.data(); c.idx = idx.data(); c.n_nodes = (int32_t)node.size();
    }
  };
  static std::vector<uint8_t> GoodFlags(const std::vector<MapPoint*>& v) {
    std::vector<uint8_t> f(v.size(), 0);
    for (size_t i = 0; i < v.size(); ++i) f[i] = v[i] && !v[i]->isBad();
    return f;
  }
  static std::vector<float> Angles(const std::vector<cv::KeyPoint>& k) {
    std::vector<float> a(k.size());
    for (size_t i = 0; i < k.size(); ++i) a[i] = k[i].angle;
    return a;
  }
  static const uint8_t* Rows(const cv::Mat& m, std::vector<uint8_t>& tmp) {  // N x 32 contiguous view
    if (m.rows == 0 || (size_t)m.step == 32) return m.data;
    tmp.resize((size_t)m.rows * 32);
    for (int r = 0; r < m.rows; ++r) memcpy(&tmp[(size_t)r * 32], m.data + (size_t)r * (size_t)m.step, 32);
    return tmp.data();
  }
};

inline int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
  const std::vector<MapPoint*> vpMapPointsKF = pKF->GetMapPointMatches();
  vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
  const int n1 = pKF->mDescriptors.rows, n2 = F.mDescriptors.rows;
  FlatFeatVec fv1(pKF->mFeatVec), fv2(F.mFeatVec);
  std::vector<uint8_t> valid1 = GoodFlags(vpMapPointsKF), t1, t2;
  valid1.resize((size_t)n1, 0);
  std::vector<float> a1 = Angles(pKF->mvKeysUn), a2 = Angles(F.mvKeys);
  std::vector<int32_t> out((size_t)n2, -1);
  int nmatches = 0;
  Check(orbm_search_bow(Device(), Rows(pKF->mDescriptors, t1), n1, Rows(F.mDescriptors, t2), n2, &fv1.c, &fv2.c, valid1.data(),
                        NULL, a1.data(), a2.data(), mfNNratio, mbCheckOrientation ? 1 : 0, 0, out.data(), &nmatches),
        "ORBmatcher::SearchByBoW");
  for (int f = 0; f < n2 && f < F.N; ++f)
    if (out[f] >= 0) vpMapPointMatches[f] = vpMapPointsKF[out[f]];
  return nmatches;
}

inline int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
  const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
  const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
  vpMatches12 = std::vector<MapPoint*>(vpMapPoints1.size(), static_cast<MapPoint*>(NULL));
  const int n1 = pKF1->mDescriptors.rows, n2 = pKF2->mDescriptors.rows;
  FlatFeatVec fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
  std::vector<uint8_t> valid1 = GoodFlags(vpMapPoints1), valid2 = GoodFlags(vpMapPoints2), t1, t2;
  valid1.resize((size_t)n1, 0);
  valid2.resize((size_t)n2, 0);
  std::vector<float> a1 = Angles(pKF1->mvKeysUn), a2 = Angles(pKF2->mvKeysUn);
  std::vector<int32_t> out((size_t)n1, -1);
  int nmatches = 0;
  Check(orbm_search_bow(Device(), Rows(pKF1->mDescriptors, t1), n1, Rows(pKF2->mDescriptors, t2), n2, &fv1.c, &fv2.c,
                        valid1.data(), valid2.data(), a1.data(), a2.data(), mfNNratio, mbCheckOrientation ? 1 : 0, 1, out.data(),
                        &nmatches),
        "ORBmatcher::SearchByBoW");
  for (size_t i = 0; i < vpMatches12.size() && i < (size_t)n1; ++i)
    if (out[i] >= 0) vpMatches12[i] = vpMapPoints2[out[i]];
  return nmatches;
}

inline void ORBmatcher::StereoSearch(const std::vector<cv::KeyPoint>& keysLeft, const cv::Mat& descLeft,
                                     const std::vector<cv::KeyPoint>& keysRight, const cv::Mat& descRight,
                                     const std::vector<float>& scaleFactors, int nRows, float minD, float maxD,
                                     std::vector<int>& bestIdxR, std::vector<int>& bestDist) {
  static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be 7 x 4 bytes");
  std::vector<uint8_t> t1, t2;
  bestIdxR.assign(keysLeft.size(), 0);
  bestDist.assign(keysLeft.size(), -1);
  Check(orbm_stereo_search(Device(), reinterpret_cast<const orbx_keypoint*>(keysLeft.data()), Rows(descLeft, t1),
                           (int)keysLeft.size(), reinterpret_cast<const orbx_keypoint*>(keysRight.data()), Rows(descRight, t2),
                           (int)keysRight.size(), scaleFactors.data(), (int)scaleFactors.size(), nRows, minD, maxD,
                           bestIdxR.data(), bestDist.data()),
        "ORBmatcher::StereoSearch");
}

}  // namespace ORB_SLAM2

#endif  // ORBMATCHER_H
