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
 *     SearchByProjection(Frame&, const vector<MapPoint*>&, float)          src/ORBmatcher.cc:46-132
 *     SearchByProjection(Frame&, Map*, double[3][3], double[3], ...)       src/ORBmatcher.cc:134-220
 *     SearchByProjection(KeyFrame*, cv::Mat Scw, ...)                      src/ORBmatcher.cc:491-604
 *     SearchByProjection(Frame&, const Frame&, float, bool)                src/ORBmatcher.cc:1529-1671
 *     SearchByProjection(Frame&, KeyFrame*, const set<MapPoint*>&, ...)    src/ORBmatcher.cc:1673-1800
 * plus the extension StereoSearch() for Frame::ComputeStereoMatches (src/Frame.cc:527-610).
 * The projection searches keep the pose / camera / MapPoint arithmetic on the host, written with
 * the includer's cv::Mat exactly as the reference evaluates it, and hand the descriptor work --
 * grid gather, Hamming top-2, acceptance and the order-dependent hiding of matched keypoints --
 * to orbm_window_search.  Same for
 *     Fuse(KeyFrame*, const vector<MapPoint*>&, float)                     src/ORBmatcher.cc:1026-1176
 *     Fuse(KeyFrame*, cv::Mat Scw, ..., vector<MapPoint*>& vpReplacePoint) src/ORBmatcher.cc:1178-1301
 *     SearchBySim3(KeyFrame*, KeyFrame*, ..., s12, R12, t12, th)           src/ORBmatcher.cc:1303-1527
 * and, through orbm_search_triangulation (vocabulary nodes + epipolar gate),
 *     SearchForTriangulation(KeyFrame*, KeyFrame*, cv::Mat F12, ...)       src/ORBmatcher.cc:858-1024
 * and, through orbm_window_candidates (window gather + all distances; the re-assignment
 * bookkeeping of :641-668 runs over the lists on the host),
 *     SearchForInitialization(Frame&, Frame&, vbPrevMatched, vnMatches12, windowSize)  src/ORBmatcher.cc:606-721
 * With these the header defines every public method of the class: the reference's
 * src/ORBmatcher.cc leaves the build.
 *
 * Like the reference header, this one includes MapPoint.h / KeyFrame.h / Frame.h from the
 * includer's path (the reference's own in its build; tests/cpp/mock in this repo's tests).
 */
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <algorithm>
#include <climits>
#include <cmath>
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
  // Extension (not in the reference): every candidate keyframe of one relocalization against the current frame in ONE
  // device call -- the loop of src/Tracking.cc:1768-1795.  vvpMapPointMatches[i] and the returned counts equal what
  // SearchByBoW(vpKFs[i], F, ...) gives; NULL or bad keyframes get an empty result and 0.
  std::vector<int> SearchByBoW(const std::vector<KeyFrame*>& vpKFs, Frame& F, std::vector<std::vector<MapPoint*> >& vvpMapPointMatches);

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

  // ---- projection searches: flat views for orbm_window_search
  struct FlatGrid {  // a Frame's or KeyFrame's keypoints + cell grid
    std::vector<float> x, y;
    std::vector<int32_t> oct, off;
    std::vector<uint32_t> idx;
    std::vector<uint8_t> tmp;
    orbm_grid_frame c;
    void Keys(const std::vector<cv::KeyPoint>& k, const std::vector<float>& uright, const cv::Mat& desc) {
      const size_t n = k.size();
      x.resize(n); y.resize(n); oct.resize(n);
      for (size_t i = 0; i < n; ++i) { x[i] = k[i].pt.x; y[i] = k[i].pt.y; oct[i] = k[i].octave; }
      c.x = x.data(); c.y = y.data(); c.octave = oct.data();
      c.u_right = uright.size() == n ? uright.data() : NULL;
      c.desc = Rows(desc, tmp);
      c.n = (int32_t)n;
    }
    // Frame: public mGrid[64][48] and the static bounds (include/Frame.h:189,211-214)
    explicit FlatGrid(const Frame& F) {
      Keys(F.mvKeysUn, F.mvuRight, F.mDescriptors);
      c.min_x = Frame::mnMinX; c.min_y = Frame::mnMinY;
      c.width_inv = Frame::mfGridElementWidthInv; c.height_inv = Frame::mfGridElementHeightInv;
      c.cols = FRAME_GRID_COLS; c.rows = FRAME_GRID_ROWS;
      off.reserve((size_t)c.cols * c.rows + 1);
      off.push_back(0);
      for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
        for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) {
          const std::vector<std::size_t>& cell = F.mGrid[ix][iy];
          for (size_t j = 0; j < cell.size(); ++j) idx.push_back((uint32_t)cell[j]);
          off.push_back((int32_t)idx.size());
        }
      c.cell_off = off.data(); c.cell_idx = idx.data();
    }
    // KeyFrame: its mGrid is protected (include/KeyFrame.h:233) but is a copy of the grid its Frame
    // built (KeyFrame.cc:55-60) with Frame::PosInGrid (Frame.cc:397-410) over the static float
    // bounds, keypoints in index order -- rebuilt here the same way.  The window arithmetic of
    // KeyFrame::GetFeaturesInArea (KeyFrame.cc:618-632) uses the KeyFrame's own (int) bounds.
    explicit FlatGrid(const KeyFrame& KF) {
      Keys(KF.mvKeysUn, KF.mvuRight, KF.mDescriptors);
      c.min_x = KF.mnMinX; c.min_y = KF.mnMinY;
      c.width_inv = KF.mfGridElementWidthInv; c.height_inv = KF.mfGridElementHeightInv;
      c.cols = KF.mnGridCols; c.rows = KF.mnGridRows;
      const size_t n_cells = (size_t)c.cols * c.rows;
      std::vector<int32_t> cell_of(KF.mvKeysUn.size(), -1);
      off.assign(n_cells + 1, 0);
      for (size_t i = 0; i < KF.mvKeysUn.size(); ++i) {
        const int px = round((KF.mvKeysUn[i].pt.x - Frame::mnMinX) * Frame::mfGridElementWidthInv);
        const int py = round((KF.mvKeysUn[i].pt.y - Frame::mnMinY) * Frame::mfGridElementHeightInv);
        if (px < 0 || px >= c.cols || py < 0 || py >= c.rows) continue;
        cell_of[i] = px * c.rows + py;
        ++off[(size_t)cell_of[i] + 1];
      }
      for (size_t k = 0; k < n_cells; ++k) off[k + 1] += off[k];
      idx.resize((size_t)off[n_cells]);
      std::vector<int32_t> fill(off.begin(), off.end() - 1);
      for (size_t i = 0; i < cell_of.size(); ++i)
        if (cell_of[i] >= 0) idx[(size_t)fill[(size_t)cell_of[i]]++] = (uint32_t)i;
      c.cell_off = off.data(); c.cell_idx = idx.data();
    }
  };
  struct WindowQueries {  // the queries of one call, in the reference's loop order
    std::vector<orbm_window_query> q;
    std::vector<uint8_t> desc;
    std::vector<MapPoint*> mp;
    std::vector<int> src;  // caller-side index of the query (keypoint of the last frame / keyframe)
    void Add(MapPoint* pMP, const cv::Mat& d, float x, float y, float r, int minLevel, int maxLevel, float uRight, float erMax,
             bool blocks, int source = -1) {
      orbm_window_query w;
      w.x = x; w.y = y; w.r = r; w.u_right = uRight; w.er_max = erMax;
      w.min_level = minLevel; w.max_level = maxLevel; w.blocks = blocks ? 1 : 0;
      q.push_back(w);
      desc.insert(desc.end(), d.data, d.data + 32);
      mp.push_back(pMP);
      src.push_back(source);
    }
    std::vector<orbm_window_result> Run(const FlatGrid& g, std::vector<uint8_t> hidden, int th, float ratio, bool claim,
                                        const char* what, const std::vector<float>* invLevelSigma2 = NULL) const {
      std::vector<orbm_window_result> res(q.size());
      int n = 0;
      hidden.resize((size_t)g.c.n, 0);
      orbm_window_params prm;
      prm.th = th; prm.ratio = ratio; prm.claim = claim ? 1 : 0;
      prm.inv_level_sigma2 = invLevelSigma2 ? invLevelSigma2->data() : NULL;
      prm.n_levels = invLevelSigma2 ? (int32_t)invLevelSigma2->size() : 0;
      prm.chi2_mono = 5.99; prm.chi2_stereo = 7.8;  // src/ORBmatcher.cc:1131,1143
      if (!q.empty())
        Check(orbm_window_search(Device(), &g.c, q.data(), desc.data(), (int)q.size(), hidden.data(), &prm, res.data(), &n, NULL),
              what);
      return res;
    }
  };
  static int RotationBin(float a1, float a2) {  // src/ORBmatcher.cc:1628-1634
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
  }
  static void ThreeMaxima(const std::vector<int>* histo, int& ind1, int& ind2, int& ind3) {  // src/ORBmatcher.cc:1802-1843
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < HISTO_LENGTH; i++) {
      const int s = (int)histo[i].size();
      if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
      else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) ind3 = -1;
  }
  // src/ORBmatcher.cc:262-330 with :222-260 folded in: frustum test of the map-wide search.
  static bool InFrustumReloc(MapPoint* pMP, const Frame& F, double R[3][3], double t[3], float viewingCosLimit);
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

inline std::vector<int> ORBmatcher::SearchByBoW(const std::vector<KeyFrame*>& vpKFs, Frame& F,
                                                std::vector<std::vector<MapPoint*> >& vvpMapPointMatches) {
  const size_t K = vpKFs.size();
  vvpMapPointMatches.assign(K, std::vector<MapPoint*>());
  std::vector<int> nmatches(K, 0);
  const int n2 = F.mDescriptors.rows;
  FlatFeatVec fv2(F.mFeatVec);
  std::vector<uint8_t> t2;
  std::vector<float> a2 = Angles(F.mvKeys);
  orbm_bow_side side2 = {Rows(F.mDescriptors, t2), n2, &fv2.c, NULL, a2.data()};
  // host-side packing per candidate; everything stays alive until the call returns
  struct Cand { std::vector<MapPoint*> mps; std::vector<uint8_t> valid, tmp; std::vector<float> ang; std::vector<int32_t> out; };
  std::vector<Cand> c(K);
  std::vector<FlatFeatVec*> fvs(K, (FlatFeatVec*)NULL);
  std::vector<orbm_bow_side> side1;
  std::vector<int32_t*> outs;
  std::vector<size_t> which;
  for (size_t i = 0; i < K; ++i) {
    KeyFrame* pKF = vpKFs[i];
    if (!pKF || pKF->isBad()) continue;  // Tracking.cc:1771-1772 discards them
    Cand& ci = c[i];
    ci.mps = pKF->GetMapPointMatches();
    const int n1 = pKF->mDescriptors.rows;
    ci.valid = GoodFlags(ci.mps);
    ci.valid.resize((size_t)n1, 0);
    ci.ang = Angles(pKF->mvKeysUn);
    ci.out.assign((size_t)n2, -1);
    fvs[i] = new FlatFeatVec(pKF->mFeatVec);
    orbm_bow_side s1 = {Rows(pKF->mDescriptors, ci.tmp), n1, &fvs[i]->c, ci.valid.data(), ci.ang.data()};
    side1.push_back(s1);
    outs.push_back(ci.out.data());
    which.push_back(i);
  }
  std::vector<int> nm(side1.size(), 0);
  const int rc = side1.empty() ? ORB_OK
                               : orbm_search_bow_batch(Device(), side1.data(), (int)side1.size(), &side2, mfNNratio,
                                                       mbCheckOrientation ? 1 : 0, 0, outs.data(), nm.data());
  for (size_t i = 0; i < K; ++i) delete fvs[i];
  Check(rc, "ORBmatcher::SearchByBoW (batch)");
  for (size_t j = 0; j < which.size(); ++j) {
    const size_t i = which[j];
    vvpMapPointMatches[i] = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));
    for (int f = 0; f < n2 && f < F.N; ++f)
      if (c[i].out[f] >= 0) vvpMapPointMatches[i][f] = c[i].mps[c[i].out[f]];
    nmatches[i] = nm[j];
  }
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

// ---- projection searches ----------------------------------------------------------------------
// Each overload walks its MapPoints in the reference's order, keeps the reference's skip tests and
// its float expressions for the projection, and queues one window query per surviving point; the
// GPU resolves all queries of the call at once (orbm_window_search), then the matches are applied
// in query order, which reproduces the reference's writes one for one.

// Local map points into the current frame (Tracking::SearchLocalPoints).  src/ORBmatcher.cc:46-132
inline int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th) {
  const bool bFactor = th != 1.0;
  WindowQueries W;
  for (size_t iMP = 0; iMP < vpMapPoints.size(); iMP++) {
    MapPoint* pMP = vpMapPoints[iMP];
    if (!pMP->mbTrackInView) continue;
    if (pMP->isBad()) continue;
    const int& nPredictedLevel = pMP->mnTrackScaleLevel;
    float r = pMP->mTrackViewCos > 0.998 ? 3.0 : 4.5;  // RadiusByViewingCos, :332-338
    if (bFactor) r *= th;
    const float window = r * F.mvScaleFactors[nPredictedLevel];
    W.Add(pMP, pMP->GetDescriptor(), pMP->mTrackProjX, pMP->mTrackProjY, window, nPredictedLevel - 1, nPredictedLevel,
          pMP->mTrackProjXR, window, pMP->Observations() > 0);
  }
  std::vector<uint8_t> hidden(F.mvpMapPoints.size(), 0);  // :81-83
  for (size_t i = 0; i < hidden.size(); ++i) hidden[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;
  const FlatGrid grid(F);
  const std::vector<orbm_window_result> res = W.Run(grid, hidden, TH_HIGH, mfNNratio, true, "ORBmatcher::SearchByProjection");
  int nmatches = 0;
  for (size_t k = 0; k < res.size(); ++k) {
    if (res[k].idx < 0) continue;
    F.mvpMapPoints[res[k].idx] = W.mp[k];  // :125-127
    F.mvpMapPointsInFrame[res[k].idx] = W.mp[k];
    F.fvMapPointSearchRadious[res[k].idx] = W.q[k].r;
    nmatches++;
  }
  return nmatches;
}

inline bool ORBmatcher::InFrustumReloc(MapPoint* pMP, const Frame& F, double R[3][3], double t[3], float viewingCosLimit) {
  cv::Mat P = pMP->GetWorldPos();
  const float ptX = P.at<float>(0), ptY = P.at<float>(1), ptZ = P.at<float>(2);
  // camera coordinates: double products rounded to float once, :273-275
  const float PcX = R[0][0] * ptX + R[0][1] * ptY + R[0][2] * ptZ + t[0];
  const float PcY = R[1][0] * ptX + R[1][1] * ptY + R[1][2] * ptZ + t[1];
  const float PcZ = R[2][0] * ptX + R[2][1] * ptY + R[2][2] * ptZ + t[2];
  if (PcZ < 0.0) return false;
  const float invz = 1.0 / PcZ;
  const float u = F.fx * PcX * invz + F.cx;
  const float v = F.fy * PcY * invz + F.cy;
  // image bounds as ints, like Frame::GetImageBounds (Frame.cc:517-524)
  const int minX = Frame::mnMinX, maxX = Frame::mnMaxX, minY = Frame::mnMinY, maxY = Frame::mnMaxY;
  if (u < minX || u > maxX) return false;
  if (v < minY || v > maxY) return false;
  const float maxDistance = pMP->GetMaxDistanceInvariance();
  const float minDistance = pMP->GetMinDistanceInvariance();
  // PO = P - Ow with Ow = -R^T t, in double (ComputeDistance, :222-260)
  double PO[3];
  for (int i = 0; i < 3; ++i) {
    const double Rtt = (-1) * R[0][i] * t[0] + (-1) * R[1][i] * t[1] + (-1) * R[2][i] * t[2];
    PO[i] = (double)(i == 0 ? ptX : i == 1 ? ptY : ptZ) - Rtt;
  }
  const float dist = sqrt(PO[0] * PO[0] + PO[1] * PO[1] + PO[2] * PO[2]);
  if (dist < (0.9 * minDistance) || dist > (maxDistance / 0.9)) return false;
  cv::Mat Pn = pMP->GetNormal();
  const float pn[] = {Pn.at<float>(0), Pn.at<float>(1), Pn.at<float>(2)};
  float viewCos = PO[0] * pn[0] + PO[1] * pn[1] + PO[2] * pn[2];
  viewCos = viewCos / dist;
  if (viewCos < viewingCosLimit) return false;
  const float ratio = dist / minDistance;
  int nPredictedLevel = (int)(std::lower_bound(F.mvScaleFactors.begin(), F.mvScaleFactors.end(), ratio) - F.mvScaleFactors.begin());
  if (nPredictedLevel >= F.mnScaleLevels) nPredictedLevel = F.mnScaleLevels - 1;
  pMP->mRelocProjX = u;
  pMP->mRelocProjY = v;
  pMP->mnRelocScaleLevel = nPredictedLevel;
  pMP->mRelocViewCos = viewCos;
  return true;
}

// Every map point against a frame under a candidate pose (PnPsolver).  src/ORBmatcher.cc:134-220
inline int ORBmatcher::SearchByProjection(Frame& F, Map* pMap, double mCamRcw[3][3], double mCamtcw[3],
                                          vector<MapPoint*>& vMatchedMPs, vector<cv::KeyPoint>& vMatchedKPs,
                                          vector<bool>& vbMatched, const float th) {
  const std::vector<MapPoint*> vpAllMapPoints = pMap->GetAllMapPoints();
  const int nFeatures = (int)F.mvKeysUn.size();
  vMatchedMPs = vector<MapPoint*>(nFeatures, static_cast<MapPoint*>(NULL));
  vMatchedKPs.reserve(nFeatures);
  vbMatched = vector<bool>(nFeatures, false);
  const bool bFactor = th != 1.0;
  WindowQueries W;
  for (size_t iMP = 0; iMP < vpAllMapPoints.size(); iMP++) {
    MapPoint* pMP = vpAllMapPoints[iMP];
    if (!InFrustumReloc(pMP, F, mCamRcw, mCamtcw, 0.5)) continue;
    const int& nPredictedLevel = pMP->mnRelocScaleLevel;
    float r = pMP->mRelocViewCos > 0.998 ? 3.0 : 4.5;
    if (bFactor) r *= th;
    W.Add(pMP, pMP->GetDescriptor(), pMP->mRelocProjX, pMP->mRelocProjY, r * F.mvScaleFactors[nPredictedLevel],
          nPredictedLevel - 1, nPredictedLevel, 0.f, -1.f, false);
  }
  std::vector<uint8_t> hidden(F.mvpMapPoints.size(), 0);  // :181-182
  for (size_t i = 0; i < hidden.size(); ++i) hidden[i] = F.mvpMapPoints[i] != NULL;
  const FlatGrid grid(F);
  // matches go to a separate vector, so no query hides a keypoint from a later one
  const std::vector<orbm_window_result> res = W.Run(grid, hidden, TH_RELOC, mfNNratio, false, "ORBmatcher::SearchByProjection");
  int nmatches = 0;
  for (size_t k = 0; k < res.size(); ++k) {
    if (res[k].idx < 0) continue;
    vMatchedMPs[res[k].idx] = W.mp[k];
    vMatchedKPs.data()[res[k].idx] = F.mvKeys[res[k].idx];  // :211 writes into the reserved storage; size() stays 0
    vbMatched[res[k].idx] = true;
    nmatches++;
  }
  return nmatches;
}

// Loop-closure candidates into a keyframe under a Sim3.  src/ORBmatcher.cc:491-604
inline int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                                          std::vector<MapPoint*>& vpMatched, int th) {
  const float &fx = pKF->fx, &fy = pKF->fy, &cx = pKF->cx, &cy = pKF->cy;
  cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  cv::Mat Rcw = sRcw / scw;
  cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
  cv::Mat Ow = -Rcw.t() * tcw;
  std::set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());
  spAlreadyFound.erase(static_cast<MapPoint*>(NULL));
  WindowQueries W;
  for (int iMP = 0, iendMP = (int)vpPoints.size(); iMP < iendMP; iMP++) {
    MapPoint* pMP = vpPoints[iMP];
    if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc = Rcw * p3Dw + tcw;
    if (p3Dc.at<float>(2) < 0.0) continue;
    const float invz = 1 / p3Dc.at<float>(2);
    const float x = p3Dc.at<float>(0) * invz;
    const float y = p3Dc.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF->IsInImage(u, v)) continue;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    cv::Mat PO = p3Dw - Ow;
    const float dist = cv::norm(PO);
    if (dist < minDistance || dist > maxDistance) continue;
    cv::Mat Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist) continue;
    const int nPredictedLevel = pMP->PredictScale(dist, pKF);
    const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
    // the level test sits inside the candidate loop (:575-578); as a window filter it is the same set
    W.Add(pMP, pMP->GetDescriptor(), u, v, radius, nPredictedLevel - 1, nPredictedLevel, 0.f, -1.f, true);
  }
  std::vector<uint8_t> hidden(vpMatched.size(), 0);  // :570-571
  for (size_t i = 0; i < hidden.size(); ++i) hidden[i] = vpMatched[i] != NULL;
  const FlatGrid grid(*pKF);
  const std::vector<orbm_window_result> res = W.Run(grid, hidden, TH_LOW, 0.f, true, "ORBmatcher::SearchByProjection");
  int nmatches = 0;
  for (size_t k = 0; k < res.size(); ++k) {
    if (res[k].idx < 0) continue;
    vpMatched[res[k].idx] = W.mp[k];
    nmatches++;
  }
  return nmatches;
}

// Last frame's points into the current frame (Tracking::TrackWithMotionModel).  src/ORBmatcher.cc:1529-1671
inline int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat twc = -Rcw.t() * tcw;
  const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat tlc = Rlw * twc + tlw;
  const bool bForward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  const bool bBackward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  WindowQueries W;
  for (int i = 0; i < LastFrame.N; i++) {
    MapPoint* pMP = LastFrame.mvpMapPoints[i];
    if (!pMP || LastFrame.mvbOutlier[i]) continue;
    cv::Mat x3Dw = pMP->GetWorldPos();
    cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float xc = x3Dc.at<float>(0);
    const float yc = x3Dc.at<float>(1);
    const float invzc = 1.0 / x3Dc.at<float>(2);
    if (invzc < 0) continue;
    const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
    const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
    if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    const int nLastOctave = LastFrame.mvKeys[i].octave;
    const float radius = th * CurrentFrame.mvScaleFactors[nLastOctave];
    int minLevel, maxLevel;  // :1583-1588
    if (bForward) { minLevel = nLastOctave; maxLevel = -1; }
    else if (bBackward) { minLevel = 0; maxLevel = nLastOctave; }
    else { minLevel = nLastOctave - 1; maxLevel = nLastOctave + 1; }
    const float ur = u - CurrentFrame.mbf * invzc;  // :1606-1611
    W.Add(pMP, pMP->GetDescriptor(), u, v, radius, minLevel, maxLevel, ur, radius, pMP->Observations() > 0, i);
  }
  std::vector<uint8_t> hidden(CurrentFrame.mvpMapPoints.size(), 0);  // :1601-1603
  for (size_t i = 0; i < hidden.size(); ++i)
    hidden[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
  const FlatGrid grid(CurrentFrame);
  const std::vector<orbm_window_result> res = W.Run(grid, hidden, TH_HIGH, 0.f, true, "ORBmatcher::SearchByProjection");
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  for (size_t k = 0; k < res.size(); ++k) {
    const int bestIdx2 = res[k].idx;
    if (bestIdx2 < 0) continue;
    CurrentFrame.mvpMapPoints[bestIdx2] = W.mp[k];
    nmatches++;
    if (mbCheckOrientation)
      rotHist[RotationBin(LastFrame.mvKeysUn[W.src[k]].angle, CurrentFrame.mvKeysUn[bestIdx2].angle)].push_back(bestIdx2);
  }
  if (mbCheckOrientation) {  // :1643-1665
    int ind1, ind2, ind3;
    ThreeMaxima(rotHist, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        CurrentFrame.mvpMapPoints[rotHist[i][j]] = static_cast<MapPoint*>(NULL);
        nmatches--;
      }
    }
  }
  return nmatches;
}

// A relocalization candidate keyframe's points into the current frame.  src/ORBmatcher.cc:1673-1800
inline int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound,
                                          const float th, const int ORBdist) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat Ow = -Rcw.t() * tcw;
  const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();
  WindowQueries W;
  for (size_t i = 0, iend = vpMPs.size(); i < iend; i++) {
    MapPoint* pMP = vpMPs[i];
    if (!pMP || pMP->isBad() || sAlreadyFound.count(pMP)) continue;
    cv::Mat x3Dw = pMP->GetWorldPos();
    cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float xc = x3Dc.at<float>(0);
    const float yc = x3Dc.at<float>(1);
    const float invzc = 1.0 / x3Dc.at<float>(2);
    const float u = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
    const float v = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX) continue;
    if (v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    cv::Mat PO = x3Dw - Ow;
    float dist3D = cv::norm(PO);
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    const int nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame);
    const float radius = th * CurrentFrame.mvScaleFactors[nPredictedLevel];
    W.Add(pMP, pMP->GetDescriptor(), u, v, radius, nPredictedLevel - 1, nPredictedLevel + 1, 0.f, -1.f, true, (int)i);
  }
  std::vector<uint8_t> hidden(CurrentFrame.mvpMapPoints.size(), 0);  // :1741-1742: any MapPoint hides the keypoint
  for (size_t i = 0; i < hidden.size(); ++i) hidden[i] = CurrentFrame.mvpMapPoints[i] != NULL;
  const FlatGrid grid(CurrentFrame);
  const std::vector<orbm_window_result> res = W.Run(grid, hidden, ORBdist, 0.f, true, "ORBmatcher::SearchByProjection");
  int nmatches = 0;
  std::vector<int> rotHist[HISTO_LENGTH];
  for (size_t k = 0; k < res.size(); ++k) {
    const int bestIdx2 = res[k].idx;
    if (bestIdx2 < 0) continue;
    CurrentFrame.mvpMapPoints[bestIdx2] = W.mp[k];
    nmatches++;
    if (mbCheckOrientation)
      rotHist[RotationBin(pKF->mvKeysUn[W.src[k]].angle, CurrentFrame.mvKeysUn[bestIdx2].angle)].push_back(bestIdx2);
  }
  if (mbCheckOrientation) {  // :1779-1797
    int ind1, ind2, ind3;
    ThreeMaxima(rotHist, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        CurrentFrame.mvpMapPoints[rotHist[i][j]] = NULL;
        nmatches--;
      }
    }
  }
  return nmatches;
}

// Local map points into a neighbouring keyframe (LocalMapping::SearchInNeighbors).
// src/ORBmatcher.cc:1026-1176.  The search reads only the keyframe's keypoints, so all points are
// searched at once; the bookkeeping (Replace / AddObservation / AddMapPoint) then runs in the
// reference's order, re-testing the two skip conditions that earlier iterations can switch on.
inline int ORBmatcher::Fuse(KeyFrame* pKF, const vector<MapPoint*>& vpMapPoints, const float th) {
  cv::Mat Rcw = pKF->GetRotation();
  cv::Mat tcw = pKF->GetTranslation();
  const float &fx = pKF->fx, &fy = pKF->fy, &cx = pKF->cx, &cy = pKF->cy, &bf = pKF->mbf;
  cv::Mat Ow = pKF->GetCameraCenter();
  const int nMPs = (int)vpMapPoints.size();
  WindowQueries W;
  for (int i = 0; i < nMPs; i++) {
    MapPoint* pMP = vpMapPoints[i];
    if (!pMP) continue;
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;  // both only ever turn on: safe to test early, re-tested below
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc = Rcw * p3Dw + tcw;
    if (p3Dc.at<float>(2) < 0.0f) continue;
    const float invz = 1 / p3Dc.at<float>(2);
    const float x = p3Dc.at<float>(0) * invz;
    const float y = p3Dc.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF->IsInImage(u, v)) continue;
    const float ur = u - bf * invz;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    cv::Mat PO = p3Dw - Ow;
    const float dist3D = cv::norm(PO);
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    cv::Mat Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist3D) continue;
    const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
    const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
    W.Add(pMP, pMP->GetDescriptor(), u, v, radius, nPredictedLevel - 1, nPredictedLevel, ur, -1.f, false, i);
  }
  const FlatGrid grid(*pKF);
  const std::vector<orbm_window_result> res =
      W.Run(grid, std::vector<uint8_t>(), TH_LOW, 0.f, false, "ORBmatcher::Fuse", &pKF->mvInvLevelSigma2);
  int nFused = 0;
  for (size_t k = 0; k < res.size(); ++k) {
    MapPoint* pMP = W.mp[k];
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;  // :1047-1048, as of this iteration
    const int bestIdx = res[k].idx;
    if (bestIdx < 0) continue;
    MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx);  // :1149-1167
    if (pMPinKF) {
      if (!pMPinKF->isBad()) {
        if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
        else pMPinKF->Replace(pMP);
      }
    } else {
      pMP->AddObservation(pKF, bestIdx);
      pKF->AddMapPoint(pMP, bestIdx);
    }
    nFused++;
  }
  return nFused;
}

// Loop-closure points into a keyframe under a corrected Sim3 (LoopClosing::SearchAndFuse).
// src/ORBmatcher.cc:1178-1301
inline int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
                            vector<MapPoint*>& vpReplacePoint) {
  const float &fx = pKF->fx, &fy = pKF->fy, &cx = pKF->cx, &cy = pKF->cy;
  cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  cv::Mat Rcw = sRcw / scw;
  cv::Mat tcw = Scw.rowRange(0, 3).col(3) / scw;
  cv::Mat Ow = -Rcw.t() * tcw;
  const std::set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();
  const int nPoints = (int)vpPoints.size();
  WindowQueries W;
  for (int iMP = 0; iMP < nPoints; iMP++) {
    MapPoint* pMP = vpPoints[iMP];
    if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
    cv::Mat p3Dw = pMP->GetWorldPos();
    cv::Mat p3Dc = Rcw * p3Dw + tcw;
    if (p3Dc.at<float>(2) < 0.0f) continue;
    const float invz = 1.0 / p3Dc.at<float>(2);
    const float x = p3Dc.at<float>(0) * invz;
    const float y = p3Dc.at<float>(1) * invz;
    const float u = fx * x + cx;
    const float v = fy * y + cy;
    if (!pKF->IsInImage(u, v)) continue;
    const float maxDistance = pMP->GetMaxDistanceInvariance();
    const float minDistance = pMP->GetMinDistanceInvariance();
    cv::Mat PO = p3Dw - Ow;
    const float dist3D = cv::norm(PO);
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    cv::Mat Pn = pMP->GetNormal();
    if (PO.dot(Pn) < 0.5 * dist3D) continue;
    const int nPredictedLevel = pMP->PredictScale(dist3D, pKF);
    const float radius = th * pKF->mvScaleFactors[nPredictedLevel];
    W.Add(pMP, pMP->GetDescriptor(), u, v, radius, nPredictedLevel - 1, nPredictedLevel, 0.f, -1.f, false, iMP);
  }
  const FlatGrid grid(*pKF);
  const std::vector<orbm_window_result> res = W.Run(grid, std::vector<uint8_t>(), TH_LOW, 0.f, false, "ORBmatcher::Fuse");
  int nFused = 0;
  for (size_t k = 0; k < res.size(); ++k) {
    const int bestIdx = res[k].idx;
    if (bestIdx < 0) continue;
    MapPoint* pMP = W.mp[k];
    MapPoint* pMPinKF = pKF->GetMapPoint(bestIdx);  // :1281-1292
    if (pMPinKF) {
      if (!pMPinKF->isBad()) vpReplacePoint[W.src[k]] = pMPinKF;
    } else {
      pMP->AddObservation(pKF, bestIdx);
      pKF->AddMapPoint(pMP, bestIdx);
    }
    nFused++;
  }
  return nFused;
}

// Mutual projection search between two keyframes under a Sim3 (LoopClosing::ComputeSim3).
// src/ORBmatcher.cc:1303-1527
inline int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
                                    const cv::Mat& R12, const cv::Mat& t12, const float th) {
  const float &fx = pKF1->fx, &fy = pKF1->fy, &cx = pKF1->cx, &cy = pKF1->cy;
  cv::Mat R1w = pKF1->GetRotation();
  cv::Mat t1w = pKF1->GetTranslation();
  cv::Mat R2w = pKF2->GetRotation();
  cv::Mat t2w = pKF2->GetTranslation();
  cv::Mat sR12 = s12 * R12;
  cv::Mat sR21 = (1.0 / s12) * R12.t();
  cv::Mat t21 = -sR21 * t12;
  const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches();
  const int N1 = (int)vpMapPoints1.size();
  const std::vector<MapPoint*> vpMapPoints2 = pKF2->GetMapPointMatches();
  const int N2 = (int)vpMapPoints2.size();
  std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);
  for (int i = 0; i < N1; i++) {
    MapPoint* pMP = vpMatches12[i];
    if (pMP) {
      vbAlreadyMatched1[i] = true;
      const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
      if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
    }
  }
  // one direction: points of `from` (already in its camera frame through Rfw, tfw) into `into`
  struct Direction {
    static std::vector<int> Search(const std::vector<MapPoint*>& vpMapPoints, const std::vector<bool>& vbAlreadyMatched,
                                   const cv::Mat& Rfw, const cv::Mat& tfw, const cv::Mat& sR, const cv::Mat& t, KeyFrame* into,
                                   float fx, float fy, float cx, float cy, float th) {
      const int N = (int)vpMapPoints.size();
      WindowQueries W;
      for (int i = 0; i < N; i++) {
        MapPoint* pMP = vpMapPoints[i];
        if (!pMP || vbAlreadyMatched[i]) continue;
        if (pMP->isBad()) continue;
        cv::Mat p3Dw = pMP->GetWorldPos();
        cv::Mat p3Dcf = Rfw * p3Dw + tfw;
        cv::Mat p3Dc = sR * p3Dcf + t;
        if (p3Dc.at<float>(2) < 0.0) continue;
        const float invz = 1.0 / p3Dc.at<float>(2);
        const float x = p3Dc.at<float>(0) * invz;
        const float y = p3Dc.at<float>(1) * invz;
        const float u = fx * x + cx;
        const float v = fy * y + cy;
        if (!into->IsInImage(u, v)) continue;
        const float maxDistance = pMP->GetMaxDistanceInvariance();
        const float minDistance = pMP->GetMinDistanceInvariance();
        const float dist3D = cv::norm(p3Dc);
        if (dist3D < minDistance || dist3D > maxDistance) continue;
        const int nPredictedLevel = pMP->PredictScale(dist3D, into);
        const float radius = th * into->mvScaleFactors[nPredictedLevel];
        W.Add(pMP, pMP->GetDescriptor(), u, v, radius, nPredictedLevel - 1, nPredictedLevel, 0.f, -1.f, false, i);
      }
      const FlatGrid grid(*into);
      const std::vector<orbm_window_result> res =
          W.Run(grid, std::vector<uint8_t>(), TH_HIGH, 0.f, false, "ORBmatcher::SearchBySim3");
      std::vector<int> vnMatch(N, -1);
      for (size_t k = 0; k < res.size(); ++k)
        if (res[k].idx >= 0) vnMatch[W.src[k]] = res[k].idx;
      return vnMatch;
    }
  };
  const std::vector<int> vnMatch1 = Direction::Search(vpMapPoints1, vbAlreadyMatched1, R1w, t1w, sR21, t21, pKF2, fx, fy, cx, cy, th);
  const std::vector<int> vnMatch2 = Direction::Search(vpMapPoints2, vbAlreadyMatched2, R2w, t2w, sR12, t12, pKF1, fx, fy, cx, cy, th);
  int nFound = 0;  // :1510-1524: keep what both directions agree on
  for (int i1 = 0; i1 < N1; i1++) {
    const int idx2 = vnMatch1[i1];
    if (idx2 >= 0 && vnMatch2[idx2] == i1) {
      vpMatches12[i1] = vpMapPoints2[idx2];
      nFound++;
    }
  }
  return nFound;
}

// Unmatched keypoints of two keyframes under the epipolar constraint (LocalMapping::CreateNewMapPoints).
// src/ORBmatcher.cc:858-1024
inline int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                              std::vector<pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo) {
  // epipole in the second image, :866-872
  cv::Mat Cw = pKF1->GetCameraCenter();
  cv::Mat R2w = pKF2->GetRotation();
  cv::Mat t2w = pKF2->GetTranslation();
  cv::Mat C2 = R2w * Cw + t2w;
  const float invz = 1.0f / C2.at<float>(2);
  const float ex = pKF2->fx * C2.at<float>(0) * invz + pKF2->cx;
  const float ey = pKF2->fy * C2.at<float>(1) * invz + pKF2->cy;
  struct Side {
    std::vector<float> x, y, angle;
    std::vector<int32_t> oct;
    std::vector<uint8_t> usable, stereo, tmp;
    FlatFeatVec fv;
    orbm_tri_keyframe c;
    Side(KeyFrame* pKF, bool bOnlyStereo) : fv(pKF->mFeatVec) {
      const size_t n = pKF->mvKeysUn.size();
      x.resize(n); y.resize(n); angle.resize(n); oct.resize(n); usable.resize(n); stereo.resize(n);
      for (size_t i = 0; i < n; ++i) {
        const cv::KeyPoint& kp = pKF->mvKeysUn[i];
        x[i] = kp.pt.x; y[i] = kp.pt.y; angle[i] = kp.angle; oct[i] = kp.octave;
        stereo[i] = i < pKF->mvuRight.size() && pKF->mvuRight[i] >= 0;
        usable[i] = !pKF->GetMapPoint(i) && (!bOnlyStereo || stereo[i]);  // :897-905, :915-925
      }
      c.desc = Rows(pKF->mDescriptors, tmp);
      c.x = x.data(); c.y = y.data(); c.octave = oct.data(); c.angle = angle.data();
      c.usable = usable.data(); c.stereo = stereo.data(); c.n = (int32_t)n; c.fv = &fv.c;
    }
  };
  const Side s1(pKF1, bOnlyStereo), s2(pKF2, bOnlyStereo);
  float F[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) F[3 * r + c] = F12.at<float>(r, c);
  std::vector<int32_t> vMatches12((size_t)s1.c.n, -1);
  int nmatches = 0;
  Check(orbm_search_triangulation(Device(), &s1.c, &s2.c, F, ex, ey, pKF2->mvScaleFactors.data(), pKF2->mvLevelSigma2.data(),
                                  (int)pKF2->mvScaleFactors.size(), mbCheckOrientation ? 1 : 0, vMatches12.data(), &nmatches),
        "ORBmatcher::SearchForTriangulation");
  vMatchedPairs.clear();
  vMatchedPairs.reserve(nmatches);
  for (size_t i = 0, iend = vMatches12.size(); i < iend; i++) {
    if (vMatches12[i] < 0) continue;
    vMatchedPairs.push_back(std::make_pair(i, (size_t)vMatches12[i]));
  }
  return nmatches;
}

// Monocular initialisation: level-0 keypoints of the first frame searched in a window of the second.
// src/ORBmatcher.cc:606-721.  A keypoint of F2 can be re-assigned to a later query with a smaller
// distance (:641-642, :661-668), so the acceptance needs every candidate, not the best two: the GPU
// gathers the windows and computes all distances (orbm_window_candidates); the re-assignment
// bookkeeping then runs over those lists in the reference's order.
inline int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                               std::vector<int>& vnMatches12, int windowSize) {
  int nmatches = 0;
  vnMatches12 = std::vector<int>(F1.mvKeysUn.size(), -1);
  std::vector<int> rotHist[HISTO_LENGTH];
  std::vector<int> vMatchedDistance(F2.mvKeysUn.size(), INT_MAX);
  std::vector<int> vnMatches21(F2.mvKeysUn.size(), -1);
  WindowQueries W;
  for (size_t i1 = 0, iend1 = F1.mvKeysUn.size(); i1 < iend1; i1++) {
    const int level1 = F1.mvKeysUn[i1].octave;
    if (level1 > 0) continue;
    W.Add(NULL, F1.mDescriptors.row((int)i1), vbPrevMatched[i1].x, vbPrevMatched[i1].y, windowSize, level1, level1, 0.f, -1.f, false,
          (int)i1);
  }
  const FlatGrid grid(F2);
  const int nq = (int)W.q.size();
  std::vector<int32_t> offset((size_t)nq), count((size_t)nq), candIdx, candDist;
  if (nq > 0) {
    int total = 0;
    candIdx.resize((size_t)nq * 64 + 1024);
    candDist.resize(candIdx.size());
    int rc = orbm_window_candidates(Device(), &grid.c, W.q.data(), W.desc.data(), nq, offset.data(), count.data(), candIdx.data(),
                                    candDist.data(), (int)candIdx.size(), &total);
    if (rc == ORB_E_CAPACITY) {
      candIdx.resize((size_t)total);
      candDist.resize((size_t)total);
      rc = orbm_window_candidates(Device(), &grid.c, W.q.data(), W.desc.data(), nq, offset.data(), count.data(), candIdx.data(),
                                  candDist.data(), total, &total);
    }
    Check(rc, "ORBmatcher::SearchForInitialization");
  }
  for (int k = 0; k < nq; ++k) {
    const int i1 = W.src[k];
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int c = offset[k]; c < offset[k] + count[k]; ++c) {
      const int i2 = candIdx[c], dist = candDist[c];
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if (bestDist <= TH_LOW) {
      if (bestDist < (float)bestDist2 * mfNNratio) {
        if (vnMatches21[bestIdx2] >= 0) {
          vnMatches12[vnMatches21[bestIdx2]] = -1;
          nmatches--;
        }
        vnMatches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (mbCheckOrientation) rotHist[RotationBin(F1.mvKeysUn[i1].angle, F2.mvKeysUn[bestIdx2].angle)].push_back(i1);
      }
    }
  }
  if (mbCheckOrientation) {  // :685-707
    int ind1, ind2, ind3;
    ThreeMaxima(rotHist, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        const int idx1 = rotHist[i][j];
        if (vnMatches12[idx1] >= 0) {
          vnMatches12[idx1] = -1;
          nmatches--;
        }
      }
    }
  }
  for (size_t i1 = 0, iend1 = vnMatches12.size(); i1 < iend1; i1++)  // :710-712
    if (vnMatches12[i1] >= 0) vbPrevMatched[i1] = F2.mvKeysUn[vnMatches12[i1]].pt;
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
