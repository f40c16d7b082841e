// oracle/match_port.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the reference's descriptor-search paths on flat arrays.
// It follows src/ORBmatcher.cc line by line:
//   hamming256          ORBmatcher.cc:1848-1864 (DescriptorDistance)
//   top2 update         ORBmatcher.cc:417-426
//   search_bow_kf_frame ORBmatcher.cc:360-489
//   search_bow_kf_kf    ORBmatcher.cc:723-856
//   three_maxima        ORBmatcher.cc:1802-1843
//   stereo_search       Frame.cc:527-610 (descriptor part of ComputeStereoMatches)
//   compute_stereo_matches Frame.cc:527-700 (whole: + sliding-window refinement, parabola, median filter)
//   features_in_area    Frame.cc:342-395 (GetFeaturesInArea)
//   window_search       ORBmatcher.cc:46-132 (and the best-only loops of :491-604, :1529-1800)
//   search_triangulation ORBmatcher.cc:858-1024 + :341-358
//   distinctive_descriptor MapPoint.cc:333-363 (ComputeDistinctiveDescriptors)
//   undistort_keypoints / image_bounds / assign_features_to_grid / stereo_from_rgbd  Frame.cc:419-485, :245-260, :703-727
//     (pinned on the reference's own Frame.cc in libframeref.so and, for cv::undistortPoints, on cv2 4.13: tests/test_frame_finish_cpu.py)
// Parity status: hamming256, search_bow (both overloads) and window_search are pinned on the
// reference's own src/ORBmatcher.cc (as is search_triangulation) compiled unmodified over mocks (oracle/_ref/libmatchref.so):
// tests/test_projection_cpu.py runs the product header over this restatement against it on 31
// scenes.  stereo_search / compute_stereo_matches, features_in_area and distinctive_descriptor are
// pinned on the reference's own src/Frame.cc and src/MapPoint.cc compiled unmodified
// (oracle/_ref/libframeref.so, tests/test_frame_ref_cpu.py: bit-equal mvuRight / mvDepth, identical
// index lists, identical chosen descriptors).  top2_bruteforce / merge_top2 have no counterpart in
// the reference (its distance and update rule are the pinned ones above): known answers + an
// independent numpy model in tests/test_oracle_match.py.
#pragma once
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "cvprims.hpp"

namespace matchport {

static const int TH_HIGH = 95, TH_LOW = 45, HISTO_LENGTH = 30;  // ORBmatcher.cc:37-40

// The SWAR bit count of the reference, on eight 32-bit words.
static inline int hamming256(const uint8_t* a, const uint8_t* b) {
  int dist = 0;
  for (int i = 0; i < 8; ++i) {
    uint32_t x, y;
    memcpy(&x, a + 4 * i, 4);
    memcpy(&y, b + 4 * i, 4);
    uint32_t v = x ^ y;
    v = v - ((v >> 1) & 0x55555555u);
    v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
    dist += (int)((((v + (v >> 4)) & 0x0F0F0F0Fu) * 0x01010101u) >> 24);
  }
  return dist;
}

struct Top2 {
  int best1 = 256, best2 = 256;
  long long idx = -1;
  inline void feed(int d, long long j) {
    if (d < best1) { best2 = best1; best1 = d; idx = j; }
    else if (d < best2) best2 = d;
  }
};

// accept rule: ORBmatcher.cc:429-431 (<= th) and :799-801 (strict <)
static inline bool accept(int best1, int best2, float ratio, int th, bool strict) {
  if (strict ? !(best1 < th) : !(best1 <= th)) return false;
  return (float)best1 < ratio * (float)best2;
}

// Brute-force top-2 of every query against rows [0,nd) of db.
static inline void top2_bruteforce(const uint8_t* q, int nq, const uint8_t* db, long long nd,
                                   int* d1, long long* i1, int* d2) {
  for (int i = 0; i < nq; ++i) {
    Top2 t;
    for (long long j = 0; j < nd; ++j) t.feed(hamming256(q + 32 * (size_t)i, db + 32 * (size_t)j), j);
    d1[i] = t.best1; i1[i] = t.idx; d2[i] = t.best2;
  }
}

static inline void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; ++i) {
    const int s = (int)histo[i].size();
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) ind3 = -1;
}

static inline int rot_bin(float a1, float a2) {
  const float factor = 1.0f / HISTO_LENGTH;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)round(rot * factor);
  if (bin == HISTO_LENGTH) bin = 0;
  return bin;
}

// A DBoW2::FeatureVector flattened: node ids ascending (std::map order), CSR offsets,
// feature indices in insertion order.
struct FeatVec {
  const uint32_t* node; const int* off; const uint32_t* idx; int n_nodes;
};
static inline int lower_bound_node(const FeatVec& f, int from, uint32_t id) {
  int lo = from, hi = f.n_nodes;
  while (lo < hi) { int m = (lo + hi) / 2; if (f.node[m] < id) lo = m + 1; else hi = m; }
  return lo;
}

// Shared body of both SearchByBoW overloads.
//   side 1 = KeyFrame (its features drive the outer loop), side 2 = Frame / KeyFrame 2.
//   valid1[i]: feature i of side 1 has a good MapPoint.  valid2 == nullptr (KF-Frame) or
//   flags for side 2 (KF-KF).  KF-Frame: match2[f] = index on side 1 claiming frame feature
//   f, and `claimed` is match2 itself; KF-KF: match1[i1] = idx2, claimed = vbMatched2.
static inline int search_bow(const uint8_t* desc1, int n1, const uint8_t* desc2, int n2,
                             const FeatVec& fv1, const FeatVec& fv2, const uint8_t* valid1,
                             const uint8_t* valid2, const float* ang1, const float* ang2,
                             float ratio, int check_ori, bool kf_kf, int* out) {
  const int n_out = kf_kf ? n1 : n2;
  for (int i = 0; i < n_out; ++i) out[i] = -1;
  std::vector<uint8_t> claimed2((size_t)n2, 0);
  std::vector<int> hist[HISTO_LENGTH];
  int nmatches = 0;
  int a = 0, b = 0;
  while (a < fv1.n_nodes && b < fv2.n_nodes) {
    if (fv1.node[a] == fv2.node[b]) {
      for (int p = fv1.off[a]; p < fv1.off[a + 1]; ++p) {
        const int i1 = (int)fv1.idx[p];
        if (!valid1[i1]) continue;
        Top2 t;
        for (int r = fv2.off[b]; r < fv2.off[b + 1]; ++r) {
          const int i2 = (int)fv2.idx[r];
          if (claimed2[i2]) continue;
          if (valid2 && !valid2[i2]) continue;
          t.feed(hamming256(desc1 + 32 * (size_t)i1, desc2 + 32 * (size_t)i2), i2);
        }
        if (accept(t.best1, t.best2, ratio, TH_LOW, kf_kf)) {
          const int i2 = (int)t.idx;
          claimed2[i2] = 1;
          if (kf_kf) out[i1] = i2; else out[i2] = i1;
          if (check_ori) hist[rot_bin(ang1[i1], ang2[i2])].push_back(kf_kf ? i1 : i2);
          ++nmatches;
        }
      }
      ++a; ++b;
    } else if (fv1.node[a] < fv2.node[b]) {
      a = lower_bound_node(fv1, a, fv2.node[b]);
    } else {
      b = lower_bound_node(fv2, b, fv1.node[a]);
    }
  }
  if (check_ori) {
    int i1 = -1, i2 = -1, i3 = -1;
    three_maxima(hist, HISTO_LENGTH, i1, i2, i3);
    for (int i = 0; i < HISTO_LENGTH; ++i) {
      if (i == i1 || i == i2 || i == i3) continue;
      for (int j : hist[i]) { out[j] = -1; --nmatches; }
    }
  }
  return nmatches;
}

// Frame.cc:537-610.  kp arrays are (x, y, octave); out best_idx/best_dist per left keypoint
// (best_dist == TH_HIGH and best_idx == 0 when nothing beat the initial bound; best_dist == -1
// when the keypoint was skipped before the scan).
static inline void stereo_search(const float* lx, const float* ly, const int* loct, const uint8_t* ldesc, int nl,
                                 const float* rx, const float* ry, const int* roct, const uint8_t* rdesc, int nr,
                                 const float* scale_factors, int n_rows, float min_d, float max_d,
                                 int* best_idx, int* best_dist) {
  std::vector<std::vector<int>> rows((size_t)n_rows);
  for (int i = 0; i < nr; ++i) {
    const float r = 2.0f * scale_factors[roct[i]];
    const int maxr = (int)ceil(ry[i] + r), minr = (int)floor(ry[i] - r);
    for (int y = minr; y <= maxr; ++y)
      if (y >= 0 && y < n_rows) rows[y].push_back(i);  // the reference indexes unchecked
  }
  for (int i = 0; i < nl; ++i) {
    best_idx[i] = 0;
    best_dist[i] = -1;
    const int row = (int)ly[i];
    if (row < 0 || row >= n_rows) continue;
    const std::vector<int>& cand = rows[row];
    if (cand.empty()) continue;
    const float minU = lx[i] - max_d, maxU = lx[i] - min_d;
    if (maxU < 0) continue;
    int bd = TH_HIGH, bi = 0;
    for (int c : cand) {
      if (roct[c] < loct[i] - 1 || roct[c] > loct[i] + 1) continue;
      if (rx[c] >= minU && rx[c] <= maxU) {
        const int d = hamming256(ldesc + 32 * (size_t)i, rdesc + 32 * (size_t)c);
        if (d < bd) { bd = d; bi = c; }
      }
    }
    best_idx[i] = bi;
    best_dist[i] = bd;
  }
}

// Frame::ComputeStereoMatches, whole: Frame.cc:527-700.  Pyramid levels as the extractor leaves them
// (mvImagePyramid[l]: pointer to pixel (0,0) of level l, row stride, width, height).
struct PyrLevel { const uint8_t* px; size_t step; int w, h; };

static inline void compute_stereo_matches(const float* lx, const float* ly, const int* loct, const uint8_t* ldesc, int nl,
                                          const float* rx, const float* ry, const int* roct, const uint8_t* rdesc, int nr,
                                          const float* mvScaleFactors, const float* mvInvScaleFactors,
                                          const PyrLevel* pyrL, const PyrLevel* pyrR, float mb, float mbf,
                                          float* mvuRight, float* mvDepth) {
  for (int i = 0; i < nl; ++i) { mvuRight[i] = -1.0f; mvDepth[i] = -1.0f; }
  const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
  const int nRows = pyrL[0].h;
  const float minZ = mb;
  const float minD = 0;
  const float maxD = mbf / minZ;
  std::vector<int> best_idx((size_t)nl), best_dist((size_t)nl);
  stereo_search(lx, ly, loct, ldesc, nl, rx, ry, roct, rdesc, nr, mvScaleFactors, nRows, minD, maxD, best_idx.data(),
                best_dist.data());
  std::vector<std::pair<int, int> > vDistIdx;
  for (int iL = 0; iL < nl; ++iL) {
    if (best_dist[iL] < 0) continue;  // skipped before the scan
    const int levelL = loct[iL];
    const float uL = lx[iL];
    if (best_dist[iL] < thOrbDist) {  // :613
      const float uR0 = rx[best_idx[iL]];
      const float scaleFactor = mvInvScaleFactors[levelL];
      const float scaleduL = round(lx[iL] * scaleFactor);
      const float scaledvL = round(ly[iL] * scaleFactor);
      const float scaleduR0 = round(uR0 * scaleFactor);
      const int w = 5;
      const PyrLevel& PL = pyrL[levelL];
      const PyrLevel& PR = pyrR[levelL];
      // IL = patch - its centre, as float (:624-626)
      float IL[11][11];
      {
        const int r0 = (int)(scaledvL - w), c0 = (int)(scaleduL - w);
        const float centre = (float)PL.px[(size_t)(r0 + w) * PL.step + c0 + w];
        for (int r = 0; r < 2 * w + 1; ++r)
          for (int c = 0; c < 2 * w + 1; ++c) IL[r][c] = (float)PL.px[(size_t)(r0 + r) * PL.step + c0 + c] - centre;
      }
      int bestDist = INT_MAX;
      int bestincR = 0;
      const int L = 5;
      std::vector<float> vDists(2 * L + 1);
      const float iniu = scaleduR0 + L - w;
      const float endu = scaleduR0 + L + w + 1;
      if (iniu < 0 || endu >= PR.w) continue;
      for (int incR = -L; incR <= +L; incR++) {
        const int r0 = (int)(scaledvL - w), c0 = (int)(scaleduR0 + incR - w);
        const float centre = (float)PR.px[(size_t)(r0 + w) * PR.step + c0 + w];
        double acc = 0;  // cv::norm(IL, IR, NORM_L1)
        for (int r = 0; r < 2 * w + 1; ++r)
          for (int c = 0; c < 2 * w + 1; ++c)
            acc += std::fabs((double)IL[r][c] - (double)((float)PR.px[(size_t)(r0 + r) * PR.step + c0 + c] - centre));
        const float dist = (float)acc;
        if (dist < bestDist) {
          bestDist = (int)dist;
          bestincR = incR;
        }
        vDists[L + incR] = dist;
      }
      if (bestincR == -L || bestincR == L) continue;
      const float dist1 = vDists[L + bestincR - 1];
      const float dist2 = vDists[L + bestincR];
      const float dist3 = vDists[L + bestincR + 1];
      const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
      if (deltaR < -1 || deltaR > 1) continue;
      float bestuR = mvScaleFactors[levelL] * ((float)scaleduR0 + (float)bestincR + deltaR);
      float disparity = (uL - bestuR);
      if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) {
          disparity = 0.01;
          bestuR = uL - 0.01;
        }
        mvDepth[iL] = mbf / disparity;
        mvuRight[iL] = bestuR;
        vDistIdx.push_back(std::pair<int, int>(bestDist, iL));
      }
    }
  }
  if (vDistIdx.empty()) return;  // the reference reads vDistIdx[0] of an empty vector here
  std::sort(vDistIdx.begin(), vDistIdx.end());
  const float median = vDistIdx[vDistIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
    if (vDistIdx[i].first < thDist) break;
    mvuRight[vDistIdx[i].second] = -1;
    mvDepth[vDistIdx[i].second] = -1;
  }
}

// Merge of per-shard top-2 results into the result of one sequential scan over the
// concatenated shards.  Shards are ordered by increasing global row index.
static inline void merge_top2(const int* d1, const long long* i1, const int* d2, int n_shards, int nq,
                              int* od1, long long* oi1, int* od2) {
  for (int q = 0; q < nq; ++q) {
    Top2 t;
    for (int s = 0; s < n_shards; ++s) {
      const size_t k = (size_t)s * nq + q;
      // feeding (best1, then best2) of each shard in order reproduces the sequential rule:
      // best1 keeps the lowest global index on ties, best2 is the multiset second minimum.
      if (i1[k] >= 0) t.feed(d1[k], i1[k]);
      if (d2[k] < 256) t.feed(d2[k], -2);
    }
    od1[q] = t.best1; oi1[q] = t.idx; od2[q] = t.best2;
  }
}

// ---- Windowed projection search ---------------------------------------------
// Flat-array restatement of the loop shared by the SearchByProjection overloads.  The searched
// frame is given as the reference holds it: keypoints, mvuRight, descriptors and the cell grid
// (cell (ix, iy) at ix*rows+iy).
struct GridFrame {
  const float* x; const float* y; const int* octave; const float* u_right; const uint8_t* desc; int n;
  float min_x, min_y, width_inv, height_inv; int cols, rows;
  const int* cell_off; const uint32_t* cell_idx;
};

// Frame::GetFeaturesInArea, Frame.cc:342-395 (KeyFrame.cc:613-652 is the same without levels).
static inline void features_in_area(const GridFrame& F, float x, float y, float r, int minLevel, int maxLevel,
                                    std::vector<int>& vIndices) {
  vIndices.clear();
  const int nMinCellX = std::max(0, (int)floor((x - F.min_x - r) * F.width_inv));
  if (nMinCellX >= F.cols) return;
  const int nMaxCellX = std::min(F.cols - 1, (int)ceil((x - F.min_x + r) * F.width_inv));
  if (nMaxCellX < 0) return;
  const int nMinCellY = std::max(0, (int)floor((y - F.min_y - r) * F.height_inv));
  if (nMinCellY >= F.rows) return;
  const int nMaxCellY = std::min(F.rows - 1, (int)ceil((y - F.min_y + r) * F.height_inv));
  if (nMaxCellY < 0) return;
  const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
  for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
      const int c = ix * F.rows + iy;
      for (int j = F.cell_off[c]; j < F.cell_off[c + 1]; j++) {
        const int k = (int)F.cell_idx[j];
        if (bCheckLevels) {
          if (F.octave[k] < minLevel) continue;
          if (maxLevel >= 0)
            if (F.octave[k] > maxLevel) continue;
        }
        const float distx = F.x[k] - x;
        const float disty = F.y[k] - y;
        if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(k);
      }
    }
  }
}

// The rest of the Frame constructors (Frame.cc:85-120) after extraction, restated:
//   UndistortKeyPoints :419-449, ComputeImageBounds :451-485, ComputeStereoFromRGBD :703-727,
//   AssignFeaturesToGrid :245-260 with PosInGrid :397-408.
// cam = fx fy cx cy; dist = k1 k2 p1 p2 [k3].  cv::undistortPoints is cvprims::undistort_points.
static inline void camera_k9(const float* cam, float* K9) {
  const float k[9] = {cam[0], 0.f, cam[2], 0.f, cam[1], cam[3], 0.f, 0.f, 1.f};
  memcpy(K9, k, sizeof(k));
}
static inline void undistort_keypoints(const float* xy, int n, const float* cam, const float* dist, int n_dist, float* xy_un) {
  if (dist[0] == 0.0) {  // mDistCoef.at<float>(0)==0.0: mvKeysUn = mvKeys
    memcpy(xy_un, xy, (size_t)n * 2 * sizeof(float));
    return;
  }
  float K9[9];
  camera_k9(cam, K9);
  cvprims::undistort_points(xy, n, K9, dist, n_dist, K9, xy_un);
}
// bounds = mnMinX mnMaxX mnMinY mnMaxY
static inline void image_bounds(int w, int h, const float* cam, const float* dist, int n_dist, float* bounds) {
  if (dist[0] != 0.0) {
    float m[8] = {0.f, 0.f, (float)w, 0.f, 0.f, (float)h, (float)w, (float)h};
    undistort_keypoints(m, 4, cam, dist, n_dist, m);
    bounds[0] = std::min(m[0], m[4]);
    bounds[1] = std::max(m[2], m[6]);
    bounds[2] = std::min(m[1], m[3]);
    bounds[3] = std::max(m[5], m[7]);
  } else {
    bounds[0] = 0.0f; bounds[1] = (float)w; bounds[2] = 0.0f; bounds[3] = (float)h;
  }
}
// cells[(ix, iy) -> ix*rows+iy] = keypoint indices in push_back order
static inline void assign_features_to_grid(const float* xy_un, int n, float min_x, float min_y, float width_inv, float height_inv,
                                           int cols, int rows, std::vector<std::vector<int>>& cells) {
  cells.assign((size_t)cols * rows, std::vector<int>());
  for (int i = 0; i < n; ++i) {
    const int posX = (int)round((xy_un[2 * i] - min_x) * width_inv);
    const int posY = (int)round((xy_un[2 * i + 1] - min_y) * height_inv);
    if (posX < 0 || posX >= cols || posY < 0 || posY >= rows) continue;
    cells[(size_t)posX * rows + posY].push_back(i);
  }
}
static inline void stereo_from_rgbd(const float* xy, const float* xy_un, int n, const float* depth, int depth_w, float mbf,
                                    float* u_right, float* depth_out) {
  for (int i = 0; i < n; ++i) {
    u_right[i] = -1.f;
    depth_out[i] = -1.f;
    const float v = xy[2 * i + 1], u = xy[2 * i];
    const float d = depth[(size_t)(int)v * depth_w + (int)u];
    if (d > 0) {
      depth_out[i] = d;
      u_right[i] = xy_un[2 * i] - mbf / d;
    }
  }
}

struct WindowQuery { float x, y, r, u_right, er_max; int min_level, max_level, blocks; };
struct WindowResult { int idx, best_idx, best_dist, best_dist2, best_level, best_level2; };

// The query loop of ORBmatcher.cc:52-129 (best + second, level-aware ratio rule), which with
// ratio <= 0 is the best-only loop of :1553-1640, :1687-1775 and :511-601.  `hidden[k]` plays
// mvpMapPoints[k] (with Observations() > 0): set before the call for keypoints already matched,
// and set by an accepted query whose MapPoint blocks (claim mode).
// Optional reprojection gate: the candidate test of ORBmatcher::Fuse, ORBmatcher.cc:1118-1146.
struct WindowGate { const float* inv_level_sigma2; double chi2_mono, chi2_stereo; };

static inline int window_search(const GridFrame& F, const WindowQuery* q, const uint8_t* qdesc, int nq,
                                const uint8_t* claimed, int th, float ratio, bool claim, WindowResult* out,
                                const WindowGate* gate = nullptr) {
  std::vector<uint8_t> hidden(F.n, 0);
  if (claimed) for (int k = 0; k < F.n; ++k) hidden[k] = claimed[k] != 0;
  std::vector<int> vIndices;
  int nmatches = 0;
  for (int i = 0; i < nq; ++i) {
    const WindowQuery& Q = q[i];
    features_in_area(F, Q.x, Q.y, Q.r, Q.min_level, Q.max_level, vIndices);
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (size_t v = 0; v < vIndices.size(); ++v) {
      const int idx = vIndices[v];
      if (hidden[idx]) continue;  // :81-83
      if (Q.er_max >= 0 && F.u_right && F.u_right[idx] > 0) {  // :85-90
        const float er = fabs(Q.u_right - F.u_right[idx]);
        if (er > Q.er_max) continue;
      }
      if (gate) {
        const int kpLevel = F.octave[idx];
        const float u = Q.x, v = Q.y, ur = Q.u_right;
        const float kpr = F.u_right ? F.u_right[idx] : -1.0f;
        if (kpr >= 0) {
          const float ex = u - F.x[idx];
          const float ey = v - F.y[idx];
          const float er = ur - kpr;
          const float e2 = ex * ex + ey * ey + er * er;
          if (e2 * gate->inv_level_sigma2[kpLevel] > gate->chi2_stereo) continue;
        } else {
          const float ex = u - F.x[idx];
          const float ey = v - F.y[idx];
          const float e2 = ex * ex + ey * ey;
          if (e2 * gate->inv_level_sigma2[kpLevel] > gate->chi2_mono) continue;
        }
      }
      const int dist = hamming256(qdesc + 32 * (size_t)i, F.desc + 32 * (size_t)idx);
      if (dist < bestDist) {  // :97-110
        bestDist2 = bestDist; bestDist = dist;
        bestLevel2 = bestLevel; bestLevel = F.octave[idx];
        bestIdx = idx;
      } else if (dist < bestDist2) {
        bestLevel2 = F.octave[idx];
        bestDist2 = dist;
      }
    }
    WindowResult R{-1, bestIdx, bestDist, bestDist2, bestLevel, bestLevel2};
    if (bestIdx >= 0 && bestDist <= th) {  // :114-126
      if (!(ratio > 0 && bestLevel == bestLevel2 && bestDist > ratio * bestDist2)) {
        R.idx = bestIdx;
        if (claim && Q.blocks) hidden[bestIdx] = 1;
        nmatches++;
      }
    }
    out[i] = R;
  }
  return nmatches;
}

// ---- SearchForTriangulation, ORBmatcher.cc:858-1024 (with CheckDistEpipolarLine, :341-358) ------
struct TriKeyFrame {
  const uint8_t* desc; const float* x; const float* y; const int* octave; const float* angle;
  const uint8_t* usable;  // GetMapPoint(i) == NULL && (!bOnlyStereo || mvuRight[i] >= 0)
  const uint8_t* stereo;  // mvuRight[i] >= 0
  int n; FeatVec fv;
};

static inline bool check_dist_epipolar_line(float x1, float y1, float x2, float y2, int octave2, const float* F12,
                                            const float* level_sigma2) {
  const float a = x1 * F12[0] + y1 * F12[3] + F12[6];
  const float b = x1 * F12[1] + y1 * F12[4] + F12[7];
  const float c = x1 * F12[2] + y1 * F12[5] + F12[8];
  const float num = a * x2 + b * y2 + c;
  const float den = a * a + b * b;
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * level_sigma2[octave2];
}

static inline int search_triangulation(const TriKeyFrame& K1, const TriKeyFrame& K2, const float* F12, float ex, float ey,
                                       const float* scale_factors2, const float* level_sigma2_2, int check_ori, int* vMatches12) {
  int nmatches = 0;
  for (int i = 0; i < K1.n; ++i) vMatches12[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  int a = 0, b = 0;
  while (a < K1.fv.n_nodes && b < K2.fv.n_nodes) {
    if (K1.fv.node[a] == K2.fv.node[b]) {
      for (int p = K1.fv.off[a]; p < K1.fv.off[a + 1]; ++p) {
        const int idx1 = (int)K1.fv.idx[p];
        if (!K1.usable[idx1]) continue;  // :894-903
        const bool bStereo1 = K1.stereo[idx1] != 0;
        int bestDist = TH_LOW, bestIdx2 = -1;
        for (int r = K2.fv.off[b]; r < K2.fv.off[b + 1]; ++r) {
          const int idx2 = (int)K2.fv.idx[r];
          if (!K2.usable[idx2]) continue;  // :913-925 (vbMatched2 is never set by the reference)
          const bool bStereo2 = K2.stereo[idx2] != 0;
          const int dist = hamming256(K1.desc + 32 * (size_t)idx1, K2.desc + 32 * (size_t)idx2);
          if (dist > TH_LOW || dist > bestDist) continue;
          if (!bStereo1 && !bStereo2) {
            const float distex = ex - K2.x[idx2];
            const float distey = ey - K2.y[idx2];
            if (distex * distex + distey * distey < 100 * scale_factors2[K2.octave[idx2]]) continue;
          }
          if (check_dist_epipolar_line(K1.x[idx1], K1.y[idx1], K2.x[idx2], K2.y[idx2], K2.octave[idx2], F12, level_sigma2_2)) {
            bestIdx2 = idx2;
            bestDist = dist;
          }
        }
        if (bestIdx2 >= 0) {
          vMatches12[idx1] = bestIdx2;
          nmatches++;
          if (check_ori) rotHist[rot_bin(K1.angle[idx1], K2.angle[bestIdx2])].push_back(idx1);
        }
      }
      ++a; ++b;
    } else if (K1.fv.node[a] < K2.fv.node[b]) {
      a = lower_bound_node(K1.fv, a, K2.fv.node[b]);
    } else {
      b = lower_bound_node(K2.fv, b, K1.fv.node[a]);
    }
  }
  if (check_ori) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        vMatches12[rotHist[i][j]] = -1;
        nmatches--;
      }
    }
  }
  return nmatches;
}

// MapPoint::ComputeDistinctiveDescriptors, MapPoint.cc:333-363, on the N descriptors of one point.
static inline int distinctive_descriptor(const uint8_t* desc, int N) {
  if (N <= 0) return -1;
  std::vector<float> Distances((size_t)N * N);
  for (int i = 0; i < N; i++) {
    Distances[(size_t)i * N + i] = 0;
    for (int j = i + 1; j < N; j++) {
      int distij = hamming256(desc + 32 * (size_t)i, desc + 32 * (size_t)j);
      Distances[(size_t)i * N + j] = distij;
      Distances[(size_t)j * N + i] = distij;
    }
  }
  int BestMedian = INT_MAX;
  int BestIdx = 0;
  for (int i = 0; i < N; i++) {
    std::vector<int> vDists(Distances.begin() + (size_t)i * N, Distances.begin() + (size_t)(i + 1) * N);
    std::sort(vDists.begin(), vDists.end());
    int median = vDists[0.5 * (N - 1)];
    if (median < BestMedian) {
      BestMedian = median;
      BestIdx = i;
    }
  }
  return BestIdx;
}

}  // namespace matchport
