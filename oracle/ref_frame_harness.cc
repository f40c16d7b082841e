// oracle/ref_frame_harness.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C-ABI wrapper around the reference's own ORB_SLAM2::Frame and ORB_SLAM2::MapPoint, compiled from
// /root/reference/src/Frame.cc and src/MapPoint.cc UNMODIFIED (by oracle/Makefile, into
// oracle/_ref/libframeref.so) with their own Frame.h / MapPoint.h / ORBmatcher.h / ORBextractor.h /
// ORBVocabulary.h, the reference's src/ORBextractor.cc and the vendored DBoW2 Bow/FeatureVector, over
// oracle/cvshim.  Only KeyFrame, Map and Converter are stand-ins (oracle/mock_frame/).
// ORBmatcher::TH_HIGH / TH_LOW / DescriptorDistance are NOT restated: libframeref.so imports them
// from oracle/_ref/libmatchref.so, i.e. from the reference's compiled src/ORBmatcher.cc.
//
// Pins (tests/test_frame_ref_cpu.py):
//   Frame::ComputeStereoMatches (Frame.cc:527-700)            -> matchport::compute_stereo_matches
//   Frame::AssignFeaturesToGrid / GetFeaturesInArea (:245-260, :342-395) -> matchport::features_in_area
//   MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:310-369) -> matchport::distinctive_descriptor
//   Frame::UndistortKeyPoints / ComputeImageBounds / ComputeStereoFromRGBD / AssignFeaturesToGrid
//   (:419-485, :703-727, :245-260; tests/test_frame_finish_cpu.py)     -> matchport::frame_finish
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <limits>
#include <list>
#include <map>
#include <mutex>
#include <numeric>
#include <set>
#include <sstream>
#include <string>
#include <thread>
#include <vector>

#include <opencv2/opencv.hpp>
#include "Thirdparty/DBoW2/DBoW2/FORB.h"
#include "Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h"

#define private public      // reach Frame::AssignFeaturesToGrid and MapPoint's fields from the harness;
#define protected public    // the class layouts are unchanged and the reference's .cc files are compiled without this
#include "Frame.h"
#include "MapPoint.h"
#include "ORBmatcher.h"
#undef private
#undef protected

using namespace ORB_SLAM2;

extern "C" {

void frameref_constants(int* out) {
  out[0] = ORBmatcher::TH_HIGH;
  out[1] = ORBmatcher::TH_LOW;
  out[2] = FRAME_GRID_COLS;
  out[3] = FRAME_GRID_ROWS;
}

// Frame::ComputeStereoMatches on given keypoints / descriptors / pyramid levels.  The pyramid levels
// are wrapped (not copied) into the two extractors' public mvImagePyramid, as the extractor leaves
// them after operator().  Needs at least one surviving match (the reference indexes an empty vector
// otherwise, Frame.cc:687-688).
int frameref_stereo(int n_features, float scale_factor, int n_levels, const float* lx, const float* ly, const int* loct,
                    const uint8_t* ldesc, int nl, const float* rx, const float* ry, const int* roct, const uint8_t* rdesc, int nr,
                    const uint8_t* const* pxL, const uint8_t* const* pxR, const int* step, const int* w, const int* h, float mb,
                    float mbf, float* u_right, float* depth) {
  ORBextractor el(n_features, scale_factor, n_levels, 20, 7), er(n_features, scale_factor, n_levels, 20, 7);
  for (int l = 0; l < n_levels; ++l) {
    el.mvImagePyramid[l] = cv::Mat(h[l], w[l], CV_8U, (void*)pxL[l], (size_t)step[l]);
    er.mvImagePyramid[l] = cv::Mat(h[l], w[l], CV_8U, (void*)pxR[l], (size_t)step[l]);
  }
  Frame F;
  F.mpORBextractorLeft = &el;
  F.mpORBextractorRight = &er;
  F.N = nl;
  F.mvKeys.resize((size_t)nl);
  for (int i = 0; i < nl; ++i) F.mvKeys[i] = cv::KeyPoint(lx[i], ly[i], 31.f, -1.f, 0.f, loct[i]);
  F.mvKeysRight.resize((size_t)nr);
  for (int i = 0; i < nr; ++i) F.mvKeysRight[i] = cv::KeyPoint(rx[i], ry[i], 31.f, -1.f, 0.f, roct[i]);
  F.mDescriptors = cv::Mat(nl, 32, CV_8U, (void*)ldesc, 32);
  F.mDescriptorsRight = cv::Mat(nr, 32, CV_8U, (void*)rdesc, 32);
  F.mvScaleFactors = el.GetScaleFactors();         // Frame.cc:66-72
  F.mvInvScaleFactors = el.GetInverseScaleFactors();
  F.mb = mb;
  F.mbf = mbf;
  F.ComputeStereoMatches();
  for (int i = 0; i < nl; ++i) { u_right[i] = F.mvuRight[i]; depth[i] = F.mvDepth[i]; }
  return 0;
}

// Frame grid: AssignFeaturesToGrid over mvKeysUn with the given image bounds, then one
// GetFeaturesInArea per query.  out_idx holds the lists back to back, out_off[nq + 1] their offsets.
int frameref_features_in_area(const float* kx, const float* ky, const int* koct, int n, float min_x, float max_x, float min_y,
                              float max_y, const float* qx, const float* qy, const float* qr, const int* qmin, const int* qmax,
                              int nq, int* out_off, int* out_idx, int cap) {
  Frame::mnMinX = min_x; Frame::mnMaxX = max_x; Frame::mnMinY = min_y; Frame::mnMaxY = max_y;
  Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / static_cast<float>(max_x - min_x);   // Frame.cc:98-99
  Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / static_cast<float>(max_y - min_y);
  Frame F;
  F.N = n;
  F.mvKeysUn.resize((size_t)n);
  for (int i = 0; i < n; ++i) F.mvKeysUn[i] = cv::KeyPoint(kx[i], ky[i], 31.f, -1.f, 0.f, koct[i]);
  F.AssignFeaturesToGrid();
  int total = 0;
  for (int q = 0; q < nq; ++q) {
    out_off[q] = total;
    const std::vector<size_t> v = F.GetFeaturesInArea(qx[q], qy[q], qr[q], qmin[q], qmax[q]);
    for (size_t j = 0; j < v.size(); ++j) {
      if (total < cap) out_idx[total] = (int)v[j];
      ++total;
    }
  }
  out_off[nq] = total;
  return total <= cap ? total : -total;
}

// The rest of the Frame constructors after extraction (Frame.cc:85-120): ComputeImageBounds on a
// w x h image, the grid-element inverses exactly as :106-107 spell them, UndistortKeyPoints,
// ComputeStereoFromRGBD (when depth != NULL) and AssignFeaturesToGrid -- all the reference's own code.
// cam = fx fy cx cy, dist = k1 k2 p1 p2 [k3] (n_dist 4 or 5, as Tracking.cc:66-77 builds mDistCoef).
// bounds_out = mnMinX mnMaxX mnMinY mnMaxY mfGridElementWidthInv mfGridElementHeightInv.
// cell_off has FRAME_GRID_COLS*FRAME_GRID_ROWS + 1 entries in (ix, iy) -> ix*ROWS+iy order.
int frameref_finish(const float* kx, const float* ky, int n, const float* cam, const float* dist, int n_dist, float bf, int w, int h,
                    const float* depth, float* bounds_out, float* xy_un, float* u_right, float* depth_out, int* cell_off,
                    int* cell_idx) {
  cv::Mat K = cv::Mat::zeros(3, 3, CV_32F);
  K.at<float>(0, 0) = cam[0]; K.at<float>(1, 1) = cam[1]; K.at<float>(0, 2) = cam[2]; K.at<float>(1, 2) = cam[3];
  K.at<float>(2, 2) = 1.f;
  cv::Mat D(n_dist, 1, CV_32F);
  for (int i = 0; i < n_dist; ++i) D.at<float>(i) = dist[i];
  Frame F;
  F.mK = K.clone();
  F.mDistCoef = D.clone();
  F.mbf = bf;
  F.N = n;
  F.mvKeys.resize((size_t)n);
  for (int i = 0; i < n; ++i) F.mvKeys[i] = cv::KeyPoint(kx[i], ky[i], 31.f, -1.f, 0.f, 0);
  cv::Mat im(h, w, CV_8U);
  F.UndistortKeyPoints();
  F.ComputeImageBounds(im);
  Frame::mfGridElementWidthInv = static_cast<float>(FRAME_GRID_COLS) / (Frame::mnMaxX - Frame::mnMinX);    // Frame.cc:106-107
  Frame::mfGridElementHeightInv = static_cast<float>(FRAME_GRID_ROWS) / (Frame::mnMaxY - Frame::mnMinY);
  bounds_out[0] = Frame::mnMinX; bounds_out[1] = Frame::mnMaxX; bounds_out[2] = Frame::mnMinY; bounds_out[3] = Frame::mnMaxY;
  bounds_out[4] = Frame::mfGridElementWidthInv; bounds_out[5] = Frame::mfGridElementHeightInv;
  if (depth) {
    cv::Mat imDepth(h, w, CV_32F, (void*)depth, (size_t)w * 4);
    F.ComputeStereoFromRGBD(imDepth);
    for (int i = 0; i < n; ++i) { u_right[i] = F.mvuRight[i]; depth_out[i] = F.mvDepth[i]; }
  }
  F.AssignFeaturesToGrid();
  for (int i = 0; i < n; ++i) { xy_un[2 * i] = F.mvKeysUn[i].pt.x; xy_un[2 * i + 1] = F.mvKeysUn[i].pt.y; }
  int total = 0;
  for (int ix = 0; ix < FRAME_GRID_COLS; ++ix)
    for (int iy = 0; iy < FRAME_GRID_ROWS; ++iy) {
      cell_off[ix * FRAME_GRID_ROWS + iy] = total;
      for (size_t j = 0; j < F.mGrid[ix][iy].size(); ++j) cell_idx[total++] = (int)F.mGrid[ix][iy][j];
    }
  cell_off[FRAME_GRID_COLS * FRAME_GRID_ROWS] = total;
  return total;
}

// MapPoint::ComputeDistinctiveDescriptors for n_points map points; point p is observed by
// offset[p+1]-offset[p] keyframes whose descriptors are rows offset[p].. of desc (one keyframe per
// observation, allocated in one array so that the std::map<KeyFrame*, size_t> iterates in row
// order).  bad[row] != 0 marks a bad keyframe (skipped, MapPoint.cc:328).  Writes the chosen
// 32-byte descriptor of each point.
int frameref_distinctive(const uint8_t* desc, const int* offset, const uint8_t* bad, int n_points, uint8_t* out) {
  Map map;
  for (int p = 0; p < n_points; ++p) {
    const int n = offset[p + 1] - offset[p];
    std::vector<KeyFrame> kfs((size_t)n);
    for (int i = 0; i < n; ++i) {
      KeyFrame& K = kfs[(size_t)i];
      K.mnId = (unsigned long)i;
      K.mDescriptors = cv::Mat(1, 32, CV_8U, (void*)(desc + 32 * (size_t)(offset[p] + i)), 32);
      K.mvuRight.assign(1, -1.f);
      K.mvpMapPoints.assign(1, nullptr);
      K.mbBad = bad && bad[offset[p] + i];
    }
    cv::Mat pos = cv::Mat::zeros(3, 1, CV_32F);
    MapPoint* mp = new MapPoint(pos, &kfs[0], &map);
    for (int i = 0; i < n; ++i) mp->AddObservation(&kfs[(size_t)i], 0);
    mp->ComputeDistinctiveDescriptors();
    cv::Mat d = mp->GetDescriptor();
    if (d.empty()) memset(out + 32 * (size_t)p, 0, 32);
    else memcpy(out + 32 * (size_t)p, d.ptr(0), 32);
    delete mp;
  }
  return 0;
}

}  // extern "C"
