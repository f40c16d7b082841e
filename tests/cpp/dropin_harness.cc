// tests/cpp/dropin_harness.cc -- exercises the PRODUCT's C++ host classes (include/ORBextractor.h,
// include/ORBmatcher.h) exactly the way the reference's callers use them (Frame.cc:265-267,
// Tracking.cc:1750-1790, LoopClosing.cc:239), behind a small C interface so pytest can compare the
// results with the oracle.  cv:: types come from oracle/cvshim (headers only -- no oracle code is
// executed by these entry points), Frame/KeyFrame/MapPoint from tests/cpp/mock.
#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <memory>

#include "ORBextractor.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM2;

struct DropKp { float x, y, size, angle, response; int octave, class_id; };

static thread_local std::string g_err;

extern "C" {

const char* dropin_error() { return g_err.c_str(); }

void* dropin_create(int nf, float sf, int nl, int ini, int mn) {
  try { return new ORBextractor(nf, sf, nl, ini, mn); }
  catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void dropin_destroy(void* h) { delete (ORBextractor*)h; }

int dropin_extract(void* h, const uint8_t* img, int w, int h_, size_t stride, DropKp* kps, uint8_t* desc, int cap) {
  ORBextractor* e = (ORBextractor*)h;
  try {
    cv::Mat im(h_, w, CV_8U, (void*)img, stride);
    std::vector<cv::KeyPoint> k(3);  // stale content must be replaced
    cv::Mat d;
    (*e)(im, cv::Mat(), k, d);
    const int n = (int)k.size();
    if (n > cap) return -n;
    if (n == 0 && !d.empty()) { g_err = "descriptors not released for zero keypoints"; return -1000000; }
    for (int i = 0; i < n; ++i) {
      kps[i] = DropKp{k[i].pt.x, k[i].pt.y, k[i].size, k[i].angle, k[i].response, k[i].octave, k[i].class_id};
      memcpy(desc + 32 * (size_t)i, d.ptr(i), 32);
    }
    return n;
  } catch (const std::exception& ex) { g_err = ex.what(); return -1000001; }
}

// bench.py configs.c1_single: `iters` calls of ORBextractor::operator() on one frame, the way Tracking drives it
// (Frame.cc:262-268); returns the median microseconds per call, with or without the mvImagePyramid download
double dropin_time_extract(void* h, const uint8_t* img, int w, int h_, size_t stride, int iters, int download_pyramid) {
  ORBextractor* e = (ORBextractor*)h;
  try {
    e->SetPyramidDownload(download_pyramid != 0);
    cv::Mat im(h_, w, CV_8U, (void*)img, stride);
    std::vector<cv::KeyPoint> k;
    cv::Mat d;
    std::vector<double> us;
    for (int i = 0; i < iters + 3; ++i) {
      const auto t0 = std::chrono::steady_clock::now();
      (*e)(im, cv::Mat(), k, d);
      const auto t1 = std::chrono::steady_clock::now();
      if (i >= 3) us.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
    }
    std::sort(us.begin(), us.end());
    return us.empty() ? -1.0 : us[us.size() / 2];
  } catch (const std::exception& ex) { g_err = ex.what(); return -1.0; }
}

int dropin_level_size(void* h, int l, int* w, int* h_) {
  ORBextractor* e = (ORBextractor*)h;
  if (l < 0 || l >= e->GetLevels() || e->mvImagePyramid[l].empty()) return -1;
  *w = e->mvImagePyramid[l].cols;
  *h_ = e->mvImagePyramid[l].rows;
  return 0;
}

// reads mvImagePyramid the way Frame::ComputeStereoMatches does: ROI pointer arithmetic incl. the border
int dropin_get_level(void* h, int l, uint8_t* dst, size_t dstep, int with_border) {
  ORBextractor* e = (ORBextractor*)h;
  if (l < 0 || l >= e->GetLevels() || e->mvImagePyramid[l].empty()) return -1;
  const cv::Mat& m = e->mvImagePyramid[l];
  const int b = with_border ? 19 : 0;
  for (int y = -b; y < m.rows + b; ++y)
    memcpy(dst + (size_t)(y + b) * dstep, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - b, (size_t)(m.cols + 2 * b));
  return 0;
}

float dropin_scale(void* h, int l) { return ((ORBextractor*)h)->GetScaleFactors()[l]; }
float dropin_table(void* h, int which, int l) {
  ORBextractor* e = (ORBextractor*)h;
  return which == 0 ? e->GetScaleFactors()[l] : which == 1 ? e->GetInverseScaleFactors()[l]
       : which == 2 ? e->GetScaleSigmaSquares()[l] : e->GetInverseScaleSigmaSquares()[l];
}
int dropin_levels(void* h) { return ((ORBextractor*)h)->GetLevels(); }
float dropin_scale_factor(void* h) { return ((ORBextractor*)h)->GetScaleFactor(); }

// Frame::Frame (stereo, Frame.cc:63-99) the drop-in way: two extractors, then the whole stereo matching on the device.
int dropin_stereo_frame(void* hl, void* hr, const uint8_t* left, const uint8_t* right, int w, int h_, size_t stride, float mb,
                        float mbf, DropKp* kps, uint8_t* desc, float* u_right, float* depth, int cap) {
  try {
    ORBextractor *el = (ORBextractor*)hl, *er = (ORBextractor*)hr;
    el->SetPyramidDownload(false);
    er->SetPyramidDownload(false);
    std::vector<cv::KeyPoint> kl, kr;
    cv::Mat dl, dr;
    (*el)(cv::Mat(h_, w, CV_8U, (void*)left, stride), cv::Mat(), kl, dl);
    (*er)(cv::Mat(h_, w, CV_8U, (void*)right, stride), cv::Mat(), kr, dr);
    std::vector<float> ur, z;
    ORBextractor::ComputeStereoMatches(*el, *er, kl, dl, kr, dr, mb, mbf, ur, z);
    const int n = (int)kl.size();
    if (n > cap) return -n;
    for (int i = 0; i < n; ++i) {
      kps[i] = DropKp{kl[i].pt.x, kl[i].pt.y, kl[i].size, kl[i].angle, kl[i].response, kl[i].octave, kl[i].class_id};
      memcpy(desc + 32 * (size_t)i, dl.ptr(i), 32);
      u_right[i] = ur[i];
      depth[i] = z[i];
    }
    return n;
  } catch (const std::exception& ex) { g_err = ex.what(); return -1000001; }
}

// ORBmatcher::SearchByBoW(vector<KeyFrame*>, Frame&, ...) -- the product's batched extension: K candidate keyframes (arrays
// of K pointers) against one frame.  out[k] as in orbm_search_bow (kf_kf = 0); a NULL d1[k] stands for a bad keyframe.
int dropin_search_bow_batch(int K, const uint8_t* const* d1, const int* n1, const uint32_t* const* node1, const int32_t* const* off1,
                            const uint32_t* const* idx1, const int* nn1, const uint8_t* const* state1, const float* const* a1,
                            const uint8_t* d2, int n2, const uint32_t* node2, const int32_t* off2, const uint32_t* idx2, int nn2,
                            const float* a2, float ratio, int check_ori, int32_t* const* out, int32_t* nm) {
  try {
    std::vector<std::unique_ptr<MapPoint> > pool;
    std::vector<std::unique_ptr<KeyFrame> > kfs;
    std::vector<KeyFrame*> ptrs;
    auto fill = [](DBoW2::FeatureVector& fv, const uint32_t* node, const int32_t* off, const uint32_t* idx, int n_nodes) {
      for (int i = 0; i < n_nodes; ++i) fv[node[i]] = std::vector<unsigned int>(idx + off[i], idx + off[i + 1]);
    };
    for (int k = 0; k < K; ++k) {
      kfs.emplace_back(new KeyFrame());
      KeyFrame& kf = *kfs.back();
      if (!d1[k]) { kf.mbBad = true; ptrs.push_back(&kf); continue; }
      kf.mDescriptors = cv::Mat(n1[k], 32, CV_8U, (void*)d1[k], 32);
      kf.mvKeysUn.resize((size_t)n1[k]);
      for (int i = 0; i < n1[k]; ++i) kf.mvKeysUn[i].angle = a1[k] ? a1[k][i] : 0.f;
      kf.mvpMapPoints.assign((size_t)n1[k], static_cast<MapPoint*>(NULL));
      for (int i = 0; i < n1[k]; ++i)
        if (state1[k][i]) { pool.emplace_back(new MapPoint(state1[k][i] == 2)); kf.mvpMapPoints[i] = pool.back().get(); }
      fill(kf.mFeatVec, node1[k], off1[k], idx1[k], nn1[k]);
      ptrs.push_back(&kf);
    }
    Frame F;
    F.N = n2;
    F.mDescriptors = cv::Mat(n2, 32, CV_8U, (void*)d2, 32);
    F.mvKeys.resize((size_t)n2);
    for (int i = 0; i < n2; ++i) F.mvKeys[i].angle = a2 ? a2[i] : 0.f;
    fill(F.mFeatVec, node2, off2, idx2, nn2);
    ORBmatcher matcher(ratio, check_ori != 0);
    std::vector<std::vector<MapPoint*> > res;
    const std::vector<int> n = matcher.SearchByBoW(ptrs, F, res);
    for (int k = 0; k < K; ++k) {
      nm[k] = n[k];
      for (int f = 0; f < n2; ++f) {
        out[k][f] = -1;
        if (!res[k].empty() && res[k][f])
          for (int i = 0; i < n1[k]; ++i)
            if (kfs[k]->mvpMapPoints[i] == res[k][f]) out[k][f] = i;
      }
    }
    return 0;
  } catch (const std::exception& ex) { g_err = ex.what(); return -1000001; }
}

int dropin_stereo(const DropKp* kl, const uint8_t* dl, int nl, const DropKp* kr, const uint8_t* dr, int nr, const float* scale,
                  int n_levels, int n_rows, float min_d, float max_d, int32_t* best_idx, int32_t* best_dist) {
  try {
    std::vector<cv::KeyPoint> L((size_t)nl), R((size_t)nr);
    memcpy((void*)L.data(), kl, sizeof(DropKp) * (size_t)nl);
    memcpy((void*)R.data(), kr, sizeof(DropKp) * (size_t)nr);
    std::vector<int> bi, bd;
    ORBmatcher::StereoSearch(L, cv::Mat(nl, 32, CV_8U, (void*)dl, 32), R, cv::Mat(nr, 32, CV_8U, (void*)dr, 32),
                             std::vector<float>(scale, scale + n_levels), n_rows, min_d, max_d, bi, bd);
    for (int i = 0; i < nl; ++i) { best_idx[i] = bi[i]; best_dist[i] = bd[i]; }
    return 0;
  } catch (const std::exception& ex) { g_err = ex.what(); return -1000001; }
}

// the rest of a Frame constructor (Frame.cc:85-120) the drop-in way: bounds once, then one device call
int dropin_finish_frame(const DropKp* keys, int n, const float* cam, const float* dist, int n_dist, float bf, int w, int h_,
                        const float* depth, float* bounds, float* xy_un, float* u_right, float* depth_out, int* cell_off, int* cell_idx) {
  try {
    cv::Mat K = cv::Mat::zeros(3, 3, CV_32F);
    K.at<float>(0, 0) = cam[0]; K.at<float>(1, 1) = cam[1]; K.at<float>(0, 2) = cam[2]; K.at<float>(1, 2) = cam[3]; K.at<float>(2, 2) = 1.f;
    cv::Mat D(n_dist, 1, CV_32F);
    for (int i = 0; i < n_dist; ++i) D.at<float>(i) = dist[i];
    float minX, maxX, minY, maxY;
    ORBextractor::ComputeImageBounds(K, D, w, h_, minX, maxX, minY, maxY);
    const float wInv = static_cast<float>(64) / (maxX - minX), hInv = static_cast<float>(48) / (maxY - minY);
    bounds[0] = minX; bounds[1] = maxX; bounds[2] = minY; bounds[3] = maxY; bounds[4] = wInv; bounds[5] = hInv;
    std::vector<cv::KeyPoint> keysv((size_t)n), un;
    memcpy((void*)keysv.data(), keys, sizeof(DropKp) * (size_t)n);
    std::vector<float> ur, dz;
    std::vector<std::size_t> grid[64][48];
    cv::Mat imDepth;
    if (depth) imDepth = cv::Mat(h_, w, CV_32F, (void*)depth, (size_t)w * 4);
    ORBextractor::FinishFrame(keysv, K, D, bf, minX, minY, wInv, hInv, imDepth, un, ur, dz, grid);
    int total = 0;
    for (int ix = 0; ix < 64; ++ix)
      for (int iy = 0; iy < 48; ++iy) {
        cell_off[ix * 48 + iy] = total;
        for (std::size_t v : grid[ix][iy]) cell_idx[total++] = (int)v;
      }
    cell_off[64 * 48] = total;
    for (int i = 0; i < n; ++i) {
      xy_un[2 * i] = un[i].pt.x; xy_un[2 * i + 1] = un[i].pt.y;
      if (un[i].octave != keysv[i].octave || un[i].angle != keysv[i].angle) { g_err = "mvKeysUn lost keypoint fields"; return -1000000; }
      if (depth) { u_right[i] = ur[i]; depth_out[i] = dz[i]; }
    }
    return total;
  } catch (const std::exception& ex) { g_err = ex.what(); return -1000001; }
}

}  // extern "C"

// ---- matcher scenes shared with the reference build (oracle/ref_match_harness.cc)
float ORB_SLAM2::Frame::mnMinX, ORB_SLAM2::Frame::mnMinY, ORB_SLAM2::Frame::mnMaxX, ORB_SLAM2::Frame::mnMaxY;
float ORB_SLAM2::Frame::mfGridElementWidthInv, ORB_SLAM2::Frame::mfGridElementHeightInv;
#define SCN(name) dropin_##name
#include "matcher_scenes.inc"
