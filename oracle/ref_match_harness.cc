// oracle/ref_match_harness.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C entry points (refm_*) over the reference's OWN matcher: /root/reference/src/ORBmatcher.cc is
// compiled UNMODIFIED next to this file, against the reference's own include/ORBmatcher.h (reached
// through a build-time symlink in oracle/_ref/env so that its `#include "MapPoint.h"` etc. resolve
// to tests/cpp/mock instead of the reference's Eigen/g2o-dependent headers) and oracle/cvshim.
// The scenes are tests/cpp/matcher_scenes.inc, the same file the product's drop-in harness builds,
// so both sides see identical objects.  Output: oracle/_ref/libmatchref.so (git-ignored).
#include "ORBmatcher.h"

float ORB_SLAM2::Frame::mnMinX, ORB_SLAM2::Frame::mnMinY, ORB_SLAM2::Frame::mnMaxX, ORB_SLAM2::Frame::mnMaxY;
float ORB_SLAM2::Frame::mfGridElementWidthInv, ORB_SLAM2::Frame::mfGridElementHeightInv;
#define SCN(name) refm_##name
#include "../tests/cpp/matcher_scenes.inc"

// Timed CPU baseline of bench.py's matching leg (kind "reference"): the brute-force relocalization scan written the way the
// reference calls its own distance function -- ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&) on one-row
// cv::Mat views (ORBmatcher.cc:413-415, :1848-1864) with the update rule of :417-426 -- rows sharded over host threads.
#include <thread>
#include <vector>
extern "C" void refm_top2_scan_mt(const uint8_t* q, int nq, const uint8_t* db, long long nd, int* d1, long long* i1, int* d2,
                                  int n_threads) {
  auto work = [&](int a, int b) {
    for (int qi = a; qi < b; ++qi) {
      const cv::Mat dQ(1, 32, CV_8U, (void*)(q + 32 * (size_t)qi), 32);
      int best1 = 256, best2 = 256;
      long long idx = -1;
      for (long long j = 0; j < nd; ++j) {
        const cv::Mat dR(1, 32, CV_8U, (void*)(db + 32 * (size_t)j), 32);
        const int dist = ORB_SLAM2::ORBmatcher::DescriptorDistance(dQ, dR);
        if (dist < best1) { best2 = best1; best1 = dist; idx = j; }
        else if (dist < best2) best2 = dist;
      }
      d1[qi] = best1; d2[qi] = best2; i1[qi] = idx;
    }
  };
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t)
    th.emplace_back(work, (int)((long long)nq * t / n_threads), (int)((long long)nq * (t + 1) / n_threads));
  for (auto& x : th) x.join();
}
