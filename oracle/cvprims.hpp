// oracle/cvprims.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the five un-vendored OpenCV primitives that the reference
// ORBextractor calls (reference: src/ORBextractor.cc:103 fastAtan2, :810/:822 FAST,
// :1094 GaussianBlur, :1128 resize, :1130-1136 copyMakeBorder), plus cv::undistortPoints
// as Frame::UndistortKeyPoints / ComputeImageBounds call it (src/Frame.cc:438, :466).  OpenCV itself is
// absent from /root/reference (CMakeLists.txt:20 `find_package(OpenCV 3.4 REQUIRED)`),
// so the arithmetic below restates the published OpenCV 4.x algorithms for CV_8U and
// is pinned against the `cv2` 4.13.0 wheel by tests/test_oracle_prims.py.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may use this file.  All functions work on raw row-major u8 buffers with an
// explicit byte stride.
#pragma once
#include <cfloat>
#include <cstddef>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace cvprims {

// cvRound: round-half-to-even (lrint under the default FP environment).
static inline int cv_round(double v) { return (int)lrint(v); }
static inline int cv_round(float v) { return (int)lrintf(v); }
static inline int cv_floor(double v) {
  int i = (int)v;
  return i - (i > v);
}
static inline int cv_ceil(double v) {
  int i = (int)v;
  return i + (i < v);
}

// ---------------------------------------------------------------------------------
// BORDER_REFLECT_101 index (gfedcb|abcdefgh|gfedcba).
static inline int reflect101(int p, int n) {
  if (n == 1) return 0;
  while (p < 0 || p >= n) p = p < 0 ? -p : 2 * n - 2 - p;
  return p;
}

// copyMakeBorder(src, dst, b, b, b, b, BORDER_REFLECT_101 | BORDER_ISOLATED)
// dst is (h+2b) x (w+2b).  src may alias the interior of dst.
static inline void make_border_reflect101(const uint8_t* src, int w, int h, size_t sstep,
                                          uint8_t* dst, size_t dstep, int b) {
  // interior first (memmove tolerates src == interior of dst)
  for (int y = 0; y < h; ++y) {
    uint8_t* drow = dst + (size_t)(y + b) * dstep + b;
    const uint8_t* srow = src + (size_t)y * sstep;
    if (drow != srow) memmove(drow, srow, (size_t)w);
  }
  for (int y = 0; y < h; ++y) {
    uint8_t* drow = dst + (size_t)(y + b) * dstep;
    for (int x = 0; x < b; ++x) {
      drow[x] = drow[b + reflect101(x - b, w)];
      drow[b + w + x] = drow[b + reflect101(w + x, w)];
    }
  }
  for (int y = 0; y < b; ++y) {
    memcpy(dst + (size_t)y * dstep, dst + (size_t)(b + reflect101(y - b, h)) * dstep,
           (size_t)(w + 2 * b));
    memcpy(dst + (size_t)(b + h + y) * dstep,
           dst + (size_t)(b + reflect101(h + y, h)) * dstep, (size_t)(w + 2 * b));
  }
}

// ---------------------------------------------------------------------------------
// resize(src, dst, INTER_LINEAR) for CV_8UC1: 11-bit fixed-point coefficients.
struct LinearAxis {
  std::vector<int> ofs;       // source index of the left/top tap
  std::vector<short> c0, c1;  // 11-bit weights of tap ofs and ofs+1
};

static inline LinearAxis linear_axis(int src_n, int dst_n, bool clamp_frac) {
  LinearAxis ax;
  ax.ofs.resize(dst_n);
  ax.c0.resize(dst_n);
  ax.c1.resize(dst_n);
  const double scale = (double)src_n / dst_n;
  for (int d = 0; d < dst_n; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = cv_floor(f);
    f -= s;
    if (clamp_frac) {  // x axis: taps and fraction clamped together
      if (s < 0) { f = 0; s = 0; }
      if (s >= src_n - 1) { f = 0; s = src_n - 1; }
    }
    ax.ofs[d] = s;
    ax.c0[d] = (short)cv_round((1.f - f) * 2048.f);
    ax.c1[d] = (short)cv_round(f * 2048.f);
  }
  return ax;
}

static inline void resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstep,
                                    uint8_t* dst, int dw, int dh, size_t dstep) {
  LinearAxis ax = linear_axis(sw, dw, true);
  LinearAxis ay = linear_axis(sh, dh, false);  // y axis: rows clamped, fraction kept
  std::vector<int> r0(dw), r1(dw);
  auto hrow = [&](int sy, std::vector<int>& out) {
    sy = sy < 0 ? 0 : (sy >= sh ? sh - 1 : sy);
    const uint8_t* s = src + (size_t)sy * sstep;
    for (int d = 0; d < dw; ++d) {
      int o = ax.ofs[d];
      int o1 = o + 1 < sw ? o + 1 : sw - 1;
      out[d] = s[o] * ax.c0[d] + s[o1] * ax.c1[d];
    }
  };
  for (int y = 0; y < dh; ++y) {
    hrow(ay.ofs[y], r0);
    hrow(ay.ofs[y] + 1, r1);
    const int b0 = ay.c0[y], b1 = ay.c1[y];
    uint8_t* d = dst + (size_t)y * dstep;
    for (int x = 0; x < dw; ++x)
      d[x] = (uint8_t)((((b0 * (r0[x] >> 4)) >> 16) + ((b1 * (r1[x] >> 4)) >> 16) + 2) >> 2);
  }
}

// ---------------------------------------------------------------------------------
// GaussianBlur(7x7, sigma=2, BORDER_REFLECT_101) for CV_8UC1: OpenCV >= 3.4's
// fixed-point path.  The 8.8 kernel is {18,34,48,56,48,34,18}/256.
static const int kGauss7[7] = {18, 34, 48, 56, 48, 34, 18};

static inline void gaussian_blur7_u8(const uint8_t* src, int w, int h, size_t sstep,
                                     uint8_t* dst, size_t dstep) {
  std::vector<uint16_t> tmp((size_t)w * h);
  for (int y = 0; y < h; ++y) {
    const uint8_t* s = src + (size_t)y * sstep;
    for (int x = 0; x < w; ++x) {
      unsigned acc = 0;
      for (int k = 0; k < 7; ++k) acc += kGauss7[k] * s[reflect101(x + k - 3, w)];
      tmp[(size_t)y * w + x] = (uint16_t)acc;  // <= 255*256
    }
  }
  for (int y = 0; y < h; ++y) {
    uint8_t* d = dst + (size_t)y * dstep;
    const uint16_t* rows[7];
    for (int k = 0; k < 7; ++k) rows[k] = &tmp[(size_t)reflect101(y + k - 3, h) * w];
    for (int x = 0; x < w; ++x) {
      unsigned acc = 0;
      for (int k = 0; k < 7; ++k) acc += kGauss7[k] * rows[k][x];
      d[x] = (uint8_t)((acc + 32768u) >> 16);
    }
  }
}

// ---------------------------------------------------------------------------------
// fastAtan2(y, x) in degrees [0,360): OpenCV's degree-7 odd polynomial, every
// operation individually rounded to float (no FMA: build with -ffp-contract=off).
static inline float fast_atan2_deg(float y, float x) {
  const float scale = (float)(180.0 / 3.14159265358979323846);
  const float p1 = 0.9997878412794807f * scale;
  const float p3 = -0.3258083974640975f * scale;
  const float p5 = 0.1555786518463281f * scale;
  const float p7 = -0.04432655554792128f * scale;
  const float eps = (float)DBL_EPSILON;
  float ax = std::fabs(x), ay = std::fabs(y);
  float a, c, c2;
  if (ax >= ay) {
    c = ay / (ax + eps);
    c2 = c * c;
    a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  } else {
    c = ax / (ay + eps);
    c2 = c * c;
    a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
  }
  if (x < 0) a = 180.f - a;
  if (y < 0) a = 360.f - a;
  return a;
}

// ---------------------------------------------------------------------------------
// FAST-9/16 (cv::FAST(img, kps, threshold, true), TYPE_9_16).
struct FastKp {
  int x, y, score;
};

static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// Threshold-independent corner strength: max over the 16 nine-long arcs of
// min(v - ring) and of min(ring - v), minus one.  A pixel is a FAST-9 corner at
// threshold T iff strength >= T, and OpenCV's cornerScore<16> equals this value for
// every corner.
static inline int fast_strength(const uint8_t* p, size_t step) {
  int d[16];
  const int v = p[0];
  for (int k = 0; k < 16; ++k) d[k] = v - p[(ptrdiff_t)kRingDy[k] * (ptrdiff_t)step + kRingDx[k]];
  int best = -256;
  for (int s = 0; s < 16; ++s) {
    int lo = 256, hi = -256;
    for (int k = 0; k < 9; ++k) {
      int q = d[(s + k) & 15];
      lo = q < lo ? q : lo;
      hi = q > hi ? q : hi;
    }
    if (lo > best) best = lo;     // arc darker than the centre by at least lo
    if (-hi > best) best = -hi;   // arc brighter than the centre by at least -hi
  }
  return best - 1;
}

// OpenCV-order detection on a w x h view: row-major scan, 3x3 strict non-max
// suppression where anything outside [3,w-3)x[3,h-3) or below threshold scores 0.
static inline void fast9_16(const uint8_t* img, int w, int h, size_t step, int threshold,
                            std::vector<FastKp>& out) {
  out.clear();
  if (w < 7 || h < 7) return;
  std::vector<uint8_t> sc((size_t)w * h, 0);
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) {
      // cheap necessary condition (every nine-long arc holds one pixel of each opposite
      // pair), then the exact strength; identical result, ~10x less work on the CPU
      const uint8_t* p = img + (size_t)y * step + x;
      const int v = p[0];
      bool dark = true, bright = true;
      for (int k = 0; k < 8 && (dark || bright); k += 2) {
        const int a = v - p[(ptrdiff_t)kRingDy[k] * (ptrdiff_t)step + kRingDx[k]];
        const int b = v - p[(ptrdiff_t)kRingDy[k + 8] * (ptrdiff_t)step + kRingDx[k + 8]];
        dark = dark && (a > threshold || b > threshold);
        bright = bright && (a < -threshold || b < -threshold);
      }
      if (!dark && !bright) continue;
      int s = fast_strength(p, step);
      if (s >= threshold) sc[(size_t)y * w + x] = (uint8_t)s;
    }
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) {
      int s = sc[(size_t)y * w + x];
      if (!s) continue;
      const uint8_t* r = &sc[(size_t)y * w + x];
      if (s > r[-1] && s > r[1] && s > r[-w - 1] && s > r[-w] && s > r[-w + 1] &&
          s > r[w - 1] && s > r[w] && s > r[w + 1])
        out.push_back({x, y, s});
    }
}

// ---------------------------------------------------------------------------
// cv::undistortPoints(src, dst, K, distCoeffs, R = Mat(), P) for CV_32FC2 points, as OpenCV 4.x
// computes it when no TermCriteria is passed (calib3d undistortPoints -> cvUndistortPointsInternal
// with TermCriteria(MAX_ITER, 5, 0.01): exactly five iterations, no epsilon test).  Everything is
// double; K and the coefficients arrive as CV_32F and are widened; the 14-entry coefficient array
// (k1 k2 p1 p2 k3 k4 k5 k6 s1 s2 s3 s4 tauX tauY) is zero beyond the n_dist given, the tilt matrix is
// the identity for tau = 0 and RR = P * I.  Reference call sites: src/Frame.cc:438 and :466
// (R empty, P = mK).  Pinned against cv2 4.13.0 by tests/test_frame_finish_cpu.py.
static inline void undistort_points(const float* src_xy, int n, const float* K9, const float* dist, int n_dist,
                                    const float* P9, float* dst_xy) {
  double A[3][3], RR[3][3], k[14] = {0};
  for (int i = 0; i < 9; ++i) A[i / 3][i % 3] = (double)K9[i];
  for (int i = 0; i < n_dist && i < 14; ++i) k[i] = (double)dist[i];
  for (int i = 0; i < 3; ++i)       // RR = PP * I  (cvMatMul)
    for (int j = 0; j < 3; ++j) {
      double acc = 0;
      for (int t = 0; t < 3; ++t) acc += (double)P9[i * 3 + t] * (t == j ? 1.0 : 0.0);
      RR[i][j] = acc;
    }
  const double fx = A[0][0], fy = A[1][1], ifx = 1. / fx, ify = 1. / fy, cx = A[0][2], cy = A[1][2];
  const double invMatTilt[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int i = 0; i < n; ++i) {
    double x = src_xy[2 * i], y = src_xy[2 * i + 1], x0, y0, u = x, v = y;
    x = (x - cx) * ifx;
    y = (y - cy) * ify;
    if (dist && n_dist > 0) {
      const double vec[3] = {x, y, 1};
      double ut[3];
      for (int r = 0; r < 3; ++r) ut[r] = invMatTilt[r][0] * vec[0] + invMatTilt[r][1] * vec[1] + invMatTilt[r][2] * vec[2];
      const double invProj = ut[2] ? 1. / ut[2] : 1;
      x0 = x = invProj * ut[0];
      y0 = y = invProj * ut[1];
      for (int j = 0; j < 5; ++j) {
        const double r2 = x * x + y * y;
        const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
        if (icdist < 0) {
          x = (u - cx) * ifx;
          y = (v - cy) * ify;
          break;
        }
        const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
        const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
        x = (x0 - deltaX) * icdist;
        y = (y0 - deltaY) * icdist;
      }
    }
    const double xx = RR[0][0] * x + RR[0][1] * y + RR[0][2];
    const double yy = RR[1][0] * x + RR[1][1] * y + RR[1][2];
    const double ww = 1. / (RR[2][0] * x + RR[2][1] * y + RR[2][2]);
    x = xx * ww;
    y = yy * ww;
    dst_xy[2 * i] = (float)x;
    dst_xy[2 * i + 1] = (float)y;
  }
}

}  // namespace cvprims
