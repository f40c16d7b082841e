// oracle/cvprims.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the five un-vendored OpenCV primitives that the reference
// ORBextractor calls (reference: src/ORBextractor.cc:103 fastAtan2, :810/:822 FAST,
// :1094 GaussianBlur, :1128 resize, :1130-1136 copyMakeBorder).  OpenCV itself is
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

}  // namespace cvprims
