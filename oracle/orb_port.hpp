// oracle/orb_port.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement ("port") of the reference ORBextractor on flat arrays, written from
// the behaviour of /root/reference/src/ORBextractor.cc; each stage cites the lines it
// follows.  It is validated (tests/test_oracle_vs_ref.py, in the build container) against
// oracle/_ref: the reference's own ORBextractor.cc compiled unmodified over oracle/cvshim.
// Parity status: the reference ships no tests or golden vectors for this path,
// so parity is pinned to (a) cv2 4.13.0 for the OpenCV primitives and (b) the reference
// source itself run here; see DESIGN.md "Oracle".
//
// Quadtree tie-break convention: the reference sorts (size, node pointer) pairs
// (src/ORBextractor.cc:684), so equal-size nodes are ordered by heap address.  Here, and in
// oracle/_ref (monotonic operator new), "address order" == creation order.
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

#include "cvprims.hpp"

namespace orbport {

static const int PATCH_SIZE = 31, HALF_PATCH = 15, EDGE = 19;  // ORBextractor.cc:72-74

static const int kPattern[1024] = {
#include "../orb_slam2_e_b200/csrc/orb_pattern.inc"
};

struct KeyPoint {  // same field order / size as cv::KeyPoint (28 bytes)
  float x, y, size, angle, response;
  int octave, class_id;
};

struct Candidate {  // FAST keypoint in border-relative coordinates (ORBextractor.cc:830-831)
  int x, y, score;
};

struct Level {
  int w = 0, h = 0;
  size_t step = 0;           // of the bordered buffer
  std::vector<uint8_t> buf;  // (h+38) x (w+38)
  uint8_t* roi() { return buf.data() + (size_t)EDGE * step + EDGE; }
  const uint8_t* roi() const { return buf.data() + (size_t)EDGE * step + EDGE; }
};

class Extractor {
 public:
  int nfeatures, nlevels, iniTh, minTh;
  double scaleFactor;  // stored as double in the reference (ORBextractor.h:99)
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> quota;  // mnFeaturesPerLevel
  int umax[HALF_PATCH + 1];
  std::vector<Level> pyr;

  // ORBextractor.cc:410-470
  Extractor(int nf, float sf, int nl, int ini, int mn)
      : nfeatures(nf), nlevels(nl), iniTh(ini), minTh(mn), scaleFactor(sf) {
    scale.resize(nl); sigma2.resize(nl); invScale.resize(nl); invSigma2.resize(nl);
    scale[0] = 1.f; sigma2[0] = 1.f;
    for (int i = 1; i < nl; ++i) {
      scale[i] = (float)(scale[i - 1] * scaleFactor);  // float*double -> double -> float
      sigma2[i] = scale[i] * scale[i];
    }
    for (int i = 0; i < nl; ++i) { invScale[i] = 1.f / scale[i]; invSigma2[i] = 1.f / sigma2[i]; }
    quota.resize(nl);
    float factor = (float)(1.0f / scaleFactor);
    float want = nf * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
      quota[l] = cvprims::cv_round(want);
      sum += quota[l];
      want *= factor;
    }
    quota[nl - 1] = std::max(nf - sum, 0);
    // circular patch row extents
    int vmax = cvprims::cv_floor(HALF_PATCH * sqrt(2.f) / 2 + 1);
    int vmin = cvprims::cv_ceil(HALF_PATCH * sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH * HALF_PATCH;
    for (int v = 0; v <= vmax; ++v) umax[v] = cvprims::cv_round(sqrt(hp2 - v * v));
    for (int v = HALF_PATCH, v0 = 0; v >= vmin; --v) {
      while (umax[v0] == umax[v0 + 1]) ++v0;
      umax[v] = v0;
      ++v0;
    }
    pyr.resize(nl);
  }

  // ORBextractor.cc:1115-1140: level l is resized from level l-1, each stored inside a
  // 19-px reflect-101 frame.
  void compute_pyramid(const uint8_t* img, int w, int h, size_t stride) {
    for (int l = 0; l < nlevels; ++l) {
      Level& L = pyr[l];
      float s = invScale[l];
      L.w = cvprims::cv_round((float)w * s);
      L.h = cvprims::cv_round((float)h * s);
      L.step = (size_t)L.w + 2 * EDGE;
      L.buf.assign(L.step * (size_t)(L.h + 2 * EDGE), 0);
      if (l == 0) {
        cvprims::make_border_reflect101(img, w, h, stride, L.buf.data(), L.step, EDGE);
      } else {
        const Level& P = pyr[l - 1];
        cvprims::resize_linear_u8(P.roi(), P.w, P.h, P.step, L.roi(), L.w, L.h, L.step);
        cvprims::make_border_reflect101(L.roi(), L.w, L.h, L.step, L.buf.data(), L.step, EDGE);
      }
    }
  }

  struct CellGrid {
    int minB, maxBX, maxBY, nCols, nRows, wCell, hCell;
  };
  // ORBextractor.cc:771-787
  CellGrid cell_grid(int l) const {
    CellGrid g;
    g.minB = EDGE - 3;
    g.maxBX = pyr[l].w - EDGE + 3;
    g.maxBY = pyr[l].h - EDGE + 3;
    const float width = (float)(g.maxBX - g.minB), height = (float)(g.maxBY - g.minB);
    g.nCols = (int)(width / 30.f);
    g.nRows = (int)(height / 30.f);
    g.wCell = (int)ceil(width / g.nCols);
    g.hCell = (int)ceil(height / g.nRows);
    return g;
  }

  // ORBextractor.cc:789-837: per-cell FAST at iniTh, falling back to minTh when the cell
  // yields nothing.  Output order: cell row, cell column, then row-major inside the cell.
  void detect_level(int l, std::vector<Candidate>& out) const {
    out.clear();
    const Level& L = pyr[l];
    const CellGrid g = cell_grid(l);
    std::vector<cvprims::FastKp> cell;
    for (int i = 0; i < g.nRows; ++i) {
      const float iniY = (float)(g.minB + i * g.hCell);
      float maxY = iniY + g.hCell + 6;
      if (iniY >= g.maxBY - 3) continue;
      if (maxY > g.maxBY) maxY = (float)g.maxBY;
      for (int j = 0; j < g.nCols; ++j) {
        const float iniX = (float)(g.minB + j * g.wCell);
        float maxX = iniX + g.wCell + 6;
        if (iniX >= g.maxBX - 6) continue;
        if (maxX > g.maxBX) maxX = (float)g.maxBX;
        const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
        const uint8_t* view = L.roi() + (ptrdiff_t)y0 * (ptrdiff_t)L.step + x0;
        cvprims::fast9_16(view, cw, ch, L.step, iniTh, cell);
        if (cell.empty()) cvprims::fast9_16(view, cw, ch, L.step, minTh, cell);
        for (const auto& k : cell) out.push_back({k.x + j * g.wCell, k.y + i * g.hCell, k.score});
      }
    }
  }

  void distribute(const std::vector<Candidate>& keys, int boxW, int boxH, int N,
                  std::vector<Candidate>& kept) const;
  float ic_angle(int l, int x, int y) const;
  void describe(const uint8_t* blurred, size_t step, int x, int y, float angle_deg,
                uint8_t* desc) const;
  void extract(const uint8_t* img, int w, int h, size_t stride, std::vector<KeyPoint>& kps,
               std::vector<uint8_t>& desc);
  // stage outputs of the last extract() call, for stage-level parity tests
  std::vector<std::vector<Candidate>> lastCandidates, lastKept;
};

}  // namespace orbport

#include "orb_port_impl.hpp"
