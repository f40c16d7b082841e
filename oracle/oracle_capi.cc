// oracle/oracle_capi.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// C ABI over the CPU restatements (cvprims.hpp, orb_port.hpp, match_port.hpp) so that
// tests/, smoke() and bench.py's cpu_baseline leg can drive them through ctypes.
// Build: oracle/Makefile -> oracle/_build/liborboracle.so (-ffp-contract=off).
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "match_port.hpp"
#include "orb_port.hpp"

using orbport::Extractor;

extern "C" {

// ---- primitives -----------------------------------------------------------------
void orc_resize_linear(const uint8_t* s, int sw, int sh, size_t sstep, uint8_t* d, int dw, int dh, size_t dstep) {
  cvprims::resize_linear_u8(s, sw, sh, sstep, d, dw, dh, dstep);
}
void orc_border101(const uint8_t* s, int w, int h, size_t sstep, uint8_t* d, size_t dstep, int b) {
  cvprims::make_border_reflect101(s, w, h, sstep, d, dstep, b);
}
void orc_blur7(const uint8_t* s, int w, int h, size_t sstep, uint8_t* d, size_t dstep) {
  cvprims::gaussian_blur7_u8(s, w, h, sstep, d, dstep);
}
float orc_fast_atan2(float y, float x) { return cvprims::fast_atan2_deg(y, x); }
void orc_fast_atan2_v(const float* y, const float* x, float* out, int n) {
  for (int i = 0; i < n; ++i) out[i] = cvprims::fast_atan2_deg(y[i], x[i]);
}
// FAST on a view; out triples (x,y,score); returns count (clipped to cap).
int orc_fast(const uint8_t* img, int w, int h, size_t step, int th, int* out, int cap) {
  std::vector<cvprims::FastKp> k;
  cvprims::fast9_16(img, w, h, step, th, k);
  int n = (int)k.size() < cap ? (int)k.size() : cap;
  for (int i = 0; i < n; ++i) { out[3 * i] = k[i].x; out[3 * i + 1] = k[i].y; out[3 * i + 2] = k[i].score; }
  return (int)k.size();
}
void orc_fast_strength_map(const uint8_t* img, int w, int h, size_t step, uint8_t* out) {
  memset(out, 0, (size_t)w * h);
  for (int y = 3; y < h - 3; ++y)
    for (int x = 3; x < w - 3; ++x) {
      int s = cvprims::fast_strength(img + (size_t)y * step + x, step);
      out[(size_t)y * w + x] = (uint8_t)(s < 0 ? 0 : s);
    }
}

// ---- extractor ------------------------------------------------------------------
void* orc_create(int nf, float sf, int nl, int ini, int mn) { return new Extractor(nf, sf, nl, ini, mn); }
void orc_destroy(void* h) { delete (Extractor*)h; }
int orc_quota(void* h, int l) { return ((Extractor*)h)->quota[l]; }
int orc_umax(void* h, int v) { return ((Extractor*)h)->umax[v]; }
float orc_scale(void* h, int l) { return ((Extractor*)h)->scale[l]; }

int orc_extract(void* h, const uint8_t* img, int w, int ht, size_t stride, orbport::KeyPoint* kps, uint8_t* desc, int cap) {
  Extractor* e = (Extractor*)h;
  std::vector<orbport::KeyPoint> k;
  std::vector<uint8_t> d;
  e->extract(img, w, ht, stride, k, d);
  if ((int)k.size() > cap) return -(int)k.size();
  if (!k.empty()) {
    memcpy(kps, k.data(), k.size() * sizeof(orbport::KeyPoint));
    memcpy(desc, d.data(), d.size());
  }
  return (int)k.size();
}
int orc_level_size(void* h, int l, int* w, int* ht) {
  Extractor* e = (Extractor*)h;
  if (l < 0 || l >= e->nlevels || e->pyr[l].buf.empty()) return -1;
  *w = e->pyr[l].w; *ht = e->pyr[l].h;
  return 0;
}
int orc_get_level(void* h, int l, uint8_t* dst, size_t dstep, int with_border) {
  Extractor* e = (Extractor*)h;
  if (l < 0 || l >= e->nlevels || e->pyr[l].buf.empty()) return -1;
  const orbport::Level& L = e->pyr[l];
  const int b = with_border ? orbport::EDGE : 0;
  for (int y = -b; y < L.h + b; ++y)
    memcpy(dst + (size_t)(y + b) * dstep, L.roi() + (ptrdiff_t)y * (ptrdiff_t)L.step - b, (size_t)(L.w + 2 * b));
  return 0;
}
// stage outputs of the last orc_extract: which = 0 candidates, 1 kept; triples (x,y,score)
// in border-relative coordinates, in the reference's order.
int orc_stage(void* h, int which, int l, int* out, int cap) {
  Extractor* e = (Extractor*)h;
  const auto& v = which ? e->lastKept : e->lastCandidates;
  if (l < 0 || l >= (int)v.size()) return -1;
  int n = (int)v[l].size();
  for (int i = 0; i < n && i < cap; ++i) { out[3 * i] = v[l][i].x; out[3 * i + 1] = v[l][i].y; out[3 * i + 2] = v[l][i].score; }
  return n;
}
// Quadtree alone on caller-supplied candidates (x,y,score triples).
int orc_distribute(void* h, const int* cand, int n, int box_w, int box_h, int N, int* out, int cap) {
  std::vector<orbport::Candidate> in((size_t)n), kept;
  for (int i = 0; i < n; ++i) in[i] = {cand[3 * i], cand[3 * i + 1], cand[3 * i + 2]};
  ((Extractor*)h)->distribute(in, box_w, box_h, N, kept);
  int m = (int)kept.size();
  for (int i = 0; i < m && i < cap; ++i) { out[3 * i] = kept[i].x; out[3 * i + 1] = kept[i].y; out[3 * i + 2] = kept[i].score; }
  return m;
}
// Batch extraction on n_threads host threads, one extractor object per thread (the way the
// reference would be scaled over host cores).  Returns total keypoints.  For cpu_baseline.
long long orc_extract_batch_mt(int nf, float sf, int nl, int ini, int mn, const uint8_t* imgs, int n_frames,
                               int w, int ht, int n_threads, int* counts) {
  std::vector<std::thread> th;
  std::vector<long long> tot((size_t)n_threads, 0);
  for (int t = 0; t < n_threads; ++t)
    th.emplace_back([&, t]() {
      Extractor e(nf, sf, nl, ini, mn);
      std::vector<orbport::KeyPoint> k;
      std::vector<uint8_t> d;
      for (int f = t; f < n_frames; f += n_threads) {
        e.extract(imgs + (size_t)f * w * ht, w, ht, (size_t)w, k, d);
        if (counts) counts[f] = (int)k.size();
        tot[t] += (long long)k.size();
      }
    });
  long long s = 0;
  for (int t = 0; t < n_threads; ++t) { th[t].join(); s += tot[t]; }
  return s;
}

// ---- matcher --------------------------------------------------------------------
int orc_hamming(const uint8_t* a, const uint8_t* b) { return matchport::hamming256(a, b); }
void orc_top2(const uint8_t* q, int nq, const uint8_t* db, long long nd, int* d1, long long* i1, int* d2) {
  matchport::top2_bruteforce(q, nq, db, nd, d1, i1, d2);
}
void orc_top2_mt(const uint8_t* q, int nq, const uint8_t* db, long long nd, int* d1, long long* i1, int* d2, int n_threads) {
  std::vector<std::thread> th;
  for (int t = 0; t < n_threads; ++t)
    th.emplace_back([&, t]() {
      int a = (int)((long long)nq * t / n_threads), b = (int)((long long)nq * (t + 1) / n_threads);
      matchport::top2_bruteforce(q + 32 * (size_t)a, b - a, db, nd, d1 + a, i1 + a, d2 + a);
    });
  for (auto& x : th) x.join();
}
int orc_accept(int b1, int b2, float ratio, int th, int strict) { return matchport::accept(b1, b2, ratio, th, strict != 0); }
void orc_merge_top2(const int* d1, const long long* i1, const int* d2, int n_shards, int nq, int* od1, long long* oi1, int* od2) {
  matchport::merge_top2(d1, i1, d2, n_shards, nq, od1, oi1, od2);
}
int orc_search_bow(const uint8_t* desc1, int n1, const uint8_t* desc2, int n2,
                   const uint32_t* node1, const int* off1, const uint32_t* idx1, int nn1,
                   const uint32_t* node2, const int* off2, const uint32_t* idx2, int nn2,
                   const uint8_t* valid1, const uint8_t* valid2, const float* ang1, const float* ang2,
                   float ratio, int check_ori, int kf_kf, int* out) {
  matchport::FeatVec f1{node1, off1, idx1, nn1}, f2{node2, off2, idx2, nn2};
  return matchport::search_bow(desc1, n1, desc2, n2, f1, f2, valid1, valid2, ang1, ang2, ratio, check_ori, kf_kf != 0, out);
}
void orc_stereo_search(const float* lx, const float* ly, const int* loct, const uint8_t* ldesc, int nl,
                       const float* rx, const float* ry, const int* roct, const uint8_t* rdesc, int nr,
                       const float* scale_factors, int n_rows, float min_d, float max_d, int* best_idx, int* best_dist) {
  matchport::stereo_search(lx, ly, loct, ldesc, nl, rx, ry, roct, rdesc, nr, scale_factors, n_rows, min_d, max_d, best_idx, best_dist);
}
int orc_window_search(const float* x, const float* y, const int* octave, const float* u_right, const uint8_t* desc, int n,
                      float min_x, float min_y, float width_inv, float height_inv, int cols, int rows,
                      const int* cell_off, const uint32_t* cell_idx, const void* queries, const uint8_t* qdesc, int nq,
                      const uint8_t* claimed, int th, float ratio, int claim, void* out, const float* inv_level_sigma2,
                      double chi2_mono, double chi2_stereo) {
  matchport::GridFrame F{x, y, octave, u_right, desc, n, min_x, min_y, width_inv, height_inv, cols, rows, cell_off, cell_idx};
  matchport::WindowGate gate{inv_level_sigma2, chi2_mono, chi2_stereo};
  return matchport::window_search(F, (const matchport::WindowQuery*)queries, qdesc, nq, claimed, th, ratio, claim != 0,
                                  (matchport::WindowResult*)out, inv_level_sigma2 ? &gate : nullptr);
}
// pyramids: n_levels entries each of (pointer, step, w, h) passed as parallel arrays
void orc_compute_stereo_matches(const float* lx, const float* ly, const int* loct, const uint8_t* ldesc, int nl,
                                const float* rx, const float* ry, const int* roct, const uint8_t* rdesc, int nr,
                                const float* scale, const float* inv_scale, int n_levels, const uint8_t* const* pxL,
                                const uint8_t* const* pxR, const int* step, const int* w, const int* h, float mb, float mbf,
                                float* u_right, float* depth) {
  std::vector<matchport::PyrLevel> L((size_t)n_levels), R((size_t)n_levels);
  for (int l = 0; l < n_levels; ++l) {
    L[l] = matchport::PyrLevel{pxL[l], (size_t)step[l], w[l], h[l]};
    R[l] = matchport::PyrLevel{pxR[l], (size_t)step[l], w[l], h[l]};
  }
  matchport::compute_stereo_matches(lx, ly, loct, ldesc, nl, rx, ry, roct, rdesc, nr, scale, inv_scale, L.data(), R.data(), mb, mbf,
                                    u_right, depth);
}
// lists back to back in out_idx, offsets in out_off[nq + 1]; returns the total (negative: cap too small)
int orc_features_in_area(const float* x, const float* y, const int* octave, int n, float min_x, float min_y, float width_inv,
                         float height_inv, int cols, int rows, const int* cell_off, const uint32_t* cell_idx, const float* qx,
                         const float* qy, const float* qr, const int* qmin, const int* qmax, int nq, int* out_off, int* out_idx,
                         int cap) {
  matchport::GridFrame F{x, y, octave, nullptr, nullptr, n, min_x, min_y, width_inv, height_inv, cols, rows, cell_off, cell_idx};
  std::vector<int> v;
  int total = 0;
  for (int q = 0; q < nq; ++q) {
    out_off[q] = total;
    matchport::features_in_area(F, qx[q], qy[q], qr[q], qmin[q], qmax[q], v);
    for (size_t j = 0; j < v.size(); ++j) {
      if (total < cap) out_idx[total] = v[j];
      ++total;
    }
  }
  out_off[nq] = total;
  return total <= cap ? total : -total;
}
// rest of Frame construction (matchport::undistort_keypoints ... stereo_from_rgbd); cell lists as CSR
void orc_undistort_points(const float* xy, int n, const float* cam, const float* dist, int n_dist, float* out) {
  matchport::undistort_keypoints(xy, n, cam, dist, n_dist, out);
}
void orc_image_bounds(int w, int h, const float* cam, const float* dist, int n_dist, float* bounds) {
  matchport::image_bounds(w, h, cam, dist, n_dist, bounds);
}
int orc_frame_finish(const float* xy, int n, const float* cam, const float* dist, int n_dist, float bf, float min_x, float min_y,
                     float width_inv, float height_inv, int cols, int rows, const float* depth, int depth_w, float* xy_un,
                     float* u_right, float* depth_out, int* cell_off, int* cell_idx) {
  matchport::undistort_keypoints(xy, n, cam, dist, n_dist, xy_un);
  if (depth) matchport::stereo_from_rgbd(xy, xy_un, n, depth, depth_w, bf, u_right, depth_out);
  std::vector<std::vector<int>> cells;
  matchport::assign_features_to_grid(xy_un, n, min_x, min_y, width_inv, height_inv, cols, rows, cells);
  int total = 0;
  for (size_t c = 0; c < cells.size(); ++c) {
    cell_off[c] = total;
    for (int i : cells[c]) cell_idx[total++] = i;
  }
  cell_off[cells.size()] = total;
  return total;
}
void orc_distinctive_descriptors(const uint8_t* desc, const int* offset, int n_points, int* best) {
  for (int p = 0; p < n_points; ++p) best[p] = matchport::distinctive_descriptor(desc + 32 * (size_t)offset[p], offset[p + 1] - offset[p]);
}
}
