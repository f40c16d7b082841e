// tests/cpp/port_scenes.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// The product's include/ORBmatcher.h compiled against a TEST-ONLY backend: the C-ABI entry points
// the header calls are defined here on top of the CPU restatement (oracle/match_port.hpp) instead
// of liborbfront.so.  Purpose, for the `-m "not gpu"` suite only:
//   * run the header's host logic (projection, query construction, result application) without a GPU
//   * pin the restatement: portm_* must equal refm_* (the reference's own src/ORBmatcher.cc,
//     oracle/_ref/libmatchref.so) on every scene, so the GPU parity tests that compare
//     orbm_window_search / orbm_search_bow with the restatement are anchored on the reference.
// Nothing links this into the product; libdropin.so (the GPU build of the same scenes) links
// liborbfront.so and has no CPU path.
#include <cstdarg>
#include <cstdio>

#include "../../oracle/match_port.hpp"
#include "ORBmatcher.h"

extern "C" {
const char* orb_last_error(void) { return "port backend"; }
int orbm_hamming(const uint8_t* a, const uint8_t* b) { return matchport::hamming256(a, b); }
int orbm_search_bow(int, const uint8_t* desc1, int n1, const uint8_t* desc2, int n2, const orbm_featvec* fv1,
                    const orbm_featvec* fv2, const uint8_t* valid1, const uint8_t* valid2, const float* angle1,
                    const float* angle2, float nnratio, int check_orientation, int kf_kf, int32_t* out, int* n_matches) {
  matchport::FeatVec f1{fv1->node, fv1->off, fv1->idx, fv1->n_nodes}, f2{fv2->node, fv2->off, fv2->idx, fv2->n_nodes};
  *n_matches = matchport::search_bow(desc1, n1, desc2, n2, f1, f2, valid1, valid2, angle1, angle2, nnratio, check_orientation,
                                     kf_kf != 0, out);
  return ORB_OK;
}
int orbm_window_search(int, const orbm_grid_frame* F, const orbm_window_query* q, const uint8_t* qd, int nq,
                       const uint8_t* claimed, const orbm_window_params* prm, orbm_window_result* out, int* n_accepted,
                       int* rounds) {
  static_assert(sizeof(orbm_window_query) == sizeof(matchport::WindowQuery), "query layout");
  static_assert(sizeof(orbm_window_result) == sizeof(matchport::WindowResult), "result layout");
  matchport::GridFrame G{F->x, F->y, F->octave, F->u_right, F->desc, F->n, F->min_x, F->min_y, F->width_inv, F->height_inv,
                         F->cols, F->rows, F->cell_off, F->cell_idx};
  matchport::WindowGate gate{prm->inv_level_sigma2, prm->chi2_mono, prm->chi2_stereo};
  const int n = matchport::window_search(G, (const matchport::WindowQuery*)q, qd, nq, claimed, prm->th, prm->ratio, prm->claim != 0,
                                         (matchport::WindowResult*)out, prm->inv_level_sigma2 ? &gate : nullptr);
  if (n_accepted) *n_accepted = n;
  if (rounds) *rounds = 1;
  return ORB_OK;
}
int orbm_window_candidates(int, const orbm_grid_frame* F, const orbm_window_query* q, const uint8_t* qd, int nq, int32_t* offset,
                           int32_t* count, int32_t* cand_idx, int32_t* cand_dist, int cap, int* n_total) {
  matchport::GridFrame G{F->x, F->y, F->octave, F->u_right, F->desc, F->n, F->min_x, F->min_y, F->width_inv, F->height_inv,
                         F->cols, F->rows, F->cell_off, F->cell_idx};
  std::vector<int> v;
  int total = 0;
  for (int i = 0; i < nq; ++i) {
    matchport::features_in_area(G, q[i].x, q[i].y, q[i].r, q[i].min_level, q[i].max_level, v);
    offset[i] = total;
    count[i] = (int)v.size();
    for (size_t k = 0; k < v.size(); ++k, ++total)
      if (total < cap) { cand_idx[total] = v[k]; cand_dist[total] = matchport::hamming256(qd + 32 * (size_t)i, F->desc + 32 * (size_t)v[k]); }
  }
  *n_total = total;
  return total > cap ? ORB_E_CAPACITY : ORB_OK;
}
int orbm_search_triangulation(int, const orbm_tri_keyframe* k1, const orbm_tri_keyframe* k2, const float* F12, float ex, float ey,
                              const float* sf2, const float* sg2, int, int check_ori, int32_t* matches12, int* n_matches) {
  auto side = [](const orbm_tri_keyframe* k) {
    return matchport::TriKeyFrame{k->desc, k->x, k->y, k->octave, k->angle, k->usable, k->stereo, k->n,
                                  matchport::FeatVec{k->fv->node, k->fv->off, k->fv->idx, k->fv->n_nodes}};
  };
  *n_matches = matchport::search_triangulation(side(k1), side(k2), F12, ex, ey, sf2, sg2, check_ori, matches12);
  return ORB_OK;
}
int orbm_stereo_search(int, const orbx_keypoint*, const uint8_t*, int, const orbx_keypoint*, const uint8_t*, int, const float*, int,
                       int, float, float, int32_t*, int32_t*) {
  return ORB_E_INVALID;  // not exercised by the scenes
}
}

float ORB_SLAM2::Frame::mnMinX, ORB_SLAM2::Frame::mnMinY, ORB_SLAM2::Frame::mnMaxX, ORB_SLAM2::Frame::mnMaxY;
float ORB_SLAM2::Frame::mfGridElementWidthInv, ORB_SLAM2::Frame::mfGridElementHeightInv;
#define SCN(name) portm_##name
#include "matcher_scenes.inc"
