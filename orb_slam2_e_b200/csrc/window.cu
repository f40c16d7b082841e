// Windowed projection search (orbm_window_search): the descriptor loop shared by the five
// ORBmatcher::SearchByProjection overloads (src/ORBmatcher.cc:46-132, :134-220, :491-604,
// :1529-1671, :1673-1800) behind Frame::GetFeaturesInArea (src/Frame.cc:342-395).
//
// The reference walks its queries one after another, and an accepted match writes
// mvpMapPoints[idx], which hides keypoint idx from every later query.  That order dependence is
// strictly lower-triangular (query q only sees the matches of queries p < q), so the sequential
// answer is the unique fixed point of "every query searches with the keypoints owned by earlier
// accepted queries hidden".  The kernel iterates exactly that map with all queries in parallel
// (one warp each) until nothing changes: pass k is right for the first k queries at least, real
// frames settle in two or three passes, and the result is bit-identical to the sequential loop.
//
// Layout: the frame is re-ordered on the host into grid order (cell (ix, iy) at ix*rows+iy, the
// reference's scan order), so a window column is ONE contiguous run of 16-byte keypoint records
// and 32-byte descriptors, and "earlier in the scan" is "smaller position", which breaks distance
// ties the way `dist < bestDist` does.  Keys are dist << 20 | position.
#include <algorithm>
#include <climits>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace orbfront {
namespace {

constexpr int kPosBits = 20;
constexpr uint32_t kEmptyKey = (256u << kPosBits) | ((1u << kPosBits) - 1);
constexpr int kWinWarps = 8;

struct WinArgs {
  const float4* rec;        // [npos] x, y, u_right, octave (as int bits), grid order
  const uint4* desc;        // [npos][2]
  const int* feat;          // [npos] keypoint index of a position
  const int* cell_off;      // [cols*rows + 1]
  const orbm_window_query* q;
  const uint4* qdesc;       // [nq][2]
  const int* owner_init;    // [npos] -1: hidden from everybody, INT_MAX: free
  int* owner;               // [3][npos] rotating: read | atomicMin | reset
  uint2* keys;              // [nq] best / second key of the last pass
  uint8_t* accepted;        // [nq]
  orbm_window_result* out;  // [nq]
  int* ctrl;                // [0..2] changed flags (rotating), [3] barrier, [4] passes, [5] accepted count
  int nq, npos, cols, rows;
  float min_x, min_y, width_inv, height_inv;
  int th;
  float ratio;
  int max_rounds;  // 1: no claims, single pass
  const float* inv_sigma2;  // Fuse's reprojection gate, or null
  double chi2_mono, chi2_stereo;
};

__device__ __forceinline__ int hamming256(const uint32_t (&q)[8], const uint4 a, const uint4 b) {
  return __popc(q[0] ^ a.x) + __popc(q[1] ^ a.y) + __popc(q[2] ^ a.z) + __popc(q[3] ^ a.w) + __popc(q[4] ^ b.x) +
         __popc(q[5] ^ b.y) + __popc(q[6] ^ b.z) + __popc(q[7] ^ b.w);
}

// All CTAs are co-resident (cooperative launch); monotonic arrival counter.
__device__ __forceinline__ void grid_barrier(int* bar, int target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1);
    while (*(volatile int*)bar < target) __nanosleep(32);
    __threadfence();
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kWinWarps * 32) window_search_kernel(const WinArgs A) {
  const int lane = threadIdx.x & 31;
  const int warp = blockIdx.x * kWinWarps + (threadIdx.x >> 5), n_warps = gridDim.x * kWinWarps;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, n_threads = gridDim.x * blockDim.x;
  int round = 0;
  for (;; ++round) {
    const int* cur = A.owner + (size_t)(round % 3) * A.npos;
    int* nxt = A.owner + (size_t)((round + 1) % 3) * A.npos;
    if (A.max_rounds > 1) {
      int* rst = A.owner + (size_t)((round + 2) % 3) * A.npos;
      for (int i = tid; i < A.npos; i += n_threads) rst[i] = A.owner_init[i];
      if (tid == 0) A.ctrl[(round + 1) % 3] = 0;
    }
    bool changed = false;
    for (int qi = warp; qi < A.nq; qi += n_warps) {
      const orbm_window_query Q = A.q[qi];
      const uint4 qa = __ldg(&A.qdesc[2 * qi]), qb = __ldg(&A.qdesc[2 * qi + 1]);
      const uint32_t qd[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
      // Frame.cc:347-361 -- same float expressions, same order
      const int cx0 = max(0, (int)floorf((Q.x - A.min_x - Q.r) * A.width_inv));
      const int cx1 = min(A.cols - 1, (int)ceilf((Q.x - A.min_x + Q.r) * A.width_inv));
      const int cy0 = max(0, (int)floorf((Q.y - A.min_y - Q.r) * A.height_inv));
      const int cy1 = min(A.rows - 1, (int)ceilf((Q.y - A.min_y + Q.r) * A.height_inv));
      uint32_t k1 = kEmptyKey, k2 = kEmptyKey;
      if (cx0 < A.cols && cx1 >= 0 && cy0 < A.rows && cy1 >= 0) {
        const bool check_levels = Q.min_level > 0 || Q.max_level >= 0;  // :363
        for (int ix = cx0; ix <= cx1; ++ix) {
          const int b = __ldg(&A.cell_off[ix * A.rows + cy0]), e = __ldg(&A.cell_off[ix * A.rows + cy1 + 1]);
          for (int pos = b + lane; pos < e; pos += 32) {
            const float4 r = __ldg(&A.rec[pos]);
            const int oct = __float_as_int(r.w);
            if (check_levels && (oct < Q.min_level || (Q.max_level >= 0 && oct > Q.max_level))) continue;  // :375-382
            if (!(fabsf(r.x - Q.x) < Q.r && fabsf(r.y - Q.y) < Q.r)) continue;                           // :384-388
            if (__ldcg(&cur[pos]) < qi) continue;  // hidden: claimed before the call, or owned by an earlier query
            if (Q.er_max >= 0.0f && r.z > 0.0f && fabsf(Q.u_right - r.z) > Q.er_max) continue;  // ORBmatcher.cc:90-95
            if (A.inv_sigma2) {  // ORBmatcher.cc:1118-1146
              const float ex = Q.x - r.x, ey = Q.y - r.y;
              if (r.z >= 0.0f) {
                const float er = Q.u_right - r.z;
                const float e2 = ex * ex + ey * ey + er * er;
                if ((double)(e2 * __ldg(&A.inv_sigma2[oct])) > A.chi2_stereo) continue;
              } else {
                const float e2 = ex * ex + ey * ey;
                if ((double)(e2 * __ldg(&A.inv_sigma2[oct])) > A.chi2_mono) continue;
              }
            }
            const uint32_t key = ((uint32_t)hamming256(qd, __ldg(&A.desc[2 * pos]), __ldg(&A.desc[2 * pos + 1])) << kPosBits) |
                                 (uint32_t)pos;
            const uint32_t hi = max(key, k1);
            k1 = min(key, k1);
            k2 = min(k2, hi);
          }
        }
      }
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) {
        const uint32_t o1 = __shfl_xor_sync(0xffffffffu, k1, s), o2 = __shfl_xor_sync(0xffffffffu, k2, s);
        const uint32_t hi = max(k1, o1);
        k1 = min(k1, o1);
        k2 = min(min(k2, o2), hi);
      }
      if (lane == 0) {
        const int d1 = (int)(k1 >> kPosBits), d2 = (int)(k2 >> kPosBits);
        bool acc = k1 != kEmptyKey && d1 <= A.th;
        if (acc && A.ratio > 0.0f && k2 != kEmptyKey) {  // ORBmatcher.cc:119-123: same level and ratio fails -> rejected
          const int l1 = __float_as_int(__ldg(&A.rec[k1 & ((1u << kPosBits) - 1)]).w);
          const int l2 = __float_as_int(__ldg(&A.rec[k2 & ((1u << kPosBits) - 1)]).w);
          if (l1 == l2 && (float)d1 > A.ratio * (float)d2) acc = false;
        }
        if (round > 0) {
          const uint2 old = A.keys[qi];
          changed |= old.x != k1 || old.y != k2 || (A.accepted[qi] != 0) != acc;
        }
        A.keys[qi] = make_uint2(k1, k2);
        A.accepted[qi] = acc;
        if (acc && Q.blocks && A.max_rounds > 1) atomicMin(&nxt[k1 & ((1u << kPosBits) - 1)], qi);
      }
    }
    if (A.max_rounds <= 1) break;
    if (changed || (round == 0 && threadIdx.x == 0)) atomicExch(&A.ctrl[round % 3], 1);  // pass 0 has nothing to compare with
    grid_barrier(&A.ctrl[3], (round + 1) * (int)gridDim.x);
    // settled when a pass changed nothing: its owners equal the previous ones, so the next pass would repeat it
    if (round > 0 && *(volatile int*)&A.ctrl[round % 3] == 0) break;
    if (round + 1 >= A.max_rounds) break;
  }
  // emit (keys -> keypoint indices and levels), by the lane that produced the keys
  int n_acc = 0;
  if (lane == 0) {
    for (int qi = warp; qi < A.nq; qi += n_warps) {
      const uint2 k = A.keys[qi];
      orbm_window_result R;
      R.idx = -1; R.best_idx = -1; R.best_dist = 256; R.best_dist2 = 256; R.best_level = -1; R.best_level2 = -1;
      if (k.x != kEmptyKey) {
        const int p = (int)(k.x & ((1u << kPosBits) - 1));
        R.best_idx = __ldg(&A.feat[p]);
        R.best_dist = (int)(k.x >> kPosBits);
        R.best_level = __float_as_int(__ldg(&A.rec[p]).w);
        if (A.accepted[qi]) { R.idx = R.best_idx; ++n_acc; }
      }
      if (k.y != kEmptyKey) {
        R.best_dist2 = (int)(k.y >> kPosBits);
        R.best_level2 = __float_as_int(__ldg(&A.rec[k.y & ((1u << kPosBits) - 1)]).w);
      }
      A.out[qi] = R;
    }
    if (n_acc) atomicAdd(&A.ctrl[5], n_acc);
  }
  if (tid == 0) A.ctrl[4] = round + 1;
}

// Candidate lists (orbm_window_candidates): same gather, every candidate written out in scan order.
// pass 0 counts and reserves a range with one atomicAdd per query; pass 1 fills.
struct CandArgs {
  const float4* rec; const uint4* desc; const int* feat; const int* cell_off;
  const orbm_window_query* q; const uint4* qdesc;
  int nq, cols, rows;
  float min_x, min_y, width_inv, height_inv;
  int* offset; int* count; int* cursor;  // [nq], [nq], [1]
  int2* out; int cap;                    // (keypoint, distance)
};

template <int kPass>
__global__ void __launch_bounds__(kWinWarps * 32) window_candidates_kernel(const CandArgs A) {
  const int lane = threadIdx.x & 31;
  const int qi = blockIdx.x * kWinWarps + (threadIdx.x >> 5);
  if (qi >= A.nq) return;
  const orbm_window_query Q = A.q[qi];
  const int cx0 = max(0, (int)floorf((Q.x - A.min_x - Q.r) * A.width_inv));
  const int cx1 = min(A.cols - 1, (int)ceilf((Q.x - A.min_x + Q.r) * A.width_inv));
  const int cy0 = max(0, (int)floorf((Q.y - A.min_y - Q.r) * A.height_inv));
  const int cy1 = min(A.rows - 1, (int)ceilf((Q.y - A.min_y + Q.r) * A.height_inv));
  uint32_t qd[8];
  int base = 0;
  if (kPass == 1) {
    const uint4 qa = __ldg(&A.qdesc[2 * qi]), qb = __ldg(&A.qdesc[2 * qi + 1]);
    qd[0] = qa.x; qd[1] = qa.y; qd[2] = qa.z; qd[3] = qa.w; qd[4] = qb.x; qd[5] = qb.y; qd[6] = qb.z; qd[7] = qb.w;
    base = A.offset[qi];
    if (base + A.count[qi] > A.cap) return;  // the host reports ORB_E_CAPACITY
  }
  int n = 0;
  if (cx0 < A.cols && cx1 >= 0 && cy0 < A.rows && cy1 >= 0) {
    const bool check_levels = Q.min_level > 0 || Q.max_level >= 0;
    for (int ix = cx0; ix <= cx1; ++ix) {
      const int b = __ldg(&A.cell_off[ix * A.rows + cy0]), e = __ldg(&A.cell_off[ix * A.rows + cy1 + 1]);
      for (int p0 = b; p0 < e; p0 += 32) {  // uniform trip count: the ballot below needs the whole warp
        const int pos = p0 + lane;
        bool in = pos < e;
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (in) {
          r = __ldg(&A.rec[pos]);
          const int oct = __float_as_int(r.w);
          if (check_levels && (oct < Q.min_level || (Q.max_level >= 0 && oct > Q.max_level))) in = false;
          if (!(fabsf(r.x - Q.x) < Q.r && fabsf(r.y - Q.y) < Q.r)) in = false;
        }
        const uint32_t m = __ballot_sync(0xffffffffu, in);
        if (kPass == 1 && in) {
          const int slot = base + n + __popc(m & ((1u << lane) - 1));
          A.out[slot] = make_int2(__ldg(&A.feat[pos]), hamming256(qd, __ldg(&A.desc[2 * pos]), __ldg(&A.desc[2 * pos + 1])));
        }
        n += __popc(m);
      }
    }
  }
  if (kPass == 0 && lane == 0) {
    A.count[qi] = n;
    A.offset[qi] = atomicAdd(A.cursor, n);
  }
}

}  // namespace
}  // namespace orbfront

using namespace orbfront;

extern "C" int orbm_window_search(int device, const orbm_grid_frame* F, const orbm_window_query* queries,
                                  const uint8_t* query_desc, int nq, const uint8_t* claimed, const orbm_window_params* prm,
                                  orbm_window_result* out, int* n_accepted, int* rounds) {
  if (!prm) return set_error(ORB_E_INVALID, "orbm_window_search: null params");
  const int th = prm->th, claim = prm->claim;
  const float ratio = prm->ratio;
  const bool gate = prm->inv_level_sigma2 != nullptr;
  if (gate && prm->n_levels < 1) return set_error(ORB_E_INVALID, "orbm_window_search: reprojection gate without levels");
  if (!F || nq < 0 || (nq > 0 && (!queries || !query_desc || !out)) || F->n < 0 || F->cols < 1 || F->rows < 1 ||
      !F->cell_off || (F->n > 0 && (!F->x || !F->y || !F->octave || !F->desc)))
    return set_error(ORB_E_INVALID, "orbm_window_search: bad argument");
  if (n_accepted) *n_accepted = 0;
  if (rounds) *rounds = 0;
  if (nq == 0) return ORB_OK;
  const long long n_cells = (long long)F->cols * F->rows;
  if (n_cells > (1 << 22)) return set_error(ORB_E_INVALID, "orbm_window_search: grid of %lld cells", n_cells);
  const int npos = F->cell_off[n_cells];
  if (F->cell_off[0] != 0 || npos < 0 || npos >= (1 << kPosBits) || (npos > 0 && !F->cell_idx))
    return set_error(ORB_E_INVALID, "orbm_window_search: bad cell table (%d entries)", npos);
  for (long long c = 0; c < n_cells; ++c)
    if (F->cell_off[c + 1] < F->cell_off[c]) return set_error(ORB_E_INVALID, "orbm_window_search: cell offsets not ascending");
  int rc = use_device(device);
  if (rc) return rc;

  Layout L;  // upload block
  const size_t o_rec = L.place((size_t)npos * 16), o_desc = L.place((size_t)npos * 32), o_feat = L.place((size_t)npos * 4),
               o_off = L.place((size_t)(n_cells + 1) * 4), o_q = L.place((size_t)nq * sizeof(orbm_window_query)),
               o_qd = L.place((size_t)nq * 32), o_init = L.place((size_t)npos * 4), o_owner = L.place((size_t)npos * 12),
               o_ctrl = L.place(32), o_sig = L.place(gate ? (size_t)prm->n_levels * 4 : 0);
  const size_t o_up_end = L.bytes;  // device-only below, then the download block
  const size_t o_keys = L.place((size_t)nq * 8), o_acc = L.place(nq);
  const size_t o_out = L.place((size_t)nq * sizeof(orbm_window_result)), o_end = L.bytes;
  const size_t o_seen = L.place((size_t)F->n);  // host only
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  uint8_t *h = S->host, *d = S->dev;

  // frame -> grid order
  float* rec = reinterpret_cast<float*>(h + o_rec);
  int* feat = reinterpret_cast<int*>(h + o_feat);
  int* init = reinterpret_cast<int*>(h + o_init);
  uint8_t* seen = h + o_seen;
  memset(seen, 0, (size_t)F->n);
  for (int p = 0; p < npos; ++p) {
    const uint32_t f = F->cell_idx[p];
    if (f >= (uint32_t)F->n) return set_error(ORB_E_INVALID, "orbm_window_search: cell entry %u out of range", f);
    if (seen[f]) return set_error(ORB_E_INVALID, "orbm_window_search: keypoint %u listed in two cells", f);
    seen[f] = 1;
    rec[4 * p + 0] = F->x[f];
    rec[4 * p + 1] = F->y[f];
    rec[4 * p + 2] = F->u_right ? F->u_right[f] : -1.0f;
    memcpy(&rec[4 * p + 3], &F->octave[f], 4);
    if (gate && (F->octave[f] < 0 || F->octave[f] >= prm->n_levels))
      return set_error(ORB_E_INVALID, "orbm_window_search: keypoint %u has octave %d outside the sigma table", f, F->octave[f]);
    memcpy(h + o_desc + (size_t)p * 32, F->desc + (size_t)f * 32, 32);
    feat[p] = (int)f;
    init[p] = (claimed && claimed[f]) ? -1 : INT_MAX;
  }
  memcpy(h + o_off, F->cell_off, (size_t)(n_cells + 1) * 4);
  memcpy(h + o_q, queries, (size_t)nq * sizeof(orbm_window_query));
  memcpy(h + o_qd, query_desc, (size_t)nq * 32);
  for (int k = 0; k < 3; ++k) memcpy(h + o_owner + (size_t)k * npos * 4, init, (size_t)npos * 4);
  memset(h + o_ctrl, 0, 32);
  if (gate) memcpy(h + o_sig, prm->inv_level_sigma2, (size_t)prm->n_levels * 4);
  ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_up_end, cudaMemcpyHostToDevice, S->stream));

  WinArgs A;
  A.rec = (const float4*)(d + o_rec); A.desc = (const uint4*)(d + o_desc); A.feat = (const int*)(d + o_feat);
  A.cell_off = (const int*)(d + o_off); A.q = (const orbm_window_query*)(d + o_q); A.qdesc = (const uint4*)(d + o_qd);
  A.owner_init = (const int*)(d + o_init); A.owner = (int*)(d + o_owner); A.keys = (uint2*)(d + o_keys);
  A.accepted = d + o_acc; A.out = (orbm_window_result*)(d + o_out); A.ctrl = (int*)(d + o_ctrl);
  A.nq = nq; A.npos = npos; A.cols = F->cols; A.rows = F->rows;
  A.min_x = F->min_x; A.min_y = F->min_y; A.width_inv = F->width_inv; A.height_inv = F->height_inv;
  A.th = th; A.ratio = ratio;
  A.inv_sigma2 = gate ? (const float*)(d + o_sig) : nullptr;
  A.chi2_mono = prm->chi2_mono; A.chi2_stereo = prm->chi2_stereo;
  A.max_rounds = claim ? nq + 1 : 1;
  int grid = div_up(nq, kWinWarps);
  if (claim) {  // every CTA must be resident: the passes meet at a grid barrier
    static thread_local int max_grid[16] = {0};
    if (!max_grid[device]) {
      int per_sm = 0, sms = 0;
      ORB_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, window_search_kernel, kWinWarps * 32, 0));
      ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
      max_grid[device] = std::max(1, per_sm * sms);
    }
    grid = std::min(grid, max_grid[device]);
    void* args[] = {(void*)&A};
    ORB_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)window_search_kernel, dim3(grid), dim3(kWinWarps * 32), args, 0, S->stream));
    count_launch();
  } else {
    window_search_kernel<<<grid, kWinWarps * 32, 0, S->stream>>>(A);
    ORB_LAUNCH_CHECK();
  }
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_out, d + o_out, o_end - o_out, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_ctrl, d + o_ctrl, 32, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  memcpy(out, h + o_out, (size_t)nq * sizeof(orbm_window_result));
  const int* ctrl = reinterpret_cast<const int*>(h + o_ctrl);
  if (n_accepted) *n_accepted = ctrl[5];
  if (rounds) *rounds = ctrl[4];
  return ORB_OK;
}

extern "C" int orbm_window_candidates(int device, const orbm_grid_frame* F, const orbm_window_query* queries,
                                      const uint8_t* query_desc, int nq, int32_t* offset, int32_t* count, int32_t* cand_idx,
                                      int32_t* cand_dist, int cap, int* n_total) {
  if (!F || nq < 0 || cap < 0 || !n_total || (nq > 0 && (!queries || !query_desc || !offset || !count)) || F->n < 0 || F->cols < 1 ||
      F->rows < 1 || !F->cell_off || (F->n > 0 && (!F->x || !F->y || !F->octave || !F->desc)) || (cap > 0 && (!cand_idx || !cand_dist)))
    return set_error(ORB_E_INVALID, "orbm_window_candidates: bad argument");
  *n_total = 0;
  if (nq == 0) return ORB_OK;
  const long long n_cells = (long long)F->cols * F->rows;
  if (n_cells > (1 << 22)) return set_error(ORB_E_INVALID, "orbm_window_candidates: grid of %lld cells", n_cells);
  const int npos = F->cell_off[n_cells];
  if (F->cell_off[0] != 0 || npos < 0 || npos >= (1 << kPosBits) || (npos > 0 && !F->cell_idx))
    return set_error(ORB_E_INVALID, "orbm_window_candidates: bad cell table (%d entries)", npos);
  for (long long c = 0; c < n_cells; ++c)
    if (F->cell_off[c + 1] < F->cell_off[c]) return set_error(ORB_E_INVALID, "orbm_window_candidates: cell offsets not ascending");
  int rc = use_device(device);
  if (rc) return rc;
  Layout L;
  const size_t o_rec = L.place((size_t)npos * 16), o_desc = L.place((size_t)npos * 32), o_feat = L.place((size_t)npos * 4),
               o_off = L.place((size_t)(n_cells + 1) * 4), o_q = L.place((size_t)nq * sizeof(orbm_window_query)),
               o_qd = L.place((size_t)nq * 32), o_cur = L.place(4);
  const size_t o_up_end = L.bytes;
  const size_t o_oc = L.place((size_t)nq * 8);  // offset[nq], count[nq]
  const size_t o_out = L.place((size_t)cap * 8), o_end = L.bytes;
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  uint8_t *h = S->host, *d = S->dev;
  float* rec = reinterpret_cast<float*>(h + o_rec);
  int* feat = reinterpret_cast<int*>(h + o_feat);
  for (int p = 0; p < npos; ++p) {
    const uint32_t f = F->cell_idx[p];
    if (f >= (uint32_t)F->n) return set_error(ORB_E_INVALID, "orbm_window_candidates: cell entry %u out of range", f);
    rec[4 * p + 0] = F->x[f];
    rec[4 * p + 1] = F->y[f];
    rec[4 * p + 2] = -1.0f;
    memcpy(&rec[4 * p + 3], &F->octave[f], 4);
    memcpy(h + o_desc + (size_t)p * 32, F->desc + (size_t)f * 32, 32);
    feat[p] = (int)f;
  }
  memcpy(h + o_off, F->cell_off, (size_t)(n_cells + 1) * 4);
  memcpy(h + o_q, queries, (size_t)nq * sizeof(orbm_window_query));
  memcpy(h + o_qd, query_desc, (size_t)nq * 32);
  memset(h + o_cur, 0, 4);
  ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_up_end, cudaMemcpyHostToDevice, S->stream));
  CandArgs A;
  A.rec = (const float4*)(d + o_rec); A.desc = (const uint4*)(d + o_desc); A.feat = (const int*)(d + o_feat);
  A.cell_off = (const int*)(d + o_off); A.q = (const orbm_window_query*)(d + o_q); A.qdesc = (const uint4*)(d + o_qd);
  A.nq = nq; A.cols = F->cols; A.rows = F->rows;
  A.min_x = F->min_x; A.min_y = F->min_y; A.width_inv = F->width_inv; A.height_inv = F->height_inv;
  A.offset = (int*)(d + o_oc); A.count = A.offset + nq; A.cursor = (int*)(d + o_cur);
  A.out = (int2*)(d + o_out); A.cap = cap;
  const int grid = div_up(nq, kWinWarps);
  window_candidates_kernel<0><<<grid, kWinWarps * 32, 0, S->stream>>>(A);
  ORB_LAUNCH_CHECK();
  window_candidates_kernel<1><<<grid, kWinWarps * 32, 0, S->stream>>>(A);
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_oc, d + o_oc, (size_t)nq * 8, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_cur, d + o_cur, 4, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  const int total = *reinterpret_cast<const int*>(h + o_cur);
  *n_total = total;
  memcpy(offset, h + o_oc, (size_t)nq * 4);
  memcpy(count, h + o_oc + (size_t)nq * 4, (size_t)nq * 4);
  if (total > cap) return set_error(ORB_E_CAPACITY, "orbm_window_candidates: %d candidates, capacity %d", total, cap);
  if (total == 0) return ORB_OK;
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_out, d + o_out, (size_t)total * 8, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  const int* pairs = reinterpret_cast<const int*>(h + o_out);
  for (int k = 0; k < total; ++k) { cand_idx[k] = pairs[2 * k]; cand_dist[k] = pairs[2 * k + 1]; }
  return ORB_OK;
}
