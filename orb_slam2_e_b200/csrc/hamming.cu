// 256-bit Hamming descriptor search for sm_100a: brute-force top-2 against a resident
// database shard, BoW-node search, stereo row-band search, pairwise distances.
//
// Replaces the scalar loops of the reference (all CPU):
//   ORBmatcher::DescriptorDistance   src/ORBmatcher.cc:1848-1864
//   top-2 scan + ratio test          src/ORBmatcher.cc:417-431
//   SearchByBoW (KF,Frame)/(KF,KF)   src/ORBmatcher.cc:360-489, 723-856
//   ComputeStereoMatches (search)    src/Frame.cc:527-610
//
// Bound: integer POPC issue rate (each database byte is reused by every query, so HBM
// traffic is ~nq times below the popc time).  Tensor cores are not used: XOR+POPC is not a
// multiply-accumulate contraction.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace orbfront {

// ------------------------------------------------------------------------------------
// top-2 scan.  One CTA = 8 warps x 32 lanes x QPL queries held in registers = 1024 queries;
// every lane walks the same database rows, which arrive in shared memory through a 4-stage
// 1-D TMA bulk pipeline and are read as broadcast LDS.128.  Per pair: 8 LOP3 (xor) + 8 POPC +
// 4 IADD3, then the reference's two-best update on packed keys (distance << 20 | local row):
// the smallest key is best1 with the lowest row on ties, the second smallest key carries the
// multiset second-best distance.
constexpr int kQPL = 4;                 // queries per lane
constexpr int kTopWarps = 8;
constexpr int kTopThreads = kTopWarps * 32;
constexpr int kQPerCta = kTopThreads * kQPL;  // 1024
constexpr int kTileRows = 256;          // 8 KB per stage
constexpr int kStages = 4;
constexpr int kRowBits = 20;            // local row index bits inside a key
constexpr uint32_t kKeyNone = 256u << kRowBits;  // "no row yet": distance 256 (ORBmatcher.cc:402-404)

// 256-bit Hamming distance (ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1848-1864).  POPC issues at an eighth of the
// warp rate on sm_100a (4 lanes / clock / scheduler: 4.43 T popc/s measured), LOP3 at half of it, so the eight XOR words
// are first reduced with carry-save adders -- a + b + c = (a ^ b ^ c) + 2 * maj(a, b, c) bit by bit -- to one word of
// weight 1 and four of weight 2: 5 POPC + 8 LOP3 instead of 8 POPC.  Same integer, by construction.
__device__ __forceinline__ int hamming8(const uint32_t (&q)[8], const uint4& a, const uint4& b) {
  const uint32_t x0 = q[0] ^ a.x, x1 = q[1] ^ a.y, x2 = q[2] ^ a.z, x3 = q[3] ^ a.w;
  const uint32_t x4 = q[4] ^ b.x, x5 = q[5] ^ b.y, x6 = q[6] ^ b.z, x7 = q[7] ^ b.w;
  const uint32_t s1 = x0 ^ x1 ^ x2, c1 = (x0 & x1) | (x2 & (x0 | x1));
  const uint32_t s2 = x3 ^ x4 ^ x5, c2 = (x3 & x4) | (x5 & (x3 | x4));
  const uint32_t s3 = s1 ^ s2 ^ x6, c3 = (s1 & s2) | (x6 & (s1 | s2));
  const uint32_t ones = s3 ^ x7, c4 = s3 & x7;
  return __popc(ones) + 2 * ((__popc(c1) + __popc(c2)) + (__popc(c3) + __popc(c4)));
}

__global__ void __launch_bounds__(kTopThreads)
top2_scan_kernel(const uint4* __restrict__ db, long long n_rows, int rows_per_cta,
                 const uint32_t* __restrict__ queries, int nq, int nq_pad, uint2* __restrict__ partial) {
  __shared__ __align__(128) uint4 tile[kStages][kTileRows * 2];
  __shared__ __align__(8) uint64_t full[kStages];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long row0 = (long long)blockIdx.x * rows_per_cta;
  const int my_rows = (int)min((long long)rows_per_cta, n_rows - row0);
  const int n_tiles = (my_rows + kTileRows - 1) / kTileRows;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&full[s], 1);
    fence_barrier_init();
  }
  __syncthreads();

  auto issue = [&](int t) {
    const int rows = min(kTileRows, my_rows - t * kTileRows);
    const uint32_t bytes = (uint32_t)rows * 32u;
    mbar_expect_tx(&full[t % kStages], bytes);
    tma_load_1d(&tile[t % kStages][0], db + (row0 + (long long)t * kTileRows) * 2, bytes, &full[t % kStages]);
  };
  if (tid == 0)
    for (int t = 0; t < kStages - 1 && t < n_tiles; ++t) issue(t);

  // queries -> registers (zero rows beyond nq; their results are never read).  The distance is computed with the
  // carry-save reduction of hamming8(), arranged so that every step is ONE three-input LOP3: with x_i = q_i ^ r_i,
  //   sum(x0,x1,x2) = (q0^q1^q2) ^ (r0^r1^r2)            maj(a,b,c) = a ^ ((a^b) & (a^c))
  // the query keeps q0, q0^q1, q0^q2, q0^q1^q2 (and the same for words 3..5) instead of q0..q5, and the row's
  // combinations r0^r1, r0^r2, r0^r1^r2 are formed once per row and shared by the lane's four queries:
  // 13 LOP3 + 1.5 shared + 5 POPC per pair instead of 8 LOP3 + 8 POPC -- the POPC pipe (an eighth of the issue rate)
  // was the limit.
  uint32_t qA[kQPL][4], qB[kQPL][4], q6[kQPL], q7[kQPL];
  uint32_t k1[kQPL], k2[kQPL];
  const int qbase = blockIdx.y * kQPerCta + warp * (32 * kQPL);
#pragma unroll
  for (int r = 0; r < kQPL; ++r) {
    const int qi = qbase + r * 32 + lane;
    const uint4* src = reinterpret_cast<const uint4*>(queries + (size_t)qi * 8);
    uint4 a = make_uint4(0, 0, 0, 0), b = a;
    if (qi < nq) { a = __ldg(src); b = __ldg(src + 1); }
    qA[r][0] = a.x; qA[r][1] = a.x ^ a.y; qA[r][2] = a.x ^ a.z; qA[r][3] = a.x ^ a.y ^ a.z;
    qB[r][0] = a.w; qB[r][1] = a.w ^ b.x; qB[r][2] = a.w ^ b.y; qB[r][3] = a.w ^ b.x ^ b.y;
    q6[r] = b.z; q7[r] = b.w;
    k1[r] = kKeyNone; k2[r] = kKeyNone;
  }

  for (int t = 0; t < n_tiles; ++t) {
    if (tid == 0 && t + kStages - 1 < n_tiles) issue(t + kStages - 1);
    mbar_wait(&full[t % kStages], (uint32_t)((t / kStages) & 1));
    const uint4* rows = &tile[t % kStages][0];
    const int n = min(kTileRows, my_rows - t * kTileRows);
    const uint32_t jbase = (uint32_t)(t * kTileRows);
#pragma unroll 2
    for (int j = 0; j < n; ++j) {
      const uint4 a = rows[2 * j], b = rows[2 * j + 1];
      const uint32_t r01 = a.x ^ a.y, r02 = a.x ^ a.z, r012 = r01 ^ a.z;
      const uint32_t r34 = a.w ^ b.x, r35 = a.w ^ b.y, r345 = r34 ^ b.y;
#pragma unroll
      for (int r = 0; r < kQPL; ++r) {
        const uint32_t s1 = qA[r][3] ^ r012;
        const uint32_t c1 = qA[r][0] ^ a.x ^ ((qA[r][1] ^ r01) & (qA[r][2] ^ r02));
        const uint32_t s2 = qB[r][3] ^ r345;
        const uint32_t c2 = qB[r][0] ^ a.w ^ ((qB[r][1] ^ r34) & (qB[r][2] ^ r35));
        const uint32_t x6 = q6[r] ^ b.z;
        const uint32_t s3 = s1 ^ s2 ^ x6, c3 = s1 ^ ((s1 ^ s2) & (s1 ^ x6));
        const uint32_t ones = s3 ^ q7[r] ^ b.w, c4 = s3 & (q7[r] ^ b.w);
        const int dist = __popc(ones) + 2 * ((__popc(c1) + __popc(c2)) + (__popc(c3) + __popc(c4)));
        const uint32_t key = ((uint32_t)dist << kRowBits) | (jbase + j);
        const uint32_t hi = max(key, k1[r]);
        k1[r] = min(key, k1[r]);
        k2[r] = min(k2[r], hi);
      }
    }
    __syncthreads();  // everyone is done with this stage before it is refilled
  }

#pragma unroll
  for (int r = 0; r < kQPL; ++r) {
    const int qi = qbase + r * 32 + lane;
    if (qi < nq_pad) partial[(size_t)blockIdx.x * nq_pad + qi] = make_uint2(k1[r], k2[r]);
  }
}

// Folds the per-CTA partials (in increasing row order) into one orbm_top2 per query.
__global__ void top2_reduce_kernel(const uint2* __restrict__ partial, int n_chunks, int rows_per_cta, int nq,
                                   int nq_pad, long long row_base, orbm_top2* __restrict__ out) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  int b1 = 256, b2 = 256;
  long long idx = -1;
  for (int c = 0; c < n_chunks; ++c) {
    const uint2 p = partial[(size_t)c * nq_pad + qi];
    const int d1 = (int)(p.x >> kRowBits), d2 = (int)(p.y >> kRowBits);
    if (d1 < b1) {  // strict: earlier chunk (lower rows) wins ties, as the sequential scan does
      b2 = b1; b1 = d1;
      idx = row_base + (long long)c * rows_per_cta + (long long)(p.x & ((1u << kRowBits) - 1));
    } else if (d1 < b2) {
      b2 = d1;
    }
    if (d2 < b2) b2 = d2;
  }
  orbm_top2 r;
  r.best1 = b1; r.best2 = b2; r.idx = idx;
  out[qi] = r;
}

__global__ void hamming_pairs_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int n,
                                     int* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 a0 = a[2 * i], a1 = a[2 * i + 1], b0 = b[2 * i], b1 = b[2 * i + 1];
  out[i] = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// POPC issue-rate probe: 8 independent chains per thread, 2 alu ops per POPC (below the
// alu pipe's 4x higher rate), result stored so nothing is eliminated.
__global__ void __launch_bounds__(256) popc_probe_kernel(uint32_t seed, int iters, uint32_t* out) {
  uint32_t x[8], acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { x[i] = seed + threadIdx.x * 8 + i + blockIdx.x * 2048; acc[i] = 0; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) { acc[i] += __popc(x[i]); x[i] += 0x9E3779B9u; }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}


// ------------------------------------------------------------------------------------
// SearchByBoW.  The reference's greedy claim (ORBmatcher.cc:410 / :777,804: a side-2 feature taken
// by an earlier side-1 feature is skipped) makes results depend on scan order -- but only inside a
// vocabulary node: a feature belongs to exactly one node of a DBoW2 FeatureVector, so claims never
// cross nodes.  One warp per common node therefore replays the reference order exactly (side-1
// features sequentially, the node's side-2 candidates across the lanes, warp top-2 merge on packed
// keys distance << 16 | list position) while the nodes run in parallel.  If a caller passes lists in
// which a feature appears in two common nodes, the host launches a single warp that walks the
// nodes in order instead.  The rotation histogram is global: atomics, then a one-warp finish kernel.
struct BowPair { int o1, l1, o2, l2; };  // common node: ranges in idx1 / idx2

__device__ __forceinline__ void warp_top2_merge(uint32_t& k1, uint32_t& k2) {
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    const uint32_t o1 = __shfl_xor_sync(0xffffffffu, k1, s), o2 = __shfl_xor_sync(0xffffffffu, k2, s);
    const uint32_t hi = max(k1, o1);
    k1 = min(k1, o1);
    k2 = min(min(k2, o2), hi);
  }
}

struct BowArgs {
  const uint4* desc1; const uint4* desc2;
  const BowPair* pairs; int n_pairs;
  const uint32_t* idx1; const uint32_t* idx2;
  const uint8_t* valid1; const uint8_t* valid2;  // valid2 may be null
  const float* ang1; const float* ang2;
  float ratio; int check_ori, kf_kf, n_out;
  uint8_t* claimed2; int* out; uint8_t* bin_of;
  int* hist;       // [ORBM_HISTO_LENGTH]
  int* n_matches;  // accepted matches (before the orientation filter)
};

__device__ __forceinline__ void bow_nodes(const BowArgs& A) {
  const int lane = threadIdx.x & 31;
  const int warp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), n_warps = gridDim.x * (blockDim.x >> 5);
  int nm = 0;
  for (int p = warp; p < A.n_pairs; p += n_warps) {
    const BowPair pr = A.pairs[p];
    for (int a = 0; a < pr.l1; ++a) {
      const int i1 = (int)A.idx1[pr.o1 + a];
      if (!A.valid1[i1]) continue;
      const uint4 qa = A.desc1[2 * i1], qb = A.desc1[2 * i1 + 1];
      const uint32_t q[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
      uint32_t k1 = 256u << 16, k2 = 256u << 16;
      for (int b = lane; b < pr.l2; b += 32) {
        const int i2 = (int)A.idx2[pr.o2 + b];
        if (A.claimed2[i2]) continue;
        if (A.valid2 && !A.valid2[i2]) continue;
        const uint32_t key = ((uint32_t)hamming8(q, A.desc2[2 * i2], A.desc2[2 * i2 + 1]) << 16) | (uint32_t)b;
        const uint32_t hi = max(key, k1);
        k1 = min(key, k1);
        k2 = min(k2, hi);
      }
      warp_top2_merge(k1, k2);
      const int best1 = (int)(k1 >> 16), best2 = min((int)(k2 >> 16), 256);
      const bool th_ok = A.kf_kf ? (best1 < ORBM_TH_LOW) : (best1 <= ORBM_TH_LOW);
      if (best1 < 256 && th_ok && (float)best1 < A.ratio * (float)best2) {  // uniform across the warp
        const int i2 = (int)A.idx2[pr.o2 + (int)(k1 & 0xffffu)];
        if (lane == 0) {
          A.claimed2[i2] = 1;
          const int slot = A.kf_kf ? i1 : i2;
          A.out[slot] = A.kf_kf ? i2 : i1;
          if (A.check_ori) {
            float rot = A.ang1[i1] - A.ang2[i2];
            if (rot < 0.0f) rot += 360.0f;
            int bin = (int)roundf(rot * (1.0f / ORBM_HISTO_LENGTH));
            if (bin == ORBM_HISTO_LENGTH) bin = 0;
            A.bin_of[slot] = (uint8_t)bin;
            atomicAdd(&A.hist[bin], 1);
          }
        }
        ++nm;
        __syncwarp();  // the claim is visible to the lanes' next candidate scan
      }
    }
  }
  if (lane == 0 && nm) atomicAdd(A.n_matches, nm);
}
__global__ void __launch_bounds__(128) bow_node_kernel(const __grid_constant__ BowArgs A) { bow_nodes(A); }
// batched form: problem blockIdx.y of K independent searches (one candidate keyframe each)
__global__ void __launch_bounds__(128) bow_node_batch_kernel(const BowArgs* __restrict__ args) {
  const BowArgs A = args[blockIdx.y];
  if (A.n_pairs > 0) bow_nodes(A);
}

// ComputeThreeMaxima (ORBmatcher.cc:1802-1843) and the removal of matches outside the three
// dominant rotation bins (:468-486).  One warp.
__device__ __forceinline__ void bow_finish(const int* __restrict__ hist, int* __restrict__ out, const uint8_t* __restrict__ bin_of,
                                           int n_out, int* __restrict__ n_matches) {
  const int lane = threadIdx.x;
  int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
  for (int i = 0; i < ORBM_HISTO_LENGTH; ++i) {
    const int s = hist[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if ((float)max3 < 0.1f * (float)max1) ind3 = -1;
  int dropped = 0;
  for (int i = lane; i < n_out; i += 32) {
    if (out[i] < 0) continue;
    const int b = bin_of[i];
    if (b != ind1 && b != ind2 && b != ind3) { out[i] = -1; ++dropped; }
  }
  for (int s = 16; s > 0; s >>= 1) dropped += __shfl_xor_sync(0xffffffffu, dropped, s);
  if (lane == 0) *n_matches -= dropped;
}
__global__ void __launch_bounds__(32) bow_finish_kernel(const int* __restrict__ hist, int* __restrict__ out,
                                                        const uint8_t* __restrict__ bin_of, int n_out, int* __restrict__ n_matches) {
  bow_finish(hist, out, bin_of, n_out, n_matches);
}
__global__ void __launch_bounds__(32) bow_finish_batch_kernel(const BowArgs* __restrict__ args) {
  const BowArgs A = args[blockIdx.x];
  if (A.n_pairs > 0 && A.check_ori) bow_finish(A.hist, A.out, A.bin_of, A.n_out, A.n_matches);
}

// ------------------------------------------------------------------------------------
// SearchForTriangulation (ORBmatcher.cc:858-1024).  No claims (vbMatched2 is never set), so one
// warp per (common node, side-1 keypoint) item; candidates across the lanes.  Key = distance << 16
// | (65535 - list position): min picks the smallest distance and, on ties, the LAST candidate.
struct TriItem { int i1, o2, l2; };
struct TriArgs {
  const uint4* desc1; const uint4* desc2;
  const TriItem* items; int n_items;
  const uint32_t* idx2;
  const float2* xy1; const float2* xy2; const int* oct2;
  const uint8_t* usable2; const uint8_t* stereo1; const uint8_t* stereo2;
  float F[9]; float ex, ey;
  const float* sf2; const float* sigma2;
  int* best;  // [n_items] side-2 index or -1
};

__global__ void __launch_bounds__(128) triangulation_kernel(const TriArgs A) {
  const int lane = threadIdx.x & 31;
  const int item = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (item >= A.n_items) return;
  const TriItem it = A.items[item];
  const uint4 qa = A.desc1[2 * it.i1], qb = A.desc1[2 * it.i1 + 1];
  const uint32_t q[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
  const float2 p1 = A.xy1[it.i1];
  const bool st1 = A.stereo1[it.i1] != 0;
  // epipolar line in image 2, ORBmatcher.cc:344-346
  const float a = p1.x * A.F[0] + p1.y * A.F[3] + A.F[6];
  const float b = p1.x * A.F[1] + p1.y * A.F[4] + A.F[7];
  const float c = p1.x * A.F[2] + p1.y * A.F[5] + A.F[8];
  const float den = a * a + b * b;
  uint32_t best = 0xffffffffu;
  if (den != 0.0f) {
    for (int k = lane; k < it.l2; k += 32) {
      const int i2 = (int)A.idx2[it.o2 + k];
      if (!A.usable2[i2]) continue;
      const int dist = hamming8(q, A.desc2[2 * i2], A.desc2[2 * i2 + 1]);
      if (dist > ORBM_TH_LOW) continue;
      const float2 p2 = A.xy2[i2];
      const int oc = A.oct2[i2];
      if (!st1 && !A.stereo2[i2]) {  // too close to the epipole, :921-927
        const float dx = A.ex - p2.x, dy = A.ey - p2.y;
        if (dx * dx + dy * dy < 100 * A.sf2[oc]) continue;
      }
      const float num = a * p2.x + b * p2.y + c;
      const float dsqr = num * num / den;
      if (!((double)dsqr < 3.84 * (double)A.sigma2[oc])) continue;
      best = min(best, ((uint32_t)dist << 16) | (uint32_t)(65535 - k));
    }
  }
  for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
  if (lane == 0) A.best[item] = best == 0xffffffffu ? -1 : (int)A.idx2[it.o2 + 65535 - (int)(best & 0xffffu)];
}

// ------------------------------------------------------------------------------------
// Stereo row-band search (Frame.cc:537-610): one warp per left keypoint; the reference's
// per-row candidate lists are replaced by evaluating the band predicate directly, which
// visits the same candidates in the same (increasing iR) order.
__device__ __forceinline__ void stereo_search_one(int i, int lane, const orbx_keypoint* __restrict__ L, const uint4* __restrict__ ldesc,
                                                  const orbx_keypoint* __restrict__ R, const uint4* __restrict__ rdesc, int nr,
                                                  const float* __restrict__ scale, int n_rows, float min_d, float max_d,
                                                  int* __restrict__ best_idx, int* __restrict__ best_dist) {
  const orbx_keypoint kl = L[i];
  const int row = (int)kl.y;
  const float minU = kl.x - max_d, maxU = kl.x - min_d;
  const uint4 qa = ldesc[2 * i], qb = ldesc[2 * i + 1];
  const uint32_t q[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
  uint32_t best = (uint32_t)ORBM_TH_HIGH << kRowBits;
  int in_band = 0;
  if (row >= 0 && row < n_rows) {
    for (int c = lane; c < nr; c += 32) {
      const orbx_keypoint kr = R[c];
      const float r = 2.0f * scale[kr.octave];
      const int maxr = (int)ceilf(kr.y + r), minr = (int)floorf(kr.y - r);
      if (row < minr || row > maxr) continue;
      in_band = 1;
      if (kr.octave < kl.octave - 1 || kr.octave > kl.octave + 1) continue;
      if (kr.x >= minU && kr.x <= maxU) {
        const uint32_t key = ((uint32_t)hamming8(q, rdesc[2 * c], rdesc[2 * c + 1]) << kRowBits) | (uint32_t)c;
        best = min(best, key);
      }
    }
  }
  in_band = __any_sync(0xffffffffu, in_band);
  for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
  if (lane == 0) {
    if (!in_band || maxU < 0) { best_idx[i] = 0; best_dist[i] = -1; }
    else {
      const int d = (int)(best >> kRowBits);
      best_dist[i] = d;
      best_idx[i] = d < ORBM_TH_HIGH ? (int)(best & ((1u << kRowBits) - 1)) : 0;
    }
  }
}


__global__ void __launch_bounds__(128)
stereo_search_kernel(const orbx_keypoint* __restrict__ L, const uint4* __restrict__ ldesc, int nl,
                     const orbx_keypoint* __restrict__ R, const uint4* __restrict__ rdesc, int nr,
                     const float* __restrict__ scale, int n_rows, float min_d, float max_d,
                     int* __restrict__ best_idx, int* __restrict__ best_dist) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= nl) return;
  stereo_search_one(i, lane, L, ldesc, R, rdesc, nr, scale, n_rows, min_d, max_d, best_idx, best_dist);
}

// ------------------------------------------------------------------------------------
// Sub-pixel stereo refinement (Frame.cc:613-684): for a left keypoint whose best right keypoint is
// closer than thOrbDist, slide an 11x11 patch of the left pyramid level over 11 columns of the
// right one (L1 distance of the centre-subtracted patches: integer arithmetic, exact), fit a
// parabola through the best offset and its neighbours, and turn the disparity into uRight / depth.
// One warp per left keypoint; both pyramids are read where the extractor left them in HBM.
struct StereoRefineArgs {
  const orbx_keypoint* L; const orbx_keypoint* R; int nl;
  const int* best_idx; const int* best_dist;
  PyramidView pl, pr;
  float min_d, max_d, mbf;
  int th_orb;
  float* u_right; float* depth; int* sad;  // [nl]; sad = -1 where nothing was assigned
};

// A = the pair's arrays; off_*: byte offsets of the pair's frames inside the pyramids of A.pl / A.pr (0 for the single-pair call)
__device__ __forceinline__ void stereo_refine_one(const StereoRefineArgs& A, int i, int lane, const orbx_keypoint* __restrict__ L,
                                                  const orbx_keypoint* __restrict__ R, const int* __restrict__ best_idx,
                                                  const int* __restrict__ best_dist, float* __restrict__ u_right, float* __restrict__ depth,
                                                  int* __restrict__ sad, size_t off_l0, size_t off_l, size_t off_r0, size_t off_r) {
  float out_u = -1.0f, out_z = -1.0f;
  int out_sad = -1;
  const int bd = best_dist[i];
  if (bd >= 0 && bd < A.th_orb) {  // warp-uniform
    const orbx_keypoint kl = L[i];
    const int oct = kl.octave;
    const float uL = kl.x;
    const float uR0 = R[best_idx[i]].x;
    const float sf = A.pl.inv_scale[oct];
    const float scaleduL = roundf(kl.x * sf), scaledvL = roundf(kl.y * sf), scaleduR0 = roundf(uR0 * sf);
    constexpr int w = 5, Ls = 5;
    const float iniu = scaleduR0 + Ls - w, endu = scaleduR0 + Ls + w + 1;
    const int x0 = (int)(scaleduL - w), y0 = (int)(scaledvL - w), r0 = (int)(scaleduR0 - Ls - w);
    const int W = A.pl.w[oct], H = A.pl.h[oct];
    // the reference's own guard (:636-638), plus staying inside the level (always true for keypoints
    // the extractor produced: they keep 16 px from every border of their level)
    const bool ok = !(iniu < 0 || endu >= A.pr.w[oct]) && x0 >= 0 && y0 >= 0 && x0 + 2 * w < W && y0 + 2 * w < H && r0 >= 0 &&
                    r0 + 2 * (w + Ls) < A.pr.w[oct] && y0 + 2 * w < A.pr.h[oct];
    if (ok) {
      const uint8_t* pl = A.pl.px[oct] + (oct == 0 ? off_l0 : off_l) + (size_t)y0 * A.pl.pitch[oct] + x0;
      const uint8_t* pr = A.pr.px[oct] + (oct == 0 ? off_r0 : off_r) + (size_t)y0 * A.pr.pitch[oct] + r0;
      const int cL = pl[(size_t)w * A.pl.pitch[oct] + w];
      int sums[2 * Ls + 1];
#pragma unroll
      for (int k = 0; k <= 2 * Ls; ++k) sums[k] = 0;
      for (int p = lane; p < (2 * w + 1) * (2 * w + 1); p += 32) {
        const int r = p / (2 * w + 1), c = p - r * (2 * w + 1);
        const int vl = (int)pl[(size_t)r * A.pl.pitch[oct] + c] - cL;
        const uint8_t* row = pr + (size_t)r * A.pr.pitch[oct] + c;
        const uint8_t* ctr = pr + (size_t)w * A.pr.pitch[oct] + w;
#pragma unroll
        for (int k = 0; k <= 2 * Ls; ++k) sums[k] += abs(vl - ((int)row[k] - (int)ctr[k]));
      }
#pragma unroll
      for (int k = 0; k <= 2 * Ls; ++k)
        for (int s = 16; s > 0; s >>= 1) sums[k] += __shfl_xor_sync(0xffffffffu, sums[k], s);
      int best = 0x7fffffff, bestinc = 0;  // :628-653, first minimum
#pragma unroll
      for (int k = 0; k <= 2 * Ls; ++k)
        if (sums[k] < best) { best = sums[k]; bestinc = k - Ls; }
      if (bestinc != -Ls && bestinc != Ls) {
        float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
        for (int k = 1; k < 2 * Ls; ++k)
          if (k == bestinc + Ls) { d1 = (float)sums[k - 1]; d2 = (float)sums[k]; d3 = (float)sums[k + 1]; }
        const float deltaR = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));  // :659-663
        if (!(deltaR < -1 || deltaR > 1)) {
          float bestuR = A.pl.scale[oct] * ((float)scaleduR0 + (float)bestinc + deltaR);
          float disparity = uL - bestuR;
          if (disparity >= A.min_d && disparity < A.max_d) {
            if (disparity <= 0) {
              disparity = 0.01;
              bestuR = (float)((double)uL - 0.01);
            }
            out_z = A.mbf / disparity;
            out_u = bestuR;
            out_sad = best;
          }
        }
      }
    }
  }
  if (lane == 0) { u_right[i] = out_u; depth[i] = out_z; sad[i] = out_sad; }
}


__global__ void __launch_bounds__(128) stereo_refine_kernel(const __grid_constant__ StereoRefineArgs A) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= A.nl) return;
  stereo_refine_one(A, i, lane, A.L, A.R, A.best_idx, A.best_dist, A.u_right, A.depth, A.sad, 0, 0, 0, 0);
}

// ------------------------------------------------------------------------------------
// Batched form: pair p = frames (2p, 2p+1) of ONE extractor handle's last batch; keypoints, descriptors and counts are read
// where the extraction left them on the device (BatchView), grid.y = pair.  Same per-keypoint code as the single-pair kernels.
struct StereoBatchArgs {
  BatchView B;
  StereoRefineArgs A;  // A.pl = A.pr = frame 0's pyramid; per-pair arrays are indexed [pair][out_cap]
  int* best_idx; int* best_dist;
  int n_rows;
};

__global__ void __launch_bounds__(128) stereo_search_batch_kernel(const __grid_constant__ StereoBatchArgs S) {
  const int lane = threadIdx.x & 31, p = blockIdx.y;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int nl = min(S.B.count[2 * p], S.B.out_cap), nr = min(S.B.count[2 * p + 1], S.B.out_cap);
  if (i >= nl) return;
  const size_t fl = (size_t)(2 * p) * S.B.out_cap, fr = fl + S.B.out_cap, o = (size_t)p * S.B.out_cap;
  stereo_search_one(i, lane, S.B.kp + fl, reinterpret_cast<const uint4*>(S.B.desc) + 2 * fl, S.B.kp + fr,
                    reinterpret_cast<const uint4*>(S.B.desc) + 2 * fr, nr, S.A.pl.scale, S.n_rows, S.A.min_d, S.A.max_d, S.best_idx + o,
                    S.best_dist + o);
}

__global__ void __launch_bounds__(128) stereo_refine_batch_kernel(const __grid_constant__ StereoBatchArgs S) {
  const int lane = threadIdx.x & 31, p = blockIdx.y;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= S.B.out_cap) return;
  const size_t fl = (size_t)(2 * p) * S.B.out_cap, fr = fl + S.B.out_cap, o = (size_t)p * S.B.out_cap;
  if (i >= min(S.B.count[2 * p], S.B.out_cap)) {  // rows past the left image's keypoints: unmatched (Frame.cc:529-530)
    if (lane == 0) { S.A.u_right[o + i] = -1.0f; S.A.depth[o + i] = -1.0f; S.A.sad[o + i] = -1; }
    return;
  }
  stereo_refine_one(S.A, i, lane, S.B.kp + fl, S.B.kp + fr, S.best_idx + o, S.best_dist + o, S.A.u_right + o, S.A.depth + o, S.A.sad + o,
                    (size_t)(2 * p) * S.B.level0_frame_stride, (size_t)(2 * p) * S.B.level_frame_stride,
                    (size_t)(2 * p + 1) * S.B.level0_frame_stride, (size_t)(2 * p + 1) * S.B.level_frame_stride);
}

// Median rejection (Frame.cc:687-700), one CTA per pair: median = the (n/2)-th smallest patch distance among the n matched
// keypoints (bisection on the value, 0 <= sad <= 121 * 510 < 2^16), matches at or above 1.5 * 1.4 * median are dropped.
__global__ void __launch_bounds__(256) stereo_median_batch_kernel(const __grid_constant__ StereoBatchArgs S) {
  __shared__ int s_cnt;
  const int p = blockIdx.x, tid = threadIdx.x;
  const size_t o = (size_t)p * S.B.out_cap;
  const int nl = min(S.B.count[2 * p], S.B.out_cap);
  const int* sad = S.A.sad + o;
  auto count_le = [&](int v) {  // matched keypoints with sad <= v (v = -1 counts nobody)
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    int c = 0;
    for (int i = tid; i < nl; i += 256) { const int d = sad[i]; c += d >= 0 && d <= v; }
    for (int s = 16; s > 0; s >>= 1) c += __shfl_xor_sync(0xffffffffu, c, s);
    if ((tid & 31) == 0 && c) atomicAdd(&s_cnt, c);
    __syncthreads();
    const int r = s_cnt;
    __syncthreads();
    return r;
  };
  const int n = count_le(0x7fffffff);
  if (n == 0) return;  // the reference indexes an empty vector here
  const int k = n / 2;
  int lo = 0, hi = 65535;  // smallest v with #(sad <= v) >= k + 1
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (count_le(mid) >= k + 1) hi = mid; else lo = mid + 1;
  }
  const float median = (float)lo;
  const float thDist = 1.5f * 1.4f * median;
  for (int i = tid; i < nl; i += 256) {
    const int d = sad[i];
    if (d >= 0 && !((float)d < thDist)) { S.A.u_right[o + i] = -1.0f; S.A.depth[o + i] = -1.0f; }
  }
}

// ------------------------------------------------------------------------------------
// MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:305-370), one warp per map point: for every
// row of the N x N distance matrix the k-th smallest entry (k = (int)(0.5*(N-1)), diagonal 0
// included) by bisection on the value range 0..256 with warp-wide counts; smallest median wins,
// first row on ties.  Rows live in shared memory (N <= 1024).
constexpr int kDistinctWarps = 4, kDistinctMaxN = 1024;

__global__ void __launch_bounds__(kDistinctWarps * 32) distinctive_kernel(const uint4* __restrict__ desc,
                                                                          const int* __restrict__ offset, int n_points,
                                                                          int* __restrict__ best) {
  __shared__ uint16_t row[kDistinctWarps][kDistinctMaxN];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int p = blockIdx.x * kDistinctWarps + w;
  if (p >= n_points) return;
  const int o = offset[p], N = offset[p + 1] - o;
  if (N <= 0) { if (lane == 0) best[p] = -1; return; }
  const int k = (int)(0.5 * (N - 1));
  int best_median = 0x7fffffff, best_idx = 0;
  for (int i = 0; i < N; ++i) {
    const uint4 qa = __ldg(&desc[2 * (o + i)]), qb = __ldg(&desc[2 * (o + i) + 1]);
    const uint32_t q[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
    for (int j = lane; j < N; j += 32) row[w][j] = (uint16_t)hamming8(q, __ldg(&desc[2 * (o + j)]), __ldg(&desc[2 * (o + j) + 1]));
    __syncwarp();
    int lo = 0, hi = 256;  // smallest v with #(d <= v) >= k + 1
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      int c = 0;
      for (int j = lane; j < N; j += 32) c += row[w][j] <= mid;
      for (int s = 16; s > 0; s >>= 1) c += __shfl_xor_sync(0xffffffffu, c, s);
      if (c >= k + 1) hi = mid; else lo = mid + 1;
    }
    if (lo < best_median) { best_median = lo; best_idx = i; }
    __syncwarp();
  }
  if (lane == 0) best[p] = best_idx;
}

}  // namespace orbfront

// ====================================================================================
// C ABI
using namespace orbfront;

namespace {
template <typename T>
struct DevBuf {  // grow-only device array
  T* p = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return ORB_OK;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    ORB_CUDA_TRY(cudaMalloc(&p, n * sizeof(T)));
    cap = n;
    return ORB_OK;
  }
  ~DevBuf() { if (p) cudaFree(p); }
};
}  // namespace

struct orbm_db {
  int device = 0;
  int64_t row_base = 0, n = 0, cap = 0;
  uint4* rows = nullptr;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  DevBuf<uint2> partial;
  DevBuf<uint32_t> dq;
  DevBuf<orbm_top2> dout;
  int last_launches = 0;
  bool timed = false;
};

extern "C" {

int orbm_hamming(const uint8_t* a, const uint8_t* b) {
  int d = 0;
  for (int i = 0; i < 4; ++i) {
    uint64_t x, y;
    memcpy(&x, a + 8 * i, 8);
    memcpy(&y, b + 8 * i, 8);
    d += __builtin_popcountll(x ^ y);
  }
  return d;
}

int orbm_hamming_pairs(int device, const uint8_t* a, const uint8_t* b, int n, int32_t* out) {
  if (n < 0 || (n > 0 && (!a || !b || !out))) return set_error(ORB_E_INVALID, "orbm_hamming_pairs: bad argument");
  if (n == 0) return ORB_OK;
  int rc = use_device(device);
  if (rc) return rc;
  // per-thread scratch: pack -> one upload -> kernel -> one download -> one sync
  Layout L;
  const size_t o_a = L.place((size_t)n * 32), o_b = L.place((size_t)n * 32), o_out = L.place((size_t)n * 4);
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  memcpy(S->host + o_a, a, (size_t)n * 32);
  memcpy(S->host + o_b, b, (size_t)n * 32);
  ORB_CUDA_TRY(cudaMemcpyAsync(S->dev, S->host, o_out, cudaMemcpyHostToDevice, S->stream));
  hamming_pairs_kernel<<<div_up(n, 256), 256, 0, S->stream>>>((const uint4*)(S->dev + o_a), (const uint4*)(S->dev + o_b), n,
                                                             (int*)(S->dev + o_out));
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(S->host + o_out, S->dev + o_out, (size_t)n * 4, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  memcpy(out, S->host + o_out, (size_t)n * 4);
  return ORB_OK;
}

int orbm_db_create(int device, int64_t row_base, orbm_db** out) {
  if (!out || row_base < 0) return set_error(ORB_E_INVALID, "orbm_db_create: bad argument");
  int rc = use_device(device);
  if (rc) return rc;
  orbm_db* db = new orbm_db();
  db->device = device;
  db->row_base = row_base;
  cudaError_t e = cudaStreamCreateWithFlags(&db->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreate(&db->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&db->ev1);
  if (e != cudaSuccess) {
    delete db;
    return set_error(ORB_E_CUDA, "orbm_db_create: %s", cudaGetErrorString(e));
  }
  *out = db;
  return ORB_OK;
}

int orbm_db_append(orbm_db* db, const uint8_t* rows, int64_t n, int memory) {
  if (!db || n < 0 || (n > 0 && !rows)) return set_error(ORB_E_INVALID, "orbm_db_append: bad argument");
  if (n == 0) return ORB_OK;
  int rc = use_device(db->device);
  if (rc) return rc;
  if (db->n + n > db->cap) {
    // grow by half (at least to what is needed, at least 64 K rows = 2 MB): a shard built by many small appends is
    // not reallocated and recopied at every call
    int64_t ncap = std::max<int64_t>(std::max<int64_t>(db->n + n, db->cap + db->cap / 2), 1 << 16);
    uint4* np = nullptr;
    ORB_CUDA_TRY(cudaMalloc(&np, (size_t)ncap * 32));
    cudaError_t ce = db->n ? cudaMemcpyAsync(np, db->rows, (size_t)db->n * 32, cudaMemcpyDeviceToDevice, db->stream) : cudaSuccess;
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(db->stream);
    if (ce != cudaSuccess) {  // the shard keeps its old rows
      cudaFree(np);
      return set_error(ORB_E_CUDA, "orbm_db_append: %s", cudaGetErrorString(ce));
    }
    if (db->rows) cudaFree(db->rows);
    db->rows = np;
    db->cap = ncap;
  }
  ORB_CUDA_TRY(cudaMemcpyAsync(db->rows + db->n * 2, rows, (size_t)n * 32,
                               memory ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, db->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
  db->n += n;
  return ORB_OK;
}

int orbm_db_size(const orbm_db* db, int64_t* n) {
  if (!db || !n) return set_error(ORB_E_INVALID, "orbm_db_size: bad argument");
  *n = db->n;
  return ORB_OK;
}

int orbm_db_destroy(orbm_db* db) {
  if (!db) return ORB_OK;
  cudaSetDevice(db->device);
  if (db->rows) cudaFree(db->rows);
  if (db->ev0) cudaEventDestroy(db->ev0);
  if (db->ev1) cudaEventDestroy(db->ev1);
  if (db->stream) cudaStreamDestroy(db->stream);
  delete db;
  return ORB_OK;
}

int orbm_top2_search(orbm_db* db, const uint8_t* q, int nq, orbm_top2* out, int memory) {
  NvtxRange nvtx_range("orbm_top2_search");
  if (!db || nq < 0 || (nq > 0 && (!q || !out))) return set_error(ORB_E_INVALID, "orbm_top2_search: bad argument");
  if (nq == 0) return ORB_OK;
  int rc = use_device(db->device);
  if (rc) return rc;
  const int nq_pad = align_up(nq, kQPerCta);
  // chunking: ~2 CTAs per SM, whole tiles, and local row indices must fit kRowBits
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, db->device);
  const int qblocks = nq_pad / kQPerCta;
  long long want = std::max(1, (2 * sms) / qblocks);
  // equal shares: the kernel's last tile may be partial, so rows per CTA need not be whole tiles (rounding a 1.25 M-row
  // shard of the 8-GPU case up to 256-row tiles left 8 of the 296 CTA slots empty and 3 % more rows on the others)
  long long rows_per_cta = std::max<long long>(div_up<long long>(std::max<int64_t>(db->n, 1), want), 1);
  rows_per_cta = std::min<long long>(rows_per_cta, 1ll << kRowBits);
  const int n_chunks = (int)std::max<long long>(div_up<long long>(db->n, rows_per_cta), 0);
  const uint32_t* dq = nullptr;
  orbm_top2* dout = nullptr;
  if (memory == 0) {
    if ((rc = db->dq.reserve((size_t)nq * 8)) || (rc = db->dout.reserve(nq))) return rc;
    ORB_CUDA_TRY(cudaMemcpyAsync(db->dq.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice, db->stream));
    dq = db->dq.p;
    dout = db->dout.p;
  } else {
    dq = reinterpret_cast<const uint32_t*>(q);
    dout = out;
  }
  if ((rc = db->partial.reserve((size_t)std::max(n_chunks, 1) * nq_pad))) return rc;
  ORB_CUDA_TRY(cudaEventRecord(db->ev0, db->stream));
  db->last_launches = 0;
  if (n_chunks > 0) {
    top2_scan_kernel<<<dim3(n_chunks, qblocks), kTopThreads, 0, db->stream>>>(db->rows, db->n, (int)rows_per_cta, dq, nq,
                                                                            nq_pad, db->partial.p);
    ORB_LAUNCH_CHECK();
    ++db->last_launches;
  }
  top2_reduce_kernel<<<div_up(nq, 128), 128, 0, db->stream>>>(db->partial.p, n_chunks, (int)rows_per_cta, nq, nq_pad,
                                                               db->row_base, dout);
  ORB_LAUNCH_CHECK();
  ++db->last_launches;
  ORB_CUDA_TRY(cudaEventRecord(db->ev1, db->stream));
  db->timed = true;
  if (memory == 0) {
    ORB_CUDA_TRY(cudaMemcpyAsync(out, dout, (size_t)nq * sizeof(orbm_top2), cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
  }
  return ORB_OK;
}

}  // extern "C"

// ---- allocation-free top-2 search ---------------------------------------------------------------------------
// The per-CTA partial results of the scan live in a file-scope device array instead of a cudaMalloc'd buffer.
// 304 x 1024 entries of 8 bytes = 2 432 KiB on every GPU that loads the library, until exit: two CTAs per SM
// (296 on a B200) x the 1024 queries of a CTA, rounded up -- the most top2_scan_kernel's grid produces here.
namespace {
constexpr size_t kNoAllocEntries = 304u * 1024u;
static __device__ __align__(256) uint2 g_noalloc_partial[kNoAllocEntries];
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
uint2* g_noalloc_ptr[kNoAllocMaxDevices];  // device address of g_noalloc_partial per device (the host-side name is not one)
int g_noalloc_sms[kNoAllocMaxDevices];

// once per device, with that device current
int noalloc_setup(int device, uint2** ptr, int* sms) {
  if (device < 0 || device >= kNoAllocMaxDevices) return set_error(ORB_E_INVALID, "orbm_top2_search_noalloc: device %d out of range", device);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_partial);
    if (e != cudaSuccess || !p) {
      return set_error(ORB_E_CUDA, "orbm_top2_search_noalloc: cudaGetSymbolAddress failed on device %d: %s", device,
                       cudaGetErrorString(e));
    }
    int n = 0;
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
    if (e != cudaSuccess || n < 1) return set_error(ORB_E_CUDA, "orbm_top2_search_noalloc: SM count: %s", cudaGetErrorString(e));
    g_noalloc_sms[device] = n;
    g_noalloc_ptr[device] = static_cast<uint2*>(p);
  }
  *ptr = g_noalloc_ptr[device];
  *sms = g_noalloc_sms[device];
  return ORB_OK;
}
}  // namespace

extern "C" {

int orbm_top2_search_noalloc(int device, const uint8_t* rows, int64_t n_rows, int64_t row_base, const uint8_t* q, int nq,
                             orbm_top2* out, void* cuda_stream) {
  if (n_rows < 0 || row_base < 0 || nq < 0 || (n_rows > 0 && !rows) || (nq > 0 && (!q || !out)))
    return set_error(ORB_E_INVALID, "orbm_top2_search_noalloc: bad argument");
  if ((reinterpret_cast<uintptr_t>(rows) & 15) || (reinterpret_cast<uintptr_t>(q) & 3) || (reinterpret_cast<uintptr_t>(out) & 7))
    return set_error(ORB_E_INVALID, "orbm_top2_search_noalloc: rows must be 16-byte, q 4-byte and out 8-byte aligned");
  if (nq == 0) return ORB_OK;
  int rc = use_device(device);
  if (rc) return rc;
  uint2* partial = nullptr;
  int sms = 0;
  if ((rc = noalloc_setup(device, &partial, &sms))) return rc;
  // the same chunking as orbm_top2_search
  const int nq_pad = align_up(nq, kQPerCta);
  const int qblocks = nq_pad / kQPerCta;
  const long long want = std::max(1, (2 * sms) / qblocks);
  long long rows_per_cta = std::max<long long>(div_up<long long>(std::max<int64_t>(n_rows, 1), want), 1);
  rows_per_cta = std::min<long long>(rows_per_cta, 1ll << kRowBits);
  const long long n_chunks = div_up<long long>(n_rows, rows_per_cta);
  if ((unsigned long long)std::max<long long>(n_chunks, 1) * (unsigned long long)nq_pad > kNoAllocEntries || n_chunks > 0x7fffffffll) {
    return set_error(ORB_E_INVALID,
                     "orbm_top2_search_noalloc: %d queries x %lld row chunks need %lld scratch entries, the static scratch holds %zu "
                     "(up to %zu queries against up to %lld rows per call on this GPU)",
                     nq, n_chunks, std::max<long long>(n_chunks, 1) * nq_pad, kNoAllocEntries, kNoAllocEntries,
                     (long long)want << kRowBits);
  }
  cudaStream_t s = static_cast<cudaStream_t>(cuda_stream);
  if (n_chunks > 0) {
    top2_scan_kernel<<<dim3((unsigned)n_chunks, qblocks), kTopThreads, 0, s>>>(reinterpret_cast<const uint4*>(rows), n_rows,
                                                                              (int)rows_per_cta, reinterpret_cast<const uint32_t*>(q),
                                                                              nq, nq_pad, partial);
    ORB_LAUNCH_CHECK();
  }
  // every entry the fold reads was written by the scan of this call: nothing in the scratch has to start at zero
  top2_reduce_kernel<<<div_up(nq, 128), 128, 0, s>>>(partial, (int)n_chunks, (int)rows_per_cta, nq, nq_pad, row_base, out);
  ORB_LAUNCH_CHECK();
  return ORB_OK;
}

int orbm_db_sync(orbm_db* db) {
  if (!db) return set_error(ORB_E_INVALID, "orbm_db_sync: null");
  int rc = use_device(db->device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
  return ORB_OK;
}

int orbm_db_last_ms(orbm_db* db, float* ms, int* launches) {
  if (!db || !db->timed) return set_error(ORB_E_STATE, "orbm_db_last_ms: no search has run");
  int rc = use_device(db->device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaEventSynchronize(db->ev1));
  if (ms) ORB_CUDA_TRY(cudaEventElapsedTime(ms, db->ev0, db->ev1));
  if (launches) *launches = db->last_launches;
  return ORB_OK;
}

int orbm_db_get_stream(orbm_db* db, void** cuda_stream) {
  if (!db || !cuda_stream) return set_error(ORB_E_INVALID, "orbm_db_get_stream: bad argument");
  *cuda_stream = (void*)db->stream;
  return ORB_OK;
}

int orbm_merge_top2(const orbm_top2* shards, int n_shards, int nq, float ratio, int th, orbm_top2* out,
                    uint8_t* accept) {
  if (!shards || !out || n_shards < 1 || nq < 0) return set_error(ORB_E_INVALID, "orbm_merge_top2: bad argument");
  for (int q = 0; q < nq; ++q) {
    int b1 = 256, b2 = 256;
    int64_t idx = -1;
    for (int s = 0; s < n_shards; ++s) {  // shards in increasing row order: strict < keeps the lowest row
      const orbm_top2& p = shards[(size_t)s * nq + q];
      if (p.idx >= 0 && p.best1 < b1) { b2 = b1; b1 = p.best1; idx = p.idx; }
      else if (p.idx >= 0 && p.best1 < b2) b2 = p.best1;
      if (p.best2 < b2) b2 = p.best2;
    }
    out[q].best1 = b1; out[q].best2 = b2; out[q].idx = idx;
    if (accept) accept[q] = (b1 <= th) && ((float)b1 < ratio * (float)b2);
  }
  return ORB_OK;
}

int orbm_popc_peak(int device, double* popc_per_s) {
  if (!popc_per_s) return set_error(ORB_E_INVALID, "orbm_popc_peak: null");
  int rc = use_device(device);
  if (rc) return rc;
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const int blocks = sms * 8, iters = 4096;
  DevBuf<uint32_t> out;
  if ((rc = out.reserve((size_t)blocks * 256))) return rc;
  cudaEvent_t e0, e1;
  ORB_CUDA_TRY(cudaEventCreate(&e0));
  ORB_CUDA_TRY(cudaEventCreate(&e1));
  double best = 0;
  for (int rep = 0; rep < 6; ++rep) {
    ORB_CUDA_TRY(cudaEventRecord(e0));
    popc_probe_kernel<<<blocks, 256>>>(rep * 7919u + 1u, iters, out.p);
    ORB_LAUNCH_CHECK();
    ORB_CUDA_TRY(cudaEventRecord(e1));
    ORB_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    ORB_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    const double rate = (double)blocks * 256.0 * 8.0 * iters / (ms * 1e-3);
    if (rep >= 2 && rate > best) best = rate;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *popc_per_s = best;
  return ORB_OK;
}


int orbm_search_bow(int device, const uint8_t* desc1, int n1, const uint8_t* desc2, int n2,
                    const orbm_featvec* fv1, const orbm_featvec* fv2, const uint8_t* valid1,
                    const uint8_t* valid2, const float* angle1, const float* angle2, float nnratio,
                    int check_orientation, int kf_kf, int32_t* out, int* n_matches) {
  if (n1 < 0 || n2 < 0 || !fv1 || !fv2 || !out || !n_matches || (n1 > 0 && (!desc1 || !valid1)) ||
      (n2 > 0 && !desc2) || (check_orientation && (!angle1 || !angle2)) || (kf_kf && !valid2))
    return set_error(ORB_E_INVALID, "orbm_search_bow: bad argument");
  const int n_out = kf_kf ? n1 : n2;
  *n_matches = 0;
  for (int i = 0; i < n_out; ++i) out[i] = -1;
  if (n1 == 0 || n2 == 0 || fv1->n_nodes == 0 || fv2->n_nodes == 0) return ORB_OK;
  const int ni1 = fv1->off[fv1->n_nodes], ni2 = fv2->off[fv2->n_nodes];
  int rc = use_device(device);
  if (rc) return rc;
  Layout L;
  const int max_pairs = std::min(fv1->n_nodes, fv2->n_nodes);
  // upload block
  const size_t o_d1 = L.place((size_t)n1 * 32), o_d2 = L.place((size_t)n2 * 32),
               o_pairs = L.place((size_t)max_pairs * sizeof(BowPair)), o_i1 = L.place((size_t)ni1 * 4),
               o_i2 = L.place((size_t)ni2 * 4), o_v1 = L.place(n1), o_v2 = L.place(n2), o_a1 = L.place((size_t)n1 * 4),
               o_a2 = L.place((size_t)n2 * 4), o_claim = L.place(n2), o_hist = L.place(ORBM_HISTO_LENGTH * 4),
               o_bin = L.place(n_out);
  // download block (initialised on the host, so it is uploaded too)
  const size_t o_out = L.place((size_t)n_out * 4), o_nm = L.place(4), o_end = L.bytes;
  // host-only flags for the one-node-per-feature check
  const size_t o_seen = L.place((size_t)std::max(n1, n2));
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  uint8_t* h = S->host;

  // ordered-map merge of the two node lists (ORBmatcher.cc:381-465) -> common-node ranges
  BowPair* pairs = reinterpret_cast<BowPair*>(h + o_pairs);
  int n_pairs = 0;
  for (int a = 0, b = 0; a < fv1->n_nodes && b < fv2->n_nodes;) {
    if (fv1->node[a] == fv2->node[b]) {
      const BowPair p{fv1->off[a], fv1->off[a + 1] - fv1->off[a], fv2->off[b], fv2->off[b + 1] - fv2->off[b]};
      if (p.l2 > 65535) return set_error(ORB_E_INVALID, "orbm_search_bow: node with more than 65535 features");
      pairs[n_pairs++] = p;
      ++a; ++b;
    } else if (fv1->node[a] < fv2->node[b]) {
      a = (int)(std::lower_bound(fv1->node + a, fv1->node + fv1->n_nodes, fv2->node[b]) - fv1->node);
    } else {
      b = (int)(std::lower_bound(fv2->node + b, fv2->node + fv2->n_nodes, fv1->node[a]) - fv2->node);
    }
  }
  if (n_pairs == 0) return ORB_OK;
  for (int i = 0; i < ni1; ++i)
    if (fv1->idx[i] >= (uint32_t)n1) return set_error(ORB_E_INVALID, "orbm_search_bow: feature index out of range");
  for (int i = 0; i < ni2; ++i)
    if (fv2->idx[i] >= (uint32_t)n2) return set_error(ORB_E_INVALID, "orbm_search_bow: feature index out of range");
  // Nodes may run in parallel only if no feature sits in two common nodes (always true for a DBoW2
  // FeatureVector; checked because the ABI takes arbitrary lists).
  bool disjoint = true;
  for (int side = 0; side < 2 && disjoint; ++side) {
    const int n = side ? n2 : n1;
    const uint32_t* idx = side ? fv2->idx : fv1->idx;
    uint8_t* seen = h + o_seen;
    memset(seen, 0, (size_t)n);
    for (int p = 0; p < n_pairs && disjoint; ++p) {
      const int o = side ? pairs[p].o2 : pairs[p].o1, l = side ? pairs[p].l2 : pairs[p].l1;
      for (int i = 0; i < l; ++i) {
        uint8_t& f = seen[idx[o + i]];
        if (f) { disjoint = false; break; }
        f = 1;
      }
    }
  }

  memcpy(h + o_d1, desc1, (size_t)n1 * 32);
  memcpy(h + o_d2, desc2, (size_t)n2 * 32);
  memcpy(h + o_i1, fv1->idx, (size_t)ni1 * 4);
  memcpy(h + o_i2, fv2->idx, (size_t)ni2 * 4);
  memcpy(h + o_v1, valid1, n1);
  if (valid2) memcpy(h + o_v2, valid2, n2);
  if (check_orientation) {
    memcpy(h + o_a1, angle1, (size_t)n1 * 4);
    memcpy(h + o_a2, angle2, (size_t)n2 * 4);
  }
  memset(h + o_claim, 0, (size_t)n2);
  memset(h + o_hist, 0, ORBM_HISTO_LENGTH * 4);
  memset(h + o_out, 0xff, (size_t)n_out * 4);  // -1
  memset(h + o_nm, 0, 4);
  uint8_t* d = S->dev;
  ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_end, cudaMemcpyHostToDevice, S->stream));
  BowArgs A;
  A.desc1 = (const uint4*)(d + o_d1); A.desc2 = (const uint4*)(d + o_d2);
  A.pairs = (const BowPair*)(d + o_pairs); A.n_pairs = n_pairs;
  A.idx1 = (const uint32_t*)(d + o_i1); A.idx2 = (const uint32_t*)(d + o_i2);
  A.valid1 = d + o_v1; A.valid2 = valid2 ? d + o_v2 : nullptr;
  A.ang1 = (const float*)(d + o_a1); A.ang2 = (const float*)(d + o_a2);
  A.ratio = nnratio; A.check_ori = check_orientation; A.kf_kf = kf_kf; A.n_out = n_out;
  A.claimed2 = d + o_claim; A.out = (int*)(d + o_out); A.bin_of = d + o_bin;
  A.hist = (int*)(d + o_hist); A.n_matches = (int*)(d + o_nm);
  if (disjoint) bow_node_kernel<<<div_up(n_pairs, 4), 128, 0, S->stream>>>(A);
  else bow_node_kernel<<<1, 32, 0, S->stream>>>(A);  // one warp walks the nodes in order
  ORB_LAUNCH_CHECK();
  if (check_orientation) {
    bow_finish_kernel<<<1, 32, 0, S->stream>>>(A.hist, A.out, A.bin_of, n_out, A.n_matches);
    ORB_LAUNCH_CHECK();
  }
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_out, d + o_out, o_end - o_out, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  memcpy(out, h + o_out, (size_t)n_out * 4);
  memcpy(n_matches, h + o_nm, 4);
  return ORB_OK;
}

// K independent searches against one shared side 2 (the current Frame; a KeyFrame for kf_kf = 1) in one upload, two launches
// (grid.y = K) and one download.  Each search has its own claim mask, histogram and output, exactly like K calls of
// orbm_search_bow -- which a problem falls back to if one of its features sits in two common nodes.
int orbm_search_bow_batch(int device, const orbm_bow_side* side1, int k, const orbm_bow_side* side2, float nnratio,
                          int check_orientation, int kf_kf, int32_t* const* out, int* n_matches) {
  if (k < 0 || !side2 || (k > 0 && (!side1 || !out || !n_matches))) return set_error(ORB_E_INVALID, "orbm_search_bow_batch: bad argument");
  if (k == 0) return ORB_OK;
  const int n2 = side2->n;
  if (n2 < 0 || !side2->fv || (n2 > 0 && !side2->desc) || (check_orientation && n2 > 0 && !side2->angle) || (kf_kf && n2 > 0 && !side2->valid))
    return set_error(ORB_E_INVALID, "orbm_search_bow_batch: bad side 2");
  const orbm_featvec* fv2 = side2->fv;
  const int ni2 = fv2->n_nodes ? fv2->off[fv2->n_nodes] : 0;
  for (int i = 0; i < ni2; ++i)
    if (fv2->idx[i] >= (uint32_t)n2) return set_error(ORB_E_INVALID, "orbm_search_bow_batch: feature index out of range");
  int rc = use_device(device);
  if (rc) return rc;
  // ---- pass 1: per problem, the common-node ranges (ordered-map merge, ORBmatcher.cc:381-465) and the layout
  struct Prob { int n1, ni1, n_out, n_pairs; size_t o_d1, o_pairs, o_i1, o_v1, o_a1, o_claim, o_hist, o_bin, o_out, o_nm; bool serial; };
  std::vector<Prob> pr((size_t)k);
  std::vector<std::vector<BowPair> > pairs((size_t)k);
  Layout L;
  const size_t o_args = L.place((size_t)k * sizeof(BowArgs));
  const size_t o_d2 = L.place((size_t)n2 * 32), o_i2 = L.place((size_t)ni2 * 4), o_v2 = L.place(n2), o_a2 = L.place((size_t)n2 * 4);
  std::vector<uint8_t> seen;
  for (int p = 0; p < k; ++p) {
    const orbm_bow_side& s1 = side1[p];
    Prob& q = pr[p];
    if (s1.n < 0 || !s1.fv || (!out[p] && (kf_kf ? s1.n : n2) > 0) || (s1.n > 0 && (!s1.desc || !s1.valid)) ||
        (check_orientation && s1.n > 0 && !s1.angle))
      return set_error(ORB_E_INVALID, "orbm_search_bow_batch: bad side 1 of problem %d", p);
    q.n1 = s1.n; q.n_out = kf_kf ? s1.n : n2; q.n_pairs = 0; q.serial = false;
    q.ni1 = s1.fv->n_nodes ? s1.fv->off[s1.fv->n_nodes] : 0;
    n_matches[p] = 0;
    for (int i = 0; i < q.n_out; ++i) out[p][i] = -1;
    if (q.n1 == 0 || n2 == 0) continue;
    const orbm_featvec* fv1 = s1.fv;
    for (int i = 0; i < q.ni1; ++i)
      if (fv1->idx[i] >= (uint32_t)q.n1) return set_error(ORB_E_INVALID, "orbm_search_bow_batch: feature index out of range");
    for (int a = 0, b = 0; a < fv1->n_nodes && b < fv2->n_nodes;) {
      if (fv1->node[a] == fv2->node[b]) {
        const BowPair bp{fv1->off[a], fv1->off[a + 1] - fv1->off[a], fv2->off[b], fv2->off[b + 1] - fv2->off[b]};
        if (bp.l2 > 65535) return set_error(ORB_E_INVALID, "orbm_search_bow_batch: node with more than 65535 features");
        pairs[p].push_back(bp);
        ++a; ++b;
      } else if (fv1->node[a] < fv2->node[b]) {
        a = (int)(std::lower_bound(fv1->node + a, fv1->node + fv1->n_nodes, fv2->node[b]) - fv1->node);
      } else {
        b = (int)(std::lower_bound(fv2->node + b, fv2->node + fv2->n_nodes, fv1->node[a]) - fv2->node);
      }
    }
    // nodes run in parallel only if no feature sits in two common nodes (always true for a DBoW2 FeatureVector)
    for (int side = 0; side < 2 && !q.serial; ++side) {
      const int n = side ? n2 : q.n1;
      const uint32_t* idx = side ? fv2->idx : fv1->idx;
      seen.assign((size_t)n, 0);
      for (const BowPair& bp : pairs[p]) {
        const int o = side ? bp.o2 : bp.o1, l = side ? bp.l2 : bp.l1;
        for (int i = 0; i < l && !q.serial; ++i) {
          uint8_t& f = seen[idx[o + i]];
          if (f) q.serial = true;
          f = 1;
        }
      }
    }
    if (q.serial || pairs[p].empty()) continue;
    q.n_pairs = (int)pairs[p].size();
    q.o_d1 = L.place((size_t)q.n1 * 32); q.o_pairs = L.place((size_t)q.n_pairs * sizeof(BowPair)); q.o_i1 = L.place((size_t)q.ni1 * 4);
    q.o_v1 = L.place(q.n1); q.o_a1 = L.place((size_t)q.n1 * 4); q.o_claim = L.place(n2); q.o_hist = L.place(ORBM_HISTO_LENGTH * 4);
    q.o_bin = L.place(q.n_out);
  }
  const size_t o_up_end = L.bytes;
  for (int p = 0; p < k; ++p)
    if (pr[p].n_pairs) { pr[p].o_out = L.place((size_t)pr[p].n_out * 4); pr[p].o_nm = L.place(4); }
  const size_t o_end = L.bytes;
  bool any = false;
  for (int p = 0; p < k; ++p) any = any || pr[p].n_pairs > 0;
  if (any) {
    CallScratch* S = nullptr;
    if ((rc = call_scratch(device, L.bytes, &S))) return rc;
    uint8_t *h = S->host, *d = S->dev;
    memcpy(h + o_d2, side2->desc, (size_t)n2 * 32);
    memcpy(h + o_i2, fv2->idx, (size_t)ni2 * 4);
    if (side2->valid) memcpy(h + o_v2, side2->valid, n2);
    if (check_orientation) memcpy(h + o_a2, side2->angle, (size_t)n2 * 4);
    BowArgs* args = reinterpret_cast<BowArgs*>(h + o_args);
    int max_pairs = 0;
    for (int p = 0; p < k; ++p) {
      const Prob& q = pr[p];
      BowArgs& A = args[p];
      memset(&A, 0, sizeof(A));
      if (!q.n_pairs) continue;
      const orbm_bow_side& s1 = side1[p];
      memcpy(h + q.o_d1, s1.desc, (size_t)q.n1 * 32);
      memcpy(h + q.o_pairs, pairs[p].data(), (size_t)q.n_pairs * sizeof(BowPair));
      memcpy(h + q.o_i1, s1.fv->idx, (size_t)q.ni1 * 4);
      memcpy(h + q.o_v1, s1.valid, q.n1);
      if (check_orientation) memcpy(h + q.o_a1, s1.angle, (size_t)q.n1 * 4);
      memset(h + q.o_claim, 0, (size_t)n2);
      memset(h + q.o_hist, 0, ORBM_HISTO_LENGTH * 4);
      memset(h + q.o_out, 0xff, (size_t)q.n_out * 4);  // -1
      memset(h + q.o_nm, 0, 4);
      A.desc1 = (const uint4*)(d + q.o_d1); A.desc2 = (const uint4*)(d + o_d2);
      A.pairs = (const BowPair*)(d + q.o_pairs); A.n_pairs = q.n_pairs;
      A.idx1 = (const uint32_t*)(d + q.o_i1); A.idx2 = (const uint32_t*)(d + o_i2);
      A.valid1 = d + q.o_v1; A.valid2 = side2->valid ? d + o_v2 : nullptr;
      A.ang1 = (const float*)(d + q.o_a1); A.ang2 = (const float*)(d + o_a2);
      A.ratio = nnratio; A.check_ori = check_orientation; A.kf_kf = kf_kf; A.n_out = q.n_out;
      A.claimed2 = d + q.o_claim; A.out = (int*)(d + q.o_out); A.bin_of = d + q.o_bin;
      A.hist = (int*)(d + q.o_hist); A.n_matches = (int*)(d + q.o_nm);
      max_pairs = std::max(max_pairs, q.n_pairs);
    }
    // the download block is initialised on the host (-1 / 0), so it is uploaded as well
    ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_end, cudaMemcpyHostToDevice, S->stream));
    bow_node_batch_kernel<<<dim3(div_up(max_pairs, 4), k), 128, 0, S->stream>>>((const BowArgs*)(d + o_args));
    ORB_LAUNCH_CHECK();
    if (check_orientation) {
      bow_finish_batch_kernel<<<k, 32, 0, S->stream>>>((const BowArgs*)(d + o_args));
      ORB_LAUNCH_CHECK();
    }
    ORB_CUDA_TRY(cudaMemcpyAsync(h + o_up_end, d + o_up_end, o_end - o_up_end, cudaMemcpyDeviceToHost, S->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
    for (int p = 0; p < k; ++p)
      if (pr[p].n_pairs) {
        memcpy(out[p], h + pr[p].o_out, (size_t)pr[p].n_out * 4);
        memcpy(&n_matches[p], h + pr[p].o_nm, 4);
      }
  }
  for (int p = 0; p < k; ++p)
    if (pr[p].serial) {
      const orbm_bow_side& s1 = side1[p];
      if ((rc = orbm_search_bow(device, s1.desc, s1.n, side2->desc, n2, s1.fv, fv2, s1.valid, side2->valid, s1.angle, side2->angle, nnratio,
                                check_orientation, kf_kf, out[p], &n_matches[p])))
        return rc;
    }
  return ORB_OK;
}

int orbm_stereo_search(int device, const orbx_keypoint* left, const uint8_t* ldesc, int nl,
                       const orbx_keypoint* right, const uint8_t* rdesc, int nr,
                       const float* scale_factors, int n_levels, int n_rows, float min_d, float max_d,
                       int32_t* best_idx, int32_t* best_dist) {
  if (nl < 0 || nr < 0 || n_levels < 1 || !scale_factors || (nl > 0 && (!left || !ldesc || !best_idx || !best_dist)) ||
      (nr > 0 && (!right || !rdesc)) || nr >= (1 << kRowBits))
    return set_error(ORB_E_INVALID, "orbm_stereo_search: bad argument");
  if (nl == 0) return ORB_OK;
  for (int i = 0; i < nr; ++i)
    if (right[i].octave < 0 || right[i].octave >= n_levels)
      return set_error(ORB_E_INVALID, "orbm_stereo_search: right keypoint %d has octave %d", i, right[i].octave);
  int rc = use_device(device);
  if (rc) return rc;
  Layout L;
  const size_t o_l = L.place((size_t)nl * sizeof(orbx_keypoint)), o_r = L.place((size_t)nr * sizeof(orbx_keypoint)),
               o_dl = L.place((size_t)nl * 32), o_dr = L.place((size_t)nr * 32), o_s = L.place((size_t)n_levels * 4),
               o_res = L.place((size_t)nl * 8);
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  uint8_t *h = S->host, *d = S->dev;
  memcpy(h + o_l, left, (size_t)nl * sizeof(orbx_keypoint));
  memcpy(h + o_dl, ldesc, (size_t)nl * 32);
  if (nr) {
    memcpy(h + o_r, right, (size_t)nr * sizeof(orbx_keypoint));
    memcpy(h + o_dr, rdesc, (size_t)nr * 32);
  }
  memcpy(h + o_s, scale_factors, (size_t)n_levels * 4);
  ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_res, cudaMemcpyHostToDevice, S->stream));
  int* res = (int*)(d + o_res);
  stereo_search_kernel<<<div_up(nl, 4), 128, 0, S->stream>>>((const orbx_keypoint*)(d + o_l), (const uint4*)(d + o_dl), nl,
                                                             (const orbx_keypoint*)(d + o_r), (const uint4*)(d + o_dr), nr,
                                                             (const float*)(d + o_s), n_rows, min_d, max_d, res, res + nl);
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_res, res, (size_t)nl * 8, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  memcpy(best_idx, h + o_res, (size_t)nl * 4);
  memcpy(best_dist, h + o_res + (size_t)nl * 4, (size_t)nl * 4);
  return ORB_OK;
}

int orbm_search_triangulation(int device, const orbm_tri_keyframe* kf1, const orbm_tri_keyframe* kf2, const float* F12,
                              float ex, float ey, const float* scale_factors2, const float* level_sigma2_2, int n_levels,
                              int check_orientation, int32_t* matches12, int* n_matches) {
  if (!kf1 || !kf2 || !F12 || !scale_factors2 || !level_sigma2_2 || n_levels < 1 || !matches12 || !n_matches || kf1->n < 0 ||
      kf2->n < 0 || !kf1->fv || !kf2->fv)
    return set_error(ORB_E_INVALID, "orbm_search_triangulation: bad argument");
  for (const orbm_tri_keyframe* k : {kf1, kf2})
    if (k->n > 0 && (!k->desc || !k->x || !k->y || !k->octave || !k->usable || !k->stereo || (check_orientation && !k->angle)))
      return set_error(ORB_E_INVALID, "orbm_search_triangulation: missing keyframe array");
  const int n1 = kf1->n, n2 = kf2->n;
  *n_matches = 0;
  for (int i = 0; i < n1; ++i) matches12[i] = -1;
  const orbm_featvec *fv1 = kf1->fv, *fv2 = kf2->fv;
  if (n1 == 0 || n2 == 0 || fv1->n_nodes == 0 || fv2->n_nodes == 0) return ORB_OK;
  const int ni1 = fv1->off[fv1->n_nodes], ni2 = fv2->off[fv2->n_nodes];
  for (int i = 0; i < ni1; ++i)
    if (fv1->idx[i] >= (uint32_t)n1) return set_error(ORB_E_INVALID, "orbm_search_triangulation: feature index out of range");
  for (int i = 0; i < ni2; ++i)
    if (fv2->idx[i] >= (uint32_t)n2) return set_error(ORB_E_INVALID, "orbm_search_triangulation: feature index out of range");
  for (int i = 0; i < n2; ++i)
    if (kf2->octave[i] < 0 || kf2->octave[i] >= n_levels)
      return set_error(ORB_E_INVALID, "orbm_search_triangulation: keypoint %d of keyframe 2 has octave %d", i, kf2->octave[i]);
  int rc = use_device(device);
  if (rc) return rc;
  Layout L;
  const size_t o_d1 = L.place((size_t)n1 * 32), o_d2 = L.place((size_t)n2 * 32), o_items = L.place((size_t)ni1 * sizeof(TriItem)),
               o_i2 = L.place((size_t)ni2 * 4), o_xy1 = L.place((size_t)n1 * 8), o_xy2 = L.place((size_t)n2 * 8),
               o_oct2 = L.place((size_t)n2 * 4), o_us2 = L.place(n2), o_st1 = L.place(n1), o_st2 = L.place(n2),
               o_sf = L.place((size_t)n_levels * 4), o_sg = L.place((size_t)n_levels * 4);
  const size_t o_up_end = L.bytes;
  const size_t o_best = L.place((size_t)ni1 * 4);
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  uint8_t *h = S->host, *d = S->dev;
  // items: usable side-1 keypoints of the common nodes, in the reference's scan order (:888-978)
  TriItem* items = reinterpret_cast<TriItem*>(h + o_items);
  int n_items = 0;
  for (int a = 0, b = 0; a < fv1->n_nodes && b < fv2->n_nodes;) {
    if (fv1->node[a] == fv2->node[b]) {
      const int l2 = fv2->off[b + 1] - fv2->off[b];
      if (l2 > 65535) return set_error(ORB_E_INVALID, "orbm_search_triangulation: node with more than 65535 features");
      for (int p = fv1->off[a]; p < fv1->off[a + 1]; ++p) {
        const int i1 = (int)fv1->idx[p];
        if (kf1->usable[i1]) items[n_items++] = TriItem{i1, fv2->off[b], l2};
      }
      ++a; ++b;
    } else if (fv1->node[a] < fv2->node[b]) {
      a = (int)(std::lower_bound(fv1->node + a, fv1->node + fv1->n_nodes, fv2->node[b]) - fv1->node);
    } else {
      b = (int)(std::lower_bound(fv2->node + b, fv2->node + fv2->n_nodes, fv1->node[a]) - fv2->node);
    }
  }
  if (n_items == 0) return ORB_OK;
  memcpy(h + o_d1, kf1->desc, (size_t)n1 * 32);
  memcpy(h + o_d2, kf2->desc, (size_t)n2 * 32);
  memcpy(h + o_i2, fv2->idx, (size_t)ni2 * 4);
  float* xy1 = reinterpret_cast<float*>(h + o_xy1);
  for (int i = 0; i < n1; ++i) { xy1[2 * i] = kf1->x[i]; xy1[2 * i + 1] = kf1->y[i]; }
  float* xy2 = reinterpret_cast<float*>(h + o_xy2);
  for (int i = 0; i < n2; ++i) { xy2[2 * i] = kf2->x[i]; xy2[2 * i + 1] = kf2->y[i]; }
  memcpy(h + o_oct2, kf2->octave, (size_t)n2 * 4);
  memcpy(h + o_us2, kf2->usable, n2);
  memcpy(h + o_st1, kf1->stereo, n1);
  memcpy(h + o_st2, kf2->stereo, n2);
  memcpy(h + o_sf, scale_factors2, (size_t)n_levels * 4);
  memcpy(h + o_sg, level_sigma2_2, (size_t)n_levels * 4);
  ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_up_end, cudaMemcpyHostToDevice, S->stream));
  TriArgs A;
  A.desc1 = (const uint4*)(d + o_d1); A.desc2 = (const uint4*)(d + o_d2);
  A.items = (const TriItem*)(d + o_items); A.n_items = n_items;
  A.idx2 = (const uint32_t*)(d + o_i2);
  A.xy1 = (const float2*)(d + o_xy1); A.xy2 = (const float2*)(d + o_xy2); A.oct2 = (const int*)(d + o_oct2);
  A.usable2 = d + o_us2; A.stereo1 = d + o_st1; A.stereo2 = d + o_st2;
  for (int i = 0; i < 9; ++i) A.F[i] = F12[i];
  A.ex = ex; A.ey = ey;
  A.sf2 = (const float*)(d + o_sf); A.sigma2 = (const float*)(d + o_sg);
  A.best = (int*)(d + o_best);
  triangulation_kernel<<<div_up(n_items, 4), 128, 0, S->stream>>>(A);
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_best, d + o_best, (size_t)n_items * 4, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  // apply in scan order, then the rotation filter (:980-1008)
  const int* best = reinterpret_cast<const int*>(h + o_best);
  std::vector<int> rot_hist[ORBM_HISTO_LENGTH];
  int nm = 0;
  for (int k = 0; k < n_items; ++k) {
    if (best[k] < 0) continue;
    const int i1 = items[k].i1;
    matches12[i1] = best[k];
    ++nm;
    if (check_orientation) {
      float rot = kf1->angle[i1] - kf2->angle[best[k]];
      if (rot < 0.0f) rot += 360.0f;
      int bin = (int)roundf(rot * (1.0f / ORBM_HISTO_LENGTH));
      if (bin == ORBM_HISTO_LENGTH) bin = 0;
      rot_hist[bin].push_back(i1);
    }
  }
  if (check_orientation) {
    int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
    for (int i = 0; i < ORBM_HISTO_LENGTH; ++i) {
      const int sz = (int)rot_hist[i].size();
      if (sz > max1) { max3 = max2; max2 = max1; max1 = sz; ind3 = ind2; ind2 = ind1; ind1 = i; }
      else if (sz > max2) { max3 = max2; max2 = sz; ind3 = ind2; ind2 = i; }
      else if (sz > max3) { max3 = sz; ind3 = i; }
    }
    if ((float)max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < 0.1f * (float)max1) ind3 = -1;
    for (int i = 0; i < ORBM_HISTO_LENGTH; ++i) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int i1 : rot_hist[i]) { matches12[i1] = -1; --nm; }
    }
  }
  *n_matches = nm;
  return ORB_OK;
}

int orbx_stereo_matches(orbx_handle* left, int frame_left, orbx_handle* right, int frame_right, const orbx_keypoint* kl,
                        const uint8_t* dl, int nl, const orbx_keypoint* kr, const uint8_t* dr, int nr, float mb, float mbf,
                        float* u_right, float* depth) {
  if (!left || !right || nl < 0 || nr < 0 || (nl > 0 && (!kl || !dl || !u_right || !depth)) || (nr > 0 && (!kr || !dr)) ||
      nr >= (1 << kRowBits))
    return set_error(ORB_E_INVALID, "orbx_stereo_matches: bad argument");
  for (int i = 0; i < nl; ++i) { u_right[i] = -1.0f; depth[i] = -1.0f; }  // Frame.cc:529-530
  if (nl == 0) return ORB_OK;
  StereoRefineArgs A;
  int rc;
  if ((rc = pyramid_view(left, frame_left, &A.pl)) || (rc = pyramid_view(right, frame_right, &A.pr))) return rc;
  if (A.pl.device != A.pr.device || A.pl.n_levels != A.pr.n_levels)
    return set_error(ORB_E_INVALID, "orbx_stereo_matches: the two extractors differ in device or level count");
  const int n_levels = A.pl.n_levels, device = A.pl.device;
  for (int l = 0; l < n_levels; ++l)
    if (A.pl.h[l] != A.pr.h[l]) return set_error(ORB_E_INVALID, "orbx_stereo_matches: left and right level %d differ in height", l);
  for (int i = 0; i < nl; ++i)
    if (kl[i].octave < 0 || kl[i].octave >= n_levels) return set_error(ORB_E_INVALID, "orbx_stereo_matches: left keypoint %d has octave %d", i, kl[i].octave);
  for (int i = 0; i < nr; ++i)
    if (kr[i].octave < 0 || kr[i].octave >= n_levels) return set_error(ORB_E_INVALID, "orbx_stereo_matches: right keypoint %d has octave %d", i, kr[i].octave);
  const int n_rows = A.pl.h[0];  // :535
  const float min_d = 0, max_d = mbf / mb;  // :557-559
  Layout L;
  const size_t o_l = L.place((size_t)nl * sizeof(orbx_keypoint)), o_r = L.place((size_t)nr * sizeof(orbx_keypoint)),
               o_dl = L.place((size_t)nl * 32), o_dr = L.place((size_t)nr * 32), o_s = L.place((size_t)n_levels * 4);
  const size_t o_up_end = L.bytes;
  const size_t o_best = L.place((size_t)nl * 8);
  const size_t o_res = L.place((size_t)nl * 12), o_end = L.bytes;
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  uint8_t *h = S->host, *d = S->dev;
  memcpy(h + o_l, kl, (size_t)nl * sizeof(orbx_keypoint));
  memcpy(h + o_dl, dl, (size_t)nl * 32);
  if (nr) {
    memcpy(h + o_r, kr, (size_t)nr * sizeof(orbx_keypoint));
    memcpy(h + o_dr, dr, (size_t)nr * 32);
  }
  memcpy(h + o_s, A.pl.scale, (size_t)n_levels * 4);
  ORB_CUDA_TRY(cudaMemcpyAsync(d, h, o_up_end, cudaMemcpyHostToDevice, S->stream));
  int* best = (int*)(d + o_best);
  stereo_search_kernel<<<div_up(nl, 4), 128, 0, S->stream>>>((const orbx_keypoint*)(d + o_l), (const uint4*)(d + o_dl), nl,
                                                             (const orbx_keypoint*)(d + o_r), (const uint4*)(d + o_dr), nr,
                                                             (const float*)(d + o_s), n_rows, min_d, max_d, best, best + nl);
  ORB_LAUNCH_CHECK();
  A.L = (const orbx_keypoint*)(d + o_l); A.R = (const orbx_keypoint*)(d + o_r); A.nl = nl;
  A.best_idx = best; A.best_dist = best + nl;
  A.min_d = min_d; A.max_d = max_d; A.mbf = mbf;
  A.th_orb = (ORBM_TH_HIGH + ORBM_TH_LOW) / 2;  // :532
  A.u_right = (float*)(d + o_res); A.depth = A.u_right + nl; A.sad = (int*)(A.depth + nl);
  stereo_refine_kernel<<<div_up(nl, 4), 128, 0, S->stream>>>(A);
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(h + o_res, d + o_res, o_end - o_res, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  const float* ur = reinterpret_cast<const float*>(h + o_res);
  const float* z = ur + nl;
  const int* sad = reinterpret_cast<const int*>(z + nl);
  // outlier rejection on the patch distance (:687-700): drop matches at or above 2.1 x the median
  std::vector<std::pair<int, int> > vDistIdx;
  for (int i = 0; i < nl; ++i)
    if (sad[i] >= 0) { u_right[i] = ur[i]; depth[i] = z[i]; vDistIdx.push_back(std::make_pair(sad[i], i)); }
  if (vDistIdx.empty()) return ORB_OK;  // the reference indexes an empty vector here
  std::sort(vDistIdx.begin(), vDistIdx.end());
  const float median = vDistIdx[vDistIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
    if (vDistIdx[i].first < thDist) break;
    u_right[vDistIdx[i].second] = -1;
    depth[vDistIdx[i].second] = -1;
  }
  return ORB_OK;
}

int orbx_stereo_matches_batch(orbx_handle* h, int n_pairs, float mb, float mbf, float* u_right, float* depth, int cap_per_frame) {
  if (!h || n_pairs < 0 || (n_pairs > 0 && (!u_right || !depth))) return set_error(ORB_E_INVALID, "orbx_stereo_matches_batch: bad argument");
  if (n_pairs == 0) return ORB_OK;
  StereoBatchArgs S;
  int rc;
  if ((rc = batch_view(h, &S.B))) return rc;
  if (2 * n_pairs > S.B.n_frames) return set_error(ORB_E_INVALID, "orbx_stereo_matches_batch: %d pairs need %d frames, the last batch had %d", n_pairs, 2 * n_pairs, S.B.n_frames);
  if (cap_per_frame < S.B.out_cap) return set_error(ORB_E_CAPACITY, "orbx_stereo_matches_batch: cap_per_frame %d < %d", cap_per_frame, S.B.out_cap);
  if (S.B.out_cap >= (1 << kRowBits)) return set_error(ORB_E_INVALID, "orbx_stereo_matches_batch: too many keypoints per frame");
  if ((rc = pyramid_view(h, 0, &S.A.pl))) return rc;  // also selects the device
  S.A.pr = S.A.pl;
  const size_t slots = (size_t)n_pairs * S.B.out_cap;
  Layout L;
  const size_t o_best = L.place(slots * 8), o_sad = L.place(slots * 4), o_res = L.place(slots * 8), o_end = L.bytes;
  CallScratch* C = nullptr;
  if ((rc = call_scratch(S.B.device, L.bytes, &C))) return rc;
  uint8_t* d = C->dev;
  S.best_idx = (int*)(d + o_best); S.best_dist = S.best_idx + slots;
  S.n_rows = S.A.pl.h[0];                       // :535
  S.A.L = S.A.R = nullptr; S.A.nl = 0; S.A.best_idx = S.A.best_dist = nullptr;
  S.A.min_d = 0; S.A.max_d = mbf / mb; S.A.mbf = mbf;  // :557-559
  S.A.th_orb = (ORBM_TH_HIGH + ORBM_TH_LOW) / 2;       // :532
  S.A.u_right = (float*)(d + o_res); S.A.depth = S.A.u_right + slots; S.A.sad = (int*)(d + o_sad);
  // the handle's stream orders this after the extraction; the per-thread scratch is free because every call that uses it
  // ends with a synchronisation
  cudaStream_t st = S.B.stream;
  const dim3 grid(div_up(S.B.out_cap, 4), n_pairs);
  stereo_search_batch_kernel<<<grid, 128, 0, st>>>(S);
  ORB_LAUNCH_CHECK();
  stereo_refine_batch_kernel<<<grid, 128, 0, st>>>(S);
  ORB_LAUNCH_CHECK();
  stereo_median_batch_kernel<<<n_pairs, 256, 0, st>>>(S);
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(C->host + o_res, d + o_res, o_end - o_res, cudaMemcpyDeviceToHost, st));
  ORB_CUDA_TRY(cudaStreamSynchronize(st));
  const float* ur = reinterpret_cast<const float*>(C->host + o_res);
  const float* z = ur + slots;
  for (int p = 0; p < n_pairs; ++p) {
    memcpy(u_right + (size_t)p * cap_per_frame, ur + (size_t)p * S.B.out_cap, sizeof(float) * (size_t)S.B.out_cap);
    memcpy(depth + (size_t)p * cap_per_frame, z + (size_t)p * S.B.out_cap, sizeof(float) * (size_t)S.B.out_cap);
    for (int i = S.B.out_cap; i < cap_per_frame; ++i) u_right[(size_t)p * cap_per_frame + i] = depth[(size_t)p * cap_per_frame + i] = -1.0f;
  }
  return ORB_OK;
}

int orbm_distinctive_descriptors(int device, const uint8_t* desc, const int32_t* offset, int n_points, int32_t* best) {
  if (n_points < 0 || (n_points > 0 && (!offset || !best))) return set_error(ORB_E_INVALID, "orbm_distinctive_descriptors: bad argument");
  if (n_points == 0) return ORB_OK;
  if (offset[0] != 0) return set_error(ORB_E_INVALID, "orbm_distinctive_descriptors: offset[0] must be 0");
  for (int p = 0; p < n_points; ++p) {
    const int n = offset[p + 1] - offset[p];
    if (n < 0 || n > kDistinctMaxN)
      return set_error(ORB_E_INVALID, "orbm_distinctive_descriptors: point %d has %d descriptors (0..%d supported)", p, n, kDistinctMaxN);
  }
  const int total = offset[n_points];
  if (total > 0 && !desc) return set_error(ORB_E_INVALID, "orbm_distinctive_descriptors: null descriptors");
  int rc = use_device(device);
  if (rc) return rc;
  Layout L;
  const size_t o_d = L.place((size_t)total * 32), o_off = L.place((size_t)(n_points + 1) * 4);
  const size_t o_up_end = L.bytes;
  const size_t o_best = L.place((size_t)n_points * 4);
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  if (total) memcpy(S->host + o_d, desc, (size_t)total * 32);
  memcpy(S->host + o_off, offset, (size_t)(n_points + 1) * 4);
  ORB_CUDA_TRY(cudaMemcpyAsync(S->dev, S->host, o_up_end, cudaMemcpyHostToDevice, S->stream));
  distinctive_kernel<<<div_up(n_points, kDistinctWarps), kDistinctWarps * 32, 0, S->stream>>>(
      (const uint4*)(S->dev + o_d), (const int*)(S->dev + o_off), n_points, (int*)(S->dev + o_best));
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(S->host + o_best, S->dev + o_best, (size_t)n_points * 4, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  memcpy(best, S->host + o_best, (size_t)n_points * 4);
  return ORB_OK;
}

namespace {
struct DbFileHeader { char magic[4]; uint32_t version; uint64_t n_rows; };
static_assert(sizeof(DbFileHeader) == 16, "16-byte header");
}  // namespace

int orbm_db_save(orbm_db* db, const char* path) {
  if (!db || !path) return set_error(ORB_E_INVALID, "orbm_db_save: bad argument");
  int rc = use_device(db->device);
  if (rc) return rc;
  std::vector<uint8_t> rows((size_t)db->n * 32);
  if (db->n) {
    ORB_CUDA_TRY(cudaMemcpyAsync(rows.data(), db->rows, rows.size(), cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
  }
  FILE* f = fopen(path, "wb");
  if (!f) return set_error(ORB_E_INVALID, "orbm_db_save: cannot open %s", path);
  DbFileHeader h = {{'O', 'R', 'B', 'D'}, 1u, (uint64_t)db->n};
  const bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && (rows.empty() || fwrite(rows.data(), 1, rows.size(), f) == rows.size());
  if (fclose(f) != 0 || !ok) return set_error(ORB_E_INVALID, "orbm_db_save: short write to %s", path);
  return ORB_OK;
}

int orbm_db_load(orbm_db* db, const char* path, int64_t first, int64_t count) {
  if (!db || !path || first < 0) return set_error(ORB_E_INVALID, "orbm_db_load: bad argument");
  FILE* f = fopen(path, "rb");
  if (!f) return set_error(ORB_E_INVALID, "orbm_db_load: cannot open %s", path);
  DbFileHeader h;
  if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, "ORBD", 4) != 0 || h.version != 1) {
    fclose(f);
    return set_error(ORB_E_INVALID, "orbm_db_load: %s is not a descriptor database file", path);
  }
  if ((uint64_t)first > h.n_rows) { fclose(f); return set_error(ORB_E_INVALID, "orbm_db_load: first row beyond the file"); }
  const int64_t avail = (int64_t)h.n_rows - first;
  if (count < 0 || count > avail) count = avail;
  std::vector<uint8_t> rows((size_t)count * 32);
  bool ok = fseek(f, (long)(sizeof(h) + (size_t)first * 32), SEEK_SET) == 0;
  ok = ok && (rows.empty() || fread(rows.data(), 1, rows.size(), f) == rows.size());
  fclose(f);
  if (!ok) return set_error(ORB_E_INVALID, "orbm_db_load: %s is truncated", path);
  return orbm_db_append(db, rows.data(), count, 0);
}

}  // extern "C"

// ====================================================================================
// Sharded relocalization search behind the C ABI: every rank (or device) scans its shard, ONE
// ncclAllGather of nq x 16 B per rank exchanges the per-shard (best1, best2, idx), a device kernel merges them with
// the rule of orbm_merge_top2 and applies the acceptance test of ORBmatcher.cc:429-431, and one transfer of
// nq x 17 B brings the answer to the host -- all on the shard's stream with no host synchronisation in between.
// NCCL is bound at run time (dlopen of the libnccl.so.2 already in the process, e.g. PyTorch's, else the system's),
// so liborbfront.so neither links it nor needs it for anything else.
#include <dlfcn.h>
#include <nccl.h>

namespace {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
};

int nccl_api(const NcclApi** out) {
  static NcclApi api;
  static int state = 0;  // 0 untried, 1 ok, -1 failed
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  if (state == 0) {
    // ORBFRONT_NCCL_LIB names a specific library; otherwise the libnccl.so.2 already in the process (PyTorch's, if PyTorch
    // was loaded first) or the system's
    const char* names[] = {getenv("ORBFRONT_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names)
      if (!api.lib && n && *n) api.lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
    state = -1;
    if (api.lib) {
      bool ok = true;
      auto sym = [&](const char* name) { void* p = dlsym(api.lib, name); ok = ok && p; return p; };
      api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
      api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
      api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
      api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
      api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
      api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
      api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
      api.GetVersion = (decltype(api.GetVersion))sym("ncclGetVersion");
      if (ok) state = 1;
    }
  }
  if (state != 1) return set_error(ORB_E_STATE, "NCCL is not available (dlopen libnccl.so.2: %s)", api.lib ? "symbols missing" : dlerror());
  *out = &api;
  return ORB_OK;
}

#define ORB_NCCL_TRY(api, expr)                                                                               \
  do {                                                                                                        \
    ncclResult_t _r = (expr);                                                                                 \
    if (_r != ncclSuccess)                                                                                    \
      return set_error(ORB_E_CUDA, "%s failed: %s (%s:%d)", #expr, (api)->GetErrorString(_r), __FILE__, __LINE__); \
  } while (0)

// shards[s * nq + q], shards in increasing row order -> one result per query + acceptance (orbm_merge_top2 on the device)
__global__ void merge_top2_kernel(const orbm_top2* __restrict__ shards, int n_shards, int nq, float ratio, int th,
                                  orbm_top2* __restrict__ out, uint8_t* __restrict__ accept) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nq) return;
  int b1 = 256, b2 = 256;
  long long idx = -1;
  for (int s = 0; s < n_shards; ++s) {  // strict <: the lowest shard (= lowest global row) wins ties
    const orbm_top2 p = shards[(size_t)s * nq + q];
    if (p.idx >= 0 && p.best1 < b1) { b2 = b1; b1 = p.best1; idx = p.idx; }
    else if (p.idx >= 0 && p.best1 < b2) b2 = p.best1;
    if (p.best2 < b2) b2 = p.best2;
  }
  orbm_top2 r;
  r.best1 = b1; r.best2 = b2; r.idx = idx;
  out[q] = r;
  accept[q] = (uint8_t)((b1 <= th) && ((float)b1 < __fmul_rn(ratio, (float)b2)));
}

}  // namespace

struct orbm_group {
  int n_local = 0, n_ranks = 0, rank0 = 0;  // shards held by this process, ranks of the communicator, rank of shards[0]
  std::vector<orbm_db*> shards;
  std::vector<ncclComm_t> comms;
  std::vector<orbm_top2*> gathered;   // per local shard: [n_ranks][cap_q] on its device
  std::vector<orbm_top2*> local_out;  // per local shard: [cap_q]
  std::vector<uint32_t*> dq;          // per local shard: queries
  orbm_top2* merged = nullptr;        // device of shards[0]
  uint8_t* accept = nullptr;
  uint8_t* host = nullptr;            // page-locked: queries in, merged + accept out
  int cap_q = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;  // shards[0]'s stream: start, scan done, answer on the host
  bool timed = false;
};

static int group_reserve(orbm_group* g, int nq) {
  if (nq <= g->cap_q) return ORB_OK;
  const int cap = (int)align_up<size_t>((size_t)nq, 256);
  for (int i = 0; i < g->n_local; ++i) {
    int rc = use_device(g->shards[i]->device);
    if (rc) return rc;
    if (g->gathered[i]) cudaFree(g->gathered[i]);
    if (g->local_out[i]) cudaFree(g->local_out[i]);
    if (g->dq[i]) cudaFree(g->dq[i]);
    g->gathered[i] = nullptr; g->local_out[i] = nullptr; g->dq[i] = nullptr;
    ORB_CUDA_TRY(cudaMalloc(&g->gathered[i], sizeof(orbm_top2) * (size_t)cap * g->n_ranks));
    ORB_CUDA_TRY(cudaMalloc(&g->local_out[i], sizeof(orbm_top2) * (size_t)cap));
    ORB_CUDA_TRY(cudaMalloc(&g->dq[i], 32 * (size_t)cap));
  }
  int rc = use_device(g->shards[0]->device);
  if (rc) return rc;
  if (g->merged) cudaFree(g->merged);
  if (g->accept) cudaFree(g->accept);
  if (g->host) cudaFreeHost(g->host);
  g->merged = nullptr; g->accept = nullptr; g->host = nullptr;
  ORB_CUDA_TRY(cudaMalloc(&g->merged, sizeof(orbm_top2) * (size_t)cap));
  ORB_CUDA_TRY(cudaMalloc(&g->accept, (size_t)cap));
  ORB_CUDA_TRY(cudaMallocHost(&g->host, (size_t)cap * (32 + sizeof(orbm_top2) + 1)));
  g->cap_q = cap;
  return ORB_OK;
}

static int group_init_common(orbm_group* g) {
  g->gathered.assign(g->n_local, nullptr);
  g->local_out.assign(g->n_local, nullptr);
  g->dq.assign(g->n_local, nullptr);
  int rc = use_device(g->shards[0]->device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaEventCreate(&g->ev0));
  ORB_CUDA_TRY(cudaEventCreate(&g->ev1));
  ORB_CUDA_TRY(cudaEventCreate(&g->ev2));
  return ORB_OK;
}

extern "C" {

int orbm_group_unique_id(uint8_t* id128) {
  if (!id128) return set_error(ORB_E_INVALID, "orbm_group_unique_id: null");
  const NcclApi* N;
  int rc = nccl_api(&N);
  if (rc) return rc;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  ncclUniqueId id;
  ORB_NCCL_TRY(N, N->GetUniqueId(&id));
  memcpy(id128, &id, 128);
  return ORB_OK;
}

int orbm_group_create(orbm_db* shard, int rank, int n_ranks, const uint8_t* id128, orbm_group** out) {
  if (!shard || !out || !id128 || n_ranks < 1 || rank < 0 || rank >= n_ranks) return set_error(ORB_E_INVALID, "orbm_group_create: bad argument");
  const NcclApi* N;
  int rc = nccl_api(&N);
  if (rc || (rc = use_device(shard->device))) return rc;
  orbm_group* g = new orbm_group();
  g->n_local = 1; g->n_ranks = n_ranks; g->rank0 = rank;
  g->shards.push_back(shard);
  g->comms.assign(1, nullptr);
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  ncclResult_t r = N->CommInitRank(&g->comms[0], n_ranks, id, rank);
  if (r != ncclSuccess) { delete g; return set_error(ORB_E_CUDA, "ncclCommInitRank failed: %s", N->GetErrorString(r)); }
  if ((rc = group_init_common(g))) { orbm_group_destroy(g); return rc; }
  *out = g;
  return ORB_OK;
}

int orbm_group_create_local(orbm_db* const* shards, int n_shards, orbm_group** out) {
  if (!shards || !out || n_shards < 1) return set_error(ORB_E_INVALID, "orbm_group_create_local: bad argument");
  const NcclApi* N;
  int rc = nccl_api(&N);
  if (rc) return rc;
  std::vector<int> devs;
  for (int i = 0; i < n_shards; ++i) {
    if (!shards[i]) return set_error(ORB_E_INVALID, "orbm_group_create_local: null shard");
    if (i && shards[i]->row_base < shards[i - 1]->row_base) return set_error(ORB_E_INVALID, "orbm_group_create_local: shards must be ordered by row_base");
    for (int d : devs)
      if (d == shards[i]->device) return set_error(ORB_E_INVALID, "orbm_group_create_local: one shard per device (NCCL ranks of one communicator need distinct devices)");
    devs.push_back(shards[i]->device);
  }
  orbm_group* g = new orbm_group();
  g->n_local = n_shards; g->n_ranks = n_shards; g->rank0 = 0;
  g->shards.assign(shards, shards + n_shards);
  g->comms.assign(n_shards, nullptr);
  ncclResult_t r = N->CommInitAll(g->comms.data(), n_shards, devs.data());
  if (r != ncclSuccess) { delete g; return set_error(ORB_E_CUDA, "ncclCommInitAll failed: %s", N->GetErrorString(r)); }
  if ((rc = group_init_common(g))) { orbm_group_destroy(g); return rc; }
  *out = g;
  return ORB_OK;
}

int orbm_group_search(orbm_group* g, const uint8_t* q, int nq, float ratio, int th, orbm_top2* out, uint8_t* accept) {
  NvtxRange nvtx_range("orbm_group_search");
  if (!g || nq < 0 || (nq > 0 && (!q || !out))) return set_error(ORB_E_INVALID, "orbm_group_search: bad argument");
  if (nq == 0) return ORB_OK;
  const NcclApi* N;
  int rc = nccl_api(&N);
  if (rc || (rc = group_reserve(g, nq))) return rc;
  orbm_db* d0 = g->shards[0];
  memcpy(g->host, q, (size_t)nq * 32);
  // queries to every local device, scan of every local shard (device-resident form of orbm_top2_search)
  for (int i = 0; i < g->n_local; ++i) {
    orbm_db* db = g->shards[i];
    if ((rc = use_device(db->device))) return rc;
    if (i == 0) ORB_CUDA_TRY(cudaEventRecord(g->ev0, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(g->dq[i], g->host, (size_t)nq * 32, cudaMemcpyHostToDevice, db->stream));
    if ((rc = orbm_top2_search(db, reinterpret_cast<const uint8_t*>(g->dq[i]), nq, g->local_out[i], 1))) return rc;
    if (i == 0) ORB_CUDA_TRY(cudaEventRecord(g->ev1, db->stream));
  }
  // one all-gather of nq x 16 B per rank, on the shards' streams
  if (g->n_ranks > 1) {
    if (g->n_local > 1) ORB_NCCL_TRY(N, N->GroupStart());
    for (int i = 0; i < g->n_local; ++i) {
      if ((rc = use_device(g->shards[i]->device))) return rc;
      ORB_NCCL_TRY(N, N->AllGather(g->local_out[i], g->gathered[i], (size_t)nq * sizeof(orbm_top2), ncclUint8, g->comms[i], g->shards[i]->stream));
    }
    if (g->n_local > 1) ORB_NCCL_TRY(N, N->GroupEnd());
  }
  if ((rc = use_device(d0->device))) return rc;
  const orbm_top2* src = g->n_ranks > 1 ? g->gathered[0] : g->local_out[0];
  merge_top2_kernel<<<div_up(nq, 128), 128, 0, d0->stream>>>(src, g->n_ranks, nq, ratio, th, g->merged, g->accept);
  ORB_LAUNCH_CHECK();
  uint8_t* h_out = g->host + (size_t)g->cap_q * 32;
  uint8_t* h_acc = h_out + (size_t)g->cap_q * sizeof(orbm_top2);
  ORB_CUDA_TRY(cudaMemcpyAsync(h_out, g->merged, (size_t)nq * sizeof(orbm_top2), cudaMemcpyDeviceToHost, d0->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(h_acc, g->accept, (size_t)nq, cudaMemcpyDeviceToHost, d0->stream));
  ORB_CUDA_TRY(cudaEventRecord(g->ev2, d0->stream));
  for (int i = g->n_local - 1; i >= 0; --i) {  // the other local devices only have to finish their all-gather
    if ((rc = use_device(g->shards[i]->device))) return rc;
    ORB_CUDA_TRY(cudaStreamSynchronize(g->shards[i]->stream));
  }
  g->timed = true;
  memcpy(out, h_out, (size_t)nq * sizeof(orbm_top2));
  if (accept) memcpy(accept, h_acc, (size_t)nq);
  return ORB_OK;
}

int orbm_group_last_ms(orbm_group* g, float* scan_ms, float* total_ms) {
  if (!g || !g->timed) return set_error(ORB_E_STATE, "orbm_group_last_ms: no search has run");
  int rc = use_device(g->shards[0]->device);
  if (rc) return rc;
  if (scan_ms) ORB_CUDA_TRY(cudaEventElapsedTime(scan_ms, g->ev0, g->ev1));
  if (total_ms) ORB_CUDA_TRY(cudaEventElapsedTime(total_ms, g->ev0, g->ev2));
  return ORB_OK;
}

int orbm_group_destroy(orbm_group* g) {
  if (!g) return ORB_OK;
  const NcclApi* N = nullptr;
  nccl_api(&N);
  for (int i = 0; i < g->n_local; ++i) {
    cudaSetDevice(g->shards[i]->device);
    cudaStreamSynchronize(g->shards[i]->stream);
    if (i < (int)g->gathered.size() && g->gathered[i]) cudaFree(g->gathered[i]);
    if (i < (int)g->local_out.size() && g->local_out[i]) cudaFree(g->local_out[i]);
    if (i < (int)g->dq.size() && g->dq[i]) cudaFree(g->dq[i]);
    if (N && g->comms[i]) N->CommDestroy(g->comms[i]);
  }
  cudaSetDevice(g->shards[0]->device);
  if (g->merged) cudaFree(g->merged);
  if (g->accept) cudaFree(g->accept);
  if (g->host) cudaFreeHost(g->host);
  if (g->ev0) cudaEventDestroy(g->ev0);
  if (g->ev1) cudaEventDestroy(g->ev1);
  if (g->ev2) cudaEventDestroy(g->ev2);
  delete g;
  return ORB_OK;
}

}  // extern "C"
