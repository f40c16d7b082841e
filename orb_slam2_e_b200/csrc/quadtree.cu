// Deterministic quadtree keypoint selection (sm_100a): one CTA per (frame, level) and kernel.
//
// Replaces ORBextractor::DistributeOctTree + ExtractorNode::DivideNode
// (src/ORBextractor.cc:481-763), a sequential std::list algorithm, by an equivalent
// level-synchronous formulation on SORTED PATH CODES (derivation in DESIGN.md "Quadtree"):
//   * every candidate carries its root strip and the 2-bit quadrant digit of each split it
//     would undergo (tabulated per level, extract.cuh); sorting by that code makes every tree node, at every
//     depth, a contiguous run, and b[i] = depth at which sorted neighbours i-1 and i separate;
//   * the list's node count after each breadth-first pass is cnt[d] = #{i : b[i] <= d}, the
//     number of still-divisible nodes is nexp[d]; the reference's stop rules (:669, :673) are
//     replayed on those two tables;
//   * the final "expand the biggest nodes first until N" passes (:676-737) sort the divisible
//     nodes by (key count, creation order) -- creation order being the reference's pointer
//     tie-break under a monotonic allocator -- and a prefix sum finds where the loop breaks (:730);
//   * each leaf keeps its max-response key, first candidate winning ties (:744-760), and
//     leaves are emitted in the std::list's final order (newest pass first, reverse creation
//     order inside a pass, single-key roots last in ascending order).
// Integer/byte work, latency-bound inside a CTA; parallelism comes from the batch:
// n_frames x n_levels independent CTAs.
//
// Two kernels implement this (launch_quadtree runs both, in this order):
//   quadtree_pyramid_kernel  round 2, the common case: no sort at all -- per-cell key counts and best keys on a grid of
//                            depth D0, the rules above replayed on the count pyramid (see its own comment below); it hands
//                            a (frame, level) back (kept_count = -1) when the loop would need a deeper grid;
//   quadtree_select_kernel   the sorted-path-code formulation described above, for the pairs handed back (and for
//                            everything when ORBX_QT_NO_PYRAMID is set): every other CTA returns at once.
// Both produce the same kept set in the same order (tests/test_gpu_quadtree_paths.py forces each of them).
#include "extract.cuh"

namespace orbfront {

constexpr int kQtThreads = 256;       // batches: many small CTAs per SM hide each other's latency
constexpr int kQtThreadsWide = 1024;  // one or a few frames: a level's CTA is the critical path, so it gets 4x the lanes

#ifdef ORBX_QT_PROFILE2  // developer build: cycles per phase summed over all CTAs, by level (tools/exp_qt_profile2.py)
__device__ unsigned long long g_qt_acc[16][12];
#define QT_STAMP(k) do { if (threadIdx.x == 0) qt_st[k] = clock64(); } while (0)
#elif defined(ORBX_QT_PROFILE)  // developer build: cycle stamps of the level-0 CTA of frame 0 (tools/exp_qt_profile.py)
__device__ long long g_qt_prof[32];
#define QT_STAMP(k) do { if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) g_qt_prof[k] = clock64(); } while (0)
#else
#define QT_STAMP(k) do { } while (0)
#endif
__device__ unsigned int g_qt_handed_back;  // (frame, level) pairs the sort-free kernel handed to the sorting kernel (orbx_debug_qt_handed_back)
constexpr uint32_t kSeqMark = 0x8000u;  // "head of a child created by the final partial pass"

struct QtCtl {  // control block in shared memory
  int hist[kMaxDepth + 3], dlt[kMaxDepth + 3], cnt[kMaxDepth + 3], nexp[kMaxDepth + 3];
  int d, size, mode, d_bfs, partial, counter, mstar, scan_tmp[kQtThreadsWide / 32], scan_total, run_base;
};
enum { kModeFull = 0, kModeOrdered = 1 };

// In-place ascending bitonic sort of np (power of two) 64-bit keys with optional 32-bit values
// (node tables: at most a few hundred entries).  One compare-exchange per thread and step.
template <int T>
__device__ void bitonic_sort(unsigned long long* a, uint32_t* v, int np) {
  for (int k = 2; k <= np; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (np >> 1); t += T) {
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1)), x = i | j;
        const unsigned long long ai = a[i], ax = a[x];
        if ((ai > ax) == ((i & k) == 0)) {
          a[i] = ax; a[x] = ai;
          if (v) { const uint32_t tv = v[i]; v[i] = v[x]; v[x] = tv; }
        }
      }
      __syncthreads();
    }
  }
}

// The same sort for tables of at most T entries, one entry per thread in REGISTERS: compare-exchanges with a partner in the
// same warp (stride < 32: 30 of the 36 steps for 256 entries) are three shuffles and a few selects -- no shared-memory
// traffic, no block barrier; only strides >= 32 go through the table.  A third of the instructions of the network above.
template <int T>
__device__ void sort_table(unsigned long long* a, uint32_t* v, int np) {
  if (np > T) { bitonic_sort<T>(a, v, np); return; }
  const int t = threadIdx.x, npe = max(np, 32);
  unsigned long long key = ~0ull;
  uint32_t val = 0;
  if (t < np) { key = a[t]; val = v[t]; }
  for (int k = 2; k <= npe; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      unsigned long long ok;
      uint32_t ov;
      if (j >= 32) {  // uniform
        __syncthreads();
        if (t < npe) { a[t] = key; v[t] = val; }
        __syncthreads();
        ok = t < npe ? a[t ^ j] : key;
        ov = t < npe ? v[t ^ j] : val;
      } else {
        ok = __shfl_xor_sync(0xffffffffu, key, j);
        ov = __shfl_xor_sync(0xffffffffu, val, j);
      }
      const bool keep_min = ((t & j) == 0) == ((t & k) == 0);
      const bool take = keep_min ? ok < key : ok > key;
      if (take) { key = ok; val = ov; }
    }
  }
  __syncthreads();
  if (t < npe && t < np) { a[t] = key; v[t] = val; }
  __syncthreads();
}

// Inclusive scan of one int per thread across the CTA; returns the thread's inclusive prefix.
template <int T>
__device__ int block_inclusive_scan(int v, QtCtl& C) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, v, s);
    if (lane >= s) v += o;
  }
  if (lane == 31) C.scan_tmp[warp] = v;
  __syncthreads();
  int base = 0;
  for (int w = 0; w < warp; ++w) base += C.scan_tmp[w];
  if (threadIdx.x == T - 1) C.scan_total = base + v;
  __syncthreads();
  return base + v;
}

// Stable LSD radix sort (8-bit digits) of n 64-bit keys on bits [32, 32 + nbits) -- the quadtree
// path code -- ping-ponging between buffers a and b (shared or global memory).  Each warp owns a
// contiguous chunk; per-(warp, digit) counters scanned in (digit, warp) order give every key its
// stable destination: base + rank among the warp's equal digits (__match_any_sync).  O(n) per pass
// instead of the O(n log^2 n) compare-exchanges of a sorting network.  Returns the sorted buffer.
template <int T>
__device__ unsigned long long* radix_sort_codes(unsigned long long* a, unsigned long long* b, int n, int nbits,
                                                uint32_t* whist /* [T / 32][256] */, QtCtl& C) {
  constexpr int kWarps = T / 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int per_warp = ((n + T - 1) / T) * 32;
  const int lo = min(n, warp * per_warp), hi = min(n, lo + per_warp);
  for (int shift = 32; shift < 32 + nbits; shift += 8) {
#ifdef ORBX_QT_PROFILE2
    long long r0 = clock64();
#define RSTAMP(k) do { if (tid == 0) { const long long t = clock64(); atomicAdd(&g_qt_acc[15][k], (unsigned long long)(t - r0)); r0 = t; } } while (0)
#else
#define RSTAMP(k) do { } while (0)
#endif
    for (int i = tid; i < kWarps * 256; i += T) whist[i] = 0;
    __syncthreads();
    RSTAMP(0);
    for (int i = lo + lane; i < hi; i += 32) atomicAdd(&whist[warp * 256 + (int)((a[i] >> shift) & 255u)], 1u);
    __syncthreads();
    RSTAMP(1);
    if constexpr (kWarps <= 8) {  // T == 256: one thread per digit value, the warps' counts stay in registers
      uint32_t c[kWarps], tot = 0;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) { c[w] = whist[w * 256 + tid]; tot += c[w]; }
      uint32_t base = (uint32_t)block_inclusive_scan<T>((int)tot, C) - tot;
#pragma unroll
      for (int w = 0; w < kWarps; ++w) { whist[w * 256 + tid] = base; base += c[w]; }
    } else {  // wider CTAs: threads beyond 255 only take part in the scan's barriers
      uint32_t tot = 0;
      if (tid < 256) {
#pragma unroll 8
        for (int w = 0; w < kWarps; ++w) tot += whist[w * 256 + tid];
      }
      uint32_t base = (uint32_t)block_inclusive_scan<T>((int)tot, C) - tot;
      if (tid < 256) {
#pragma unroll 8
        for (int w = 0; w < kWarps; ++w) { const uint32_t c = whist[w * 256 + tid]; whist[w * 256 + tid] = base; base += c; }
      }
    }
    __syncthreads();
    RSTAMP(2);
    for (int i0 = lo; i0 < hi; i0 += 32) {
      const int i = i0 + lane;
      const bool act = i < hi;
      const unsigned long long key = act ? a[i] : 0ull;
      const int d = act ? (int)((key >> shift) & 255u) : 256 + lane;  // idle lanes match nobody
      const uint32_t peers = __match_any_sync(0xffffffffu, d);
      const int r = __popc(peers & lt_mask);
      uint32_t base = 0;
      if (act) {
        base = whist[warp * 256 + d];
        b[base + r] = key;
      }
      __syncwarp();
      if (act && r == 0) whist[warp * 256 + d] = base + __popc(peers);
      __syncwarp();
    }
    __syncthreads();
    RSTAMP(3);
    unsigned long long* t = a; a = b; b = t;
  }
  return a;
}

// Creation-order key of a depth-k node made by breadth-first passes only: the path prefix with
// every second digit (counting back from the last) complemented, the root strip complemented when k
// is even.  Ascending key == order of creation within pass k.
__device__ __forceinline__ uint32_t bfs_creation_key(uint32_t code, int k, int dmax) {
  const uint32_t prefix = code >> (2 * (dmax - k));
  const uint32_t low = (1u << (2 * k)) - 1u;
  uint32_t mask = 0xCCCCCCCCu & low;
  if ((k & 1) == 0) mask |= ((1u << kRootBits) - 1u) << (2 * k);
  return prefix ^ mask;
}

template <int T>
__global__ void __launch_bounds__(T)
quadtree_select_kernel(const __grid_constant__ ExtractParams P, int S, int NP, int only_flagged) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int l = blockIdx.x, f = blockIdx.y, tid = threadIdx.x;
  const LevelGeom& G = P.lv[l];
  const int N = G.quota, dmax = G.dmax;

  // shared-memory carve-up
  unsigned long long* s_keys = reinterpret_cast<unsigned long long*>(smem);
  unsigned long long* nkey = s_keys + S;
  uint32_t* nval = reinterpret_cast<uint32_t*>(nkey + NP);
  uint32_t* nhead = nval + NP;
  uint16_t* s_seq = reinterpret_cast<uint16_t*>(nhead + NP);
  uint8_t* ngain = reinterpret_cast<uint8_t*>(s_seq + S);
  uint8_t* s_b = ngain + NP;
  QtCtl& C = *reinterpret_cast<QtCtl*>(smem + align_up<size_t>((size_t)S * 10 + (size_t)NP * 17 + S + 1, 16));  // NP >= 512: radix counters alias the node tables

#ifdef ORBX_QT_PROFILE2
  long long qt_st[10];
#endif
  QT_STAMP(0);
  if (only_flagged && P.kept_count[f * P.nlevels + l] != -1) return;  // done by quadtree_pyramid_kernel
  if (only_flagged && tid == 0) atomicAdd(&g_qt_handed_back, 1u);
  int n = P.cand_count[f * P.nlevels + l];
  n = min(n, G.cand_cap);
  uint32_t* kept = P.kept + (size_t)f * P.kept_frame + G.kept_off;
  if (n == 0) {
    if (tid == 0) P.kept_count[f * P.nlevels + l] = 0;
    return;
  }
  unsigned long long* gkeys = P.cand + (size_t)f * P.cand_frame + G.cand_off;
  const bool in_smem = n <= S;
  uint8_t* b = in_smem ? s_b : P.qt_b + ((size_t)f * (P.cand_frame + kMaxLevels) + G.cand_off + l);
  // sequence numbers: 16-bit in shared memory, or the same width in the global scratch
  uint16_t* seq = in_smem ? s_seq : reinterpret_cast<uint16_t*>(P.qt_seq + (size_t)f * P.cand_frame + G.cand_off);

  // FAST stores the payload only; attach the path code (all lanes busy), in place
  for (int i = tid; i < n; i += T) {
    const uint32_t pay = (uint32_t)gkeys[i];
    const uint32_t code = __ldg(&P.qcode[G.qx + (int)(pay & 0xfff)]) | __ldg(&P.qcode[G.qy + (int)((pay >> 12) & 0xfff)]);
    gkeys[i] = ((unsigned long long)code << 32) | pay;
  }
  __syncthreads();
  QT_STAMP(1);
  // sort by path code: global <-> shared ping-pong when the level fits, global <-> global scratch else
  int root_bits = 0;
  while ((1 << root_bits) < G.nIni) ++root_bits;
  const int nbits = 2 * dmax + root_bits;
  unsigned long long* other = in_smem ? s_keys : P.qt_tmp + (size_t)f * P.cand_frame + G.cand_off;
  unsigned long long* keys = radix_sort_codes<T>(gkeys, other, n, nbits, reinterpret_cast<uint32_t*>(nkey), C);
  if (in_smem && keys != s_keys) {  // even number of passes: bring the result into shared memory
    for (int i = tid; i < n; i += T) s_keys[i] = keys[i];
    keys = s_keys;
    __syncthreads();
  }

  QT_STAMP(2);
  // boundary depths b[0..n], tables cnt[] / nexp[]
  for (int i = tid; i < kMaxDepth + 3; i += T) { C.hist[i] = 0; C.dlt[i] = 0; }
  for (int i = tid; i <= n; i += T) {
    int v = 0;
    if (i > 0 && i < n) {
      const uint32_t x = (uint32_t)(keys[i - 1] >> 32) ^ (uint32_t)(keys[i] >> 32);
      if (x == 0) v = dmax + 1;  // duplicate position: never separates (cannot happen for FAST output)
      else { const int msb = 31 - __clz(x); v = msb >= 2 * dmax ? 0 : dmax - (msb >> 1); }
    }
    b[i] = (uint8_t)v;
    if (i < n) seq[i] = 0;
  }
  __syncthreads();
  for (int i = tid; i < n; i += T) {
    const int a = b[i], e = b[i + 1];
    atomicAdd(&C.hist[a], 1);
    if (a < e) { atomicAdd(&C.dlt[a], 1); atomicAdd(&C.dlt[e], -1); }
  }
  __syncthreads();

  QT_STAMP(3);
  // replay of the reference's pass loop (ORBextractor.cc:593-739) on the tables
  if (tid < 32) {  // lane <-> depth (kMaxDepth + 3 <= 32): prefix sums by shuffles, the stop rules of every pass at once
    const int dd = tid;
    int c = dd <= dmax ? C.hist[dd] : 0, e = dd <= dmax ? C.dlt[dd] : 0;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const int oc = __shfl_up_sync(0xffffffffu, c, s), oe = __shfl_up_sync(0xffffffffu, e, s);
      if (dd >= s) { c += oc; e += oe; }
    }
    if (dd > dmax) e = 0;  // cnt[d > dmax] = cnt[dmax] (the scan of zeros), nexp[d > dmax] = 0
    if (dd < kMaxDepth + 3) { C.cnt[dd] = c; C.nexp[dd] = e; }
    // pass d (1..dmax) ends the loop if the list reached N nodes or did not grow (:669), or if the next pass could
    // overshoot (:673); if no pass does, the loop ends at dmax because the list no longer grows
    const int prev = __shfl_up_sync(0xffffffffu, c, 1);
    const bool stop = c >= N || c == prev, ordered = !stop && c + 3 * e > N;
    const uint32_t m = __ballot_sync(0xffffffffu, dd >= 1 && dd <= dmax && (stop || ordered));
    const int d = m ? __ffs(m) - 1 : dmax;
    const int size = __shfl_sync(0xffffffffu, c, d);
    const int mode = m && __shfl_sync(0xffffffffu, (int)ordered, d) ? kModeOrdered : kModeFull;
    if (tid == 0) { C.d = d; C.size = size; C.mode = mode; C.d_bfs = d; C.partial = 0; }
  }
  __syncthreads();

  QT_STAMP(4);
  // ordered passes: biggest nodes first, stop as soon as the list holds N nodes (:676-737)
  bool first_ordered = true;
  while (C.mode == kModeOrdered) {
    const int d = C.d, size = C.size;
    if (tid == 0) { C.counter = 0; C.mstar = 0x7fffffff; }
    __syncthreads();
    for (int i = tid; i < n; i += T)
      if (b[i] <= d && b[i + 1] > d) nhead[atomicAdd(&C.counter, 1)] = (uint32_t)i;
    __syncthreads();
    const int M = C.counter;
    int mp = 1;
    while (mp < M) mp <<= 1;
    for (int m = tid; m < mp; m += T) {
      if (m >= M) { nkey[m] = ~0ull; nval[m] = 0; continue; }
      const int h = (int)nhead[m];
      int e = h + 1, children = 1;
      while (b[e] > d) { children += b[e] == d + 1; ++e; }
      ngain[m] = (uint8_t)(children - 1);
      const uint32_t code = (uint32_t)(keys[h] >> 32);
      const uint32_t rank = first_ordered ? bfs_creation_key(code, d, dmax) : (uint32_t)seq[h];
      nkey[m] = ~(((unsigned long long)(e - h) << 32) | rank);  // ascending ~key == (count, rank) descending
      nval[m] = (uint32_t)m;
    }
    __syncthreads();
    if (first_ordered) QT_STAMP(16);
    sort_table<T>(nkey, nval, mp);
    if (first_ordered) QT_STAMP(17);
    // where does the running node count reach N?
    int carried = 0;
    for (int base = 0; base < M; base += T) {
      const int j = base + tid;
      const int g = j < M ? ngain[nval[j]] : 0;
      const int incl = block_inclusive_scan<T>(g, C) + carried;
      if (j < M && size + incl >= N) atomicMin(&C.mstar, j);
      carried += C.scan_total;
      __syncthreads();
    }
    const bool stops = C.mstar != 0x7fffffff;
    const int n_split = stops ? C.mstar + 1 : M;
    // the split nodes' children get their creation sequence (parent's place in the order, quadrant)
    for (int j = tid; j < n_split; j += T) {
      const int h = (int)nhead[nval[j]];
      int p = h;
      do {
        if (p == h || b[p] == d + 1) {
          const uint32_t q = ((uint32_t)(keys[p] >> 32) >> (2 * (dmax - d - 1))) & 3u;
          seq[p] = (uint16_t)(((uint32_t)j << 2 | q) | (stops ? kSeqMark : 0u));
        }
        ++p;
      } while (b[p] > d);
    }
    __syncthreads();
    if (tid == 0) {
      if (stops) { C.partial = 1; C.mode = kModeFull; }
      else {
        const int nsize = C.cnt[d + 1];
        C.d = d + 1; C.size = nsize;
        if (nsize >= N || nsize == size || d + 1 >= dmax) C.mode = kModeFull;  // :735
      }
    }
    if (first_ordered) QT_STAMP(18);
#ifdef ORBX_QT_PROFILE
    if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) { g_qt_prof[19] += 1; if (first_ordered) g_qt_prof[20] = M; }
#endif
    first_ordered = false;
    __syncthreads();
  }

  QT_STAMP(5);
  // leaves: runs at depth dfin, plus the children of the nodes split by the final partial pass
  const int dfin = C.d, partial = C.partial, d_bfs = C.d_bfs;
  const uint32_t mag_h = (1u << 20) / (uint32_t)G.hCell + 1u, mag_w = (1u << 20) / (uint32_t)G.wCell + 1u;
  if (tid == 0) C.counter = 0;
  __syncthreads();
  for (int i = tid; i < n; i += T)
    if (b[i] <= dfin || (partial && (seq[i] & kSeqMark))) nhead[atomicAdd(&C.counter, 1)] = (uint32_t)i;
  __syncthreads();
  QT_STAMP(6);
  const int L = min(C.counter, G.kept_cap);
  int lp = 1;
  while (lp < L) lp <<= 1;
  for (int m = tid; m < lp; m += T) {
    if (m >= L) { nkey[m] = ~0ull; nval[m] = 0; continue; }
    const int p = (int)nhead[m];
    const bool child = partial && (seq[p] & kSeqMark);
    const int leaf_d = child ? dfin + 1 : dfin;
    // best key of the run: highest response, then earliest candidate in the reference's
    // (cell row, cell col, y, x) detection order
    int e = p;
    unsigned long long best = 0;
    uint32_t best_pay = 0;
    do {
      const uint32_t pay = (uint32_t)keys[e];
      const int x = pay & 0xfff, y = (pay >> 12) & 0xfff;
      // (y-3) / hCell and (x-3) / wCell by multiplication: exact for operands < 4096 and divisors < 64
      const unsigned long long cell = (unsigned long long)((((uint32_t)(y - 3) * mag_h) >> 20) * (uint32_t)G.nCols + (((uint32_t)(x - 3) * mag_w) >> 20));
      const unsigned long long order = (cell << 24) | (unsigned long long)(pay & 0xffffffu);
      const unsigned long long cmp = ((unsigned long long)(pay >> 24) << 48) | (0xffffffffffffull - order);
      if (cmp > best || e == p) { best = cmp; best_pay = pay; }
      ++e;
    } while (e < n && !(b[e] <= dfin || (partial && (seq[e] & kSeqMark))));
    const int created = (e - p == 1) ? max((int)b[p], (int)b[p + 1]) : leaf_d;
    const uint32_t code = (uint32_t)(keys[p] >> 32);
    unsigned long long within;
    if (created == 0) within = code >> (2 * dmax);  // single-key roots keep their push_back order
    else if (created <= d_bfs) within = 0xffffffffull - bfs_creation_key(code, created, dmax);
    else within = 0xffffffffull - (uint32_t)(seq[p] & ~kSeqMark);
    nkey[m] = ((unsigned long long)(31 - created) << 40) | within;
    nval[m] = best_pay;
  }
  __syncthreads();
  QT_STAMP(7);
  sort_table<T>(nkey, nval, lp);
  QT_STAMP(8);
  for (int m = tid; m < L; m += T) kept[m] = nval[m];
  if (tid == 0) P.kept_count[f * P.nlevels + l] = L;
  QT_STAMP(9);
#ifdef ORBX_QT_PROFILE2
  if (tid == 0) {
    for (int k = 0; k < 9; ++k) atomicAdd(&g_qt_acc[l][k], (unsigned long long)(qt_st[k + 1] - qt_st[k]));
    atomicAdd(&g_qt_acc[l][9], 1ull);
    atomicAdd(&g_qt_acc[l][10], (unsigned long long)n);
  }
#endif
#ifdef ORBX_QT_PROFILE
  if (blockIdx.x == 0 && blockIdx.y == 0 && tid == 0) { g_qt_prof[10] = n; g_qt_prof[11] = L; g_qt_prof[12] = C.d_bfs; g_qt_prof[13] = C.d; g_qt_prof[14] = dmax; g_qt_prof[15] = nbits; }
#endif
}

// ------------------------------------------------------------------------------------------------------------------
// Round 2: the same selection WITHOUT sorting the candidates, for the common case that the reference's loop ends at a
// shallow depth.  Everything the replay, the ordered passes and the leaf selection need is a function of per-node key
// counts and per-node best keys, and the nodes of depth d are the cells of a 2^d x 2^d grid per root (index = the path
// code's prefix).  So: one pass of shared-memory atomics counts the candidates per cell of the depth-D0 grid and keeps
// each cell's best key with a 64-bit atomicMax; the coarser grids are sums of four children; cnt[d] / nexp[d] are the
// numbers of cells with >= 1 / >= 2 keys; an ordered pass enumerates the cells of depth d with >= 2 keys (children =
// non-empty cells below them); a leaf's best key is the maximum over its depth-D0 descendants, and the depth at which a
// single key became alone is found by walking up while the parent still holds one key.  D0 is chosen per level on the
// host (about 6 x quota cells).  Whenever the loop would need depth D0 + 1 (sparse or clustered levels, very large
// quotas), the CTA writes kept_count = -1 and the sorting kernel above redoes that (frame, level): results are
// identical either way, only the time differs.  Replaces three radix passes, the boundary pass and two serial run
// walks per candidate set: the level-0 CTA of a textured 640x480 frame took 205 k cycles with the sorting kernel.

__host__ __device__ inline int qp_level_off(int R, int D0, int d) {  // first cell of depth d in the pyramid arrays: D0 first (word aligned), then D0-1 .. 0
  int off = 0;
  for (int k = D0; k > d; --k) off += R << (2 * k);
  return off;
}

template <int T>
__global__ void __launch_bounds__(T)
quadtree_pyramid_kernel(const __grid_constant__ ExtractParams P, int NP, int count_words, int split_words, int min_ratio4) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int l = blockIdx.x, f = blockIdx.y, tid = threadIdx.x, lane = tid & 31;
  const LevelGeom& G = P.lv[l];
  const int N = G.quota, dmax = G.dmax, D0 = G.qt_d0;
  int* const kc = &P.kept_count[f * P.nlevels + l];
  int n = min(P.cand_count[f * P.nlevels + l], G.cand_cap);
#ifdef ORBX_QT_PROFILE2
  long long qt_st[10];
#endif
  QT_STAMP(0);
  if (n == 0) { if (tid == 0) *kc = 0; return; }
  // sparse levels (few candidates per wanted keypoint) end deep in the tree: not worth trying here
  if (D0 <= 0 || n >= 65535 || 4ll * n < (long long)min_ratio4 * N) { if (tid == 0) *kc = -1; return; }  // the sorting kernel takes this level

  uint32_t* split = reinterpret_cast<uint32_t*>(smem) + count_words;   // bitmap over the cells of one depth: split by the final partial pass
  unsigned long long* nkey = reinterpret_cast<unsigned long long*>(split + split_words);
  uint32_t* nval = reinterpret_cast<uint32_t*>(nkey + NP);
  uint32_t* nhead = nval + NP;
  uint8_t* ngain = reinterpret_cast<uint8_t*>(nhead + NP);
  QtCtl& C = *reinterpret_cast<QtCtl*>(smem + align_up<size_t>((size_t)count_words * 4 + (size_t)split_words * 4 + (size_t)NP * 17, 16));
  int root_bits = 0;
  while ((1 << root_bits) < G.nIni) ++root_bits;
  const int R = 1 << root_bits;
  const int cellsD0 = R << (2 * D0);
  const int total_cells = qp_level_off(R, D0, 0) + R, my_words = (total_cells + 3) / 4 * 2;
  const unsigned long long* gkeys = P.cand + (size_t)f * P.cand_frame + G.cand_off;
  unsigned long long* best = P.qt_tmp + (size_t)f * P.cand_frame + G.cand_off;   // best key of every depth-D0 cell
  uint32_t* seq = P.qt_seq + (size_t)f * P.cand_frame + G.cand_off;              // creation sequence per pyramid cell (written before read)
  // key counts, u16 per cell, whole pyramid: shared memory, or (large grids) the level's global scratch behind the sequence numbers
  uint32_t* cw = G.qt_counts_global ? seq + (total_cells + 3) / 4 * 4 : reinterpret_cast<uint32_t*>(smem);
  uint16_t* c16 = reinterpret_cast<uint16_t*>(cw);
  uint32_t* kept = P.kept + (size_t)f * P.kept_frame + G.kept_off;
  const uint32_t mag_h = (1u << 20) / (uint32_t)G.hCell + 1u, mag_w = (1u << 20) / (uint32_t)G.wCell + 1u;

  for (int i = tid; i < my_words; i += T) cw[i] = 0;
  for (int i = tid; i < split_words; i += T) split[i] = 0;
  for (int i = tid; i < cellsD0; i += T) best[i] = 0ull;
  for (int i = tid; i < kMaxDepth + 3; i += T) { C.cnt[i] = 0; C.nexp[i] = 0; }
  __syncthreads();
  QT_STAMP(1);
  // one pass over the candidates: count and best key per depth-D0 cell
  const int shD0 = 2 * (dmax - D0);
  for (int i = tid; i < n; i += T) {
    const uint32_t pay = (uint32_t)gkeys[i];
    const int x = pay & 0xfff, y = (pay >> 12) & 0xfff;
    const uint32_t code = __ldg(&P.qcode[G.qx + x]) | __ldg(&P.qcode[G.qy + y]);
    const uint32_t c = code >> shD0;
    atomicAdd(&cw[c >> 1], 1u << ((c & 1u) * 16));
    // highest response, then earliest candidate in the reference's (cell row, cell col, y, x) detection order (:744-760)
    const unsigned long long cell = (unsigned long long)((((uint32_t)(y - 3) * mag_h) >> 20) * (uint32_t)G.nCols + (((uint32_t)(x - 3) * mag_w) >> 20));
    const unsigned long long order = (cell << 24) | (unsigned long long)(pay & 0xffffffu);
    atomicMax(&best[c], ((unsigned long long)(pay >> 24) << 48) | (0xffffffffffffull - order));
  }
  __syncthreads();
  QT_STAMP(2);
  // coarser grids (sums of four children) and, per depth, the number of nodes (cells with keys) and of divisible nodes
  for (int d = D0; d >= 0; --d) {
    const int cells = R << (2 * d), off = qp_level_off(R, D0, d), offc = d < D0 ? qp_level_off(R, D0, d + 1) : 0;
    int nz = 0, ge2 = 0;
    for (int c = tid; c < cells; c += T) {
      int v;
      if (d == D0) v = c16[off + c];
      else {
        const uint2 q = *reinterpret_cast<const uint2*>(&c16[offc + 4 * c]);  // four u16 children (offsets are multiples of 4 cells)
        v = (int)(q.x & 0xffffu) + (int)(q.x >> 16) + (int)(q.y & 0xffffu) + (int)(q.y >> 16);
        c16[off + c] = (uint16_t)v;
      }
      nz += v > 0;
      ge2 += v > 1;
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) { nz += __shfl_xor_sync(0xffffffffu, nz, s); ge2 += __shfl_xor_sync(0xffffffffu, ge2, s); }
    if (lane == 0 && (nz | ge2)) { atomicAdd(&C.cnt[d], nz); atomicAdd(&C.nexp[d], ge2); }
    __syncthreads();
  }
  // replay of the reference's pass loop (:593-739) on cnt[] / nexp[], depths 1 .. Dr; the loop must end there
  QT_STAMP(3);
  const int Dr = min(D0 - 1, dmax);
  if (tid < 32) {
    const int dd = tid;
    const int c = dd <= D0 ? C.cnt[dd] : 0, e = dd <= D0 ? C.nexp[dd] : 0;
    const int prev = __shfl_up_sync(0xffffffffu, c, 1);
    const bool stop = c >= N || c == prev, ordered = !stop && c + 3 * e > N;
    const uint32_t m = __ballot_sync(0xffffffffu, dd >= 1 && dd <= Dr && (stop || ordered));
    int d = -1, mode = kModeFull;
    if (m) { d = __ffs(m) - 1; mode = __shfl_sync(0xffffffffu, (int)ordered, d) ? kModeOrdered : kModeFull; }
    else if (Dr == dmax) d = dmax;  // the list stops growing at the last depth
    const int size = __shfl_sync(0xffffffffu, c, max(d, 0));
    if (tid == 0) { C.d = d; C.size = size; C.mode = mode; C.d_bfs = d; C.partial = 0; }
  }
  __syncthreads();
  if (C.d < 0) { if (tid == 0) *kc = -1; return; }
  QT_STAMP(4);

  // ordered passes: biggest nodes first, stop as soon as the list holds N nodes (:676-737)
  bool first_ordered = true;
  while (C.mode == kModeOrdered) {
    const int d = C.d, size = C.size;
    if (d + 1 > D0) { if (tid == 0) *kc = -1; return; }  // children below the pyramid (uniform: C.d is shared)
    const int cells = R << (2 * d), off = qp_level_off(R, D0, d), offc = qp_level_off(R, D0, d + 1);
    if (tid == 0) { C.counter = 0; C.mstar = 0x7fffffff; }
    __syncthreads();
    for (int c = tid; c < cells; c += T)
      if (c16[off + c] > 1) nhead[atomicAdd(&C.counter, 1)] = (uint32_t)c;
    __syncthreads();
    const int M = C.counter;
    int mp = 1;
    while (mp < M) mp <<= 1;
    for (int m = tid; m < mp; m += T) {
      if (m >= M) { nkey[m] = ~0ull; nval[m] = 0; continue; }
      const int c = (int)nhead[m];
      int children = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) children += c16[offc + 4 * c + q] > 0;
      ngain[m] = (uint8_t)(children - 1);
      const uint32_t rank = first_ordered ? bfs_creation_key((uint32_t)c << (2 * (dmax - d)), d, dmax) : seq[off + c];
      nkey[m] = ~(((unsigned long long)c16[off + c] << 32) | rank);  // ascending ~key == (count, rank) descending
      nval[m] = (uint32_t)m;
    }
    __syncthreads();
    sort_table<T>(nkey, nval, mp);
    int carried = 0;
    for (int base = 0; base < M; base += T) {  // where does the running node count reach N?
      const int j = base + tid;
      const int g = j < M ? ngain[nval[j]] : 0;
      const int incl = block_inclusive_scan<T>(g, C) + carried;
      if (j < M && size + incl >= N) atomicMin(&C.mstar, j);
      carried += C.scan_total;
      __syncthreads();
    }
    const bool stops = C.mstar != 0x7fffffff;
    const int n_split = stops ? C.mstar + 1 : M;
    // the split nodes' children get their creation sequence (parent's place in the order, quadrant)
    for (int j = tid; j < n_split; j += T) {
      const int c = (int)nhead[nval[j]];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (c16[offc + 4 * c + q] > 0) seq[offc + 4 * c + q] = (uint32_t)j << 2 | (uint32_t)q;
      if (stops) atomicOr(&split[c >> 5], 1u << (c & 31));
    }
    __syncthreads();
    if (tid == 0) {
      if (stops) { C.partial = 1; C.mode = kModeFull; }
      else {
        const int nsize = C.cnt[d + 1];
        C.d = d + 1; C.size = nsize;
        if (nsize >= N || nsize == size || d + 1 >= dmax) C.mode = kModeFull;  // :735
      }
    }
    first_ordered = false;
    __syncthreads();
  }

  // leaves: the nodes of depth dfin, and the children of those split by the final partial pass
  QT_STAMP(5);
  const int dfin = C.d, partial = C.partial, d_bfs = C.d_bfs;
  if (dfin + partial > D0 || 2 * (D0 - dfin) > 8) { if (tid == 0) *kc = -1; return; }  // (leaf scans cover at most 256 depth-D0 cells)
  {
    const int cells = R << (2 * dfin), off = qp_level_off(R, D0, dfin), offc = partial ? qp_level_off(R, D0, dfin + 1) : 0;
    if (tid == 0) C.counter = 0;
    __syncthreads();
    for (int c = tid; c < cells; c += T) {
      if (c16[off + c] == 0) continue;
      if (partial && (split[c >> 5] >> (c & 31) & 1u)) {
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (c16[offc + 4 * c + q] > 0) nhead[min(atomicAdd(&C.counter, 1), NP - 1)] = (uint32_t)(4 * c + q) | 0x80000000u;
      } else {
        nhead[min(atomicAdd(&C.counter, 1), NP - 1)] = (uint32_t)c;
      }
    }
    __syncthreads();
  }
  const int L = min(C.counter, G.kept_cap);
  QT_STAMP(6);
  int lp = 1;
  while (lp < L) lp <<= 1;
  for (int m = tid; m < lp; m += T) {
    if (m >= L) { nkey[m] = ~0ull; nval[m] = 0; continue; }
    const uint32_t hd = nhead[m];
    const int dl = dfin + (int)(hd >> 31);  // leaf depth
    const int c = (int)(hd & 0x7fffffffu);
    const int count = c16[qp_level_off(R, D0, dl) + c];
    // best key of the leaf: maximum over its depth-D0 cells
    unsigned long long bk = 0;
    const int span = 1 << (2 * (D0 - dl));
    for (int k = 0; k < span; ++k) bk = max(bk, best[(size_t)c * span + k]);
    const uint32_t best_pay = (uint32_t)(bk >> 48) << 24 | (uint32_t)(0xffffffu - (uint32_t)(bk & 0xffffffull));
    // depth at which this list node was created: a single key became alone where its ancestors stop holding one key
    int created = dl;
    if (count == 1)
      while (created > 0 && c16[qp_level_off(R, D0, created - 1) + (c >> (2 * (dl - created + 1)))] == 1) --created;
    const int ca = c >> (2 * (dl - created));  // ancestor cell at that depth
    unsigned long long within;
    if (created == 0) within = (unsigned long long)ca;  // single-key roots keep their push_back order
    else if (created <= d_bfs) within = 0xffffffffull - bfs_creation_key((uint32_t)ca << (2 * (dmax - created)), created, dmax);
    else within = 0xffffffffull - seq[qp_level_off(R, D0, created) + ca];
    nkey[m] = ((unsigned long long)(31 - created) << 40) | within;
    nval[m] = best_pay;
  }
  __syncthreads();
  QT_STAMP(7);
  sort_table<T>(nkey, nval, lp);
  QT_STAMP(8);
  for (int m = tid; m < L; m += T) kept[m] = nval[m];
  if (tid == 0) *kc = L;
  QT_STAMP(9);
#ifdef ORBX_QT_PROFILE2
  if (tid == 0) {
    for (int k = 0; k < 9; ++k) atomicAdd(&g_qt_acc[l][k], (unsigned long long)(qt_st[k + 1] - qt_st[k]));
    atomicAdd(&g_qt_acc[l][9], 1ull);
    atomicAdd(&g_qt_acc[l][10], (unsigned long long)n);
  }
#endif
}

int quadtree_pyramid_smem_bytes(int count_words, int split_words, int NP) {
  return (int)(align_up<size_t>((size_t)count_words * 4 + (size_t)split_words * 4 + (size_t)NP * 17, 16) + sizeof(QtCtl));
}

int quadtree_smem_bytes(int S, int NP) {
  return (int)(align_up<size_t>((size_t)S * 10 + (size_t)NP * 17 + S + 1, 16) + sizeof(QtCtl));
}

int launch_quadtree(const ExtractParams& P, cudaStream_t s, int* launches) {
  NvtxRange nvtx_range("orbx:quadtree");
  int nmax = 1;
  for (int l = 0; l < P.nlevels; ++l) nmax = std::max(nmax, P.lv[l].kept_cap);
  // One or a few frames (at most two CTAs per SM in all): every (frame, level) CTA runs at once and the
  // level-0 CTA is the critical path, so CTAs are 1024 wide and keep up to 8192 keys in shared memory.
  // Batches: 256-thread CTAs with 2048 keys, many per SM (occupancy beats residency there).
  static const int wide_env = getenv("ORBX_QT_WIDE") ? atoi(getenv("ORBX_QT_WIDE")) : -1;
  // (a chunk of the host-batch pipeline -- no grid-barrier counter -- shares the GPU with the chunks around it: 1024-thread CTAs
  // with 80 KB of shared memory would wait for half an SM to drain)
  const bool wide = wide_env >= 0 ? wide_env != 0 : ((long long)P.n_frames * P.nlevels <= 2 * 148 && P.pyr_bar != nullptr);
  const int threads = wide ? kQtThreadsWide : kQtThreads;
  int NP = threads * 2;  // the radix counters ((threads / 32) x 256 x 4 B) alias the node tables (NP x 16 B)
  while (NP < nmax) NP <<= 1;
  int S = wide ? 8192 : 2048;  // levels with more candidates sort through global scratch (L2)
  if (const char* e = getenv("ORBX_QT_SMEM_KEYS")) S = std::max(256, atoi(e));
  while (S & (S - 1)) S &= S - 1;  // power of two
  S = std::max(S, NP);             // node arrays alias nothing, but n_nodes <= n keys
  const int smem = quadtree_smem_bytes(S, NP);
  if (smem > 227 * 1024) return set_error(ORB_E_INVALID, "quadtree: %d bytes of shared memory needed (nfeatures too large)", smem);
  // (measured, round 2: a second launch with twice the shared-memory keys for the levels that overflow S -- the two
  // largest of a textured 640x480 frame -- makes those CTAs 30 % shorter but runs them at 4 per SM: 0.605 vs 0.613 ms
  // for the kernel and a worse overlap with the blur; not kept.  tools/exp_qt_profile2.py has the per-phase cycles.)
  // the sort-free kernel first; the sorting kernel then redoes the (frame, level) pairs it handed back (kept_count == -1)
  static const bool no_pyramid = getenv("ORBX_QT_NO_PYRAMID") != nullptr;
  int cells_max = 0, split_words = 1;
  bool any_level = false;
  for (int l = 0; l < P.nlevels; ++l) {
    int rb = 0;
    while ((1 << rb) < P.lv[l].nIni) ++rb;
    if (P.lv[l].qt_d0 > 0) {
      any_level = true;
      split_words = std::max(split_words, (((1 << rb) << (2 * P.lv[l].qt_d0)) / 4 + 31) / 32);
      if (!P.lv[l].qt_counts_global) cells_max = std::max(cells_max, qp_level_off(1 << rb, P.lv[l].qt_d0, 0) + (1 << rb));
    }
  }
  split_words = (split_words + 3) & ~3;  // the 64-bit node keys follow the bitmap in shared memory
  const int only_flagged = (!no_pyramid && any_level) ? 1 : 0;
  static const int min_ratio4 = getenv("ORBX_QP_RATIO4") ? atoi(getenv("ORBX_QP_RATIO4")) : 12;  // candidates per wanted keypoint x 4 below which the sort-free kernel does not try
  if (only_flagged) {
    const int count_words = (cells_max + 3) / 4 * 2;
    const int psmem = quadtree_pyramid_smem_bytes(count_words, split_words, NP);
    if (wide) {
      ORB_CUDA_TRY(cudaFuncSetAttribute(quadtree_pyramid_kernel<kQtThreadsWide>, cudaFuncAttributeMaxDynamicSharedMemorySize, psmem));
      quadtree_pyramid_kernel<kQtThreadsWide><<<dim3(P.nlevels, P.n_frames), kQtThreadsWide, psmem, s>>>(P, NP, count_words, split_words, min_ratio4);
    } else {
      ORB_CUDA_TRY(cudaFuncSetAttribute(quadtree_pyramid_kernel<kQtThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, psmem));
      quadtree_pyramid_kernel<kQtThreads><<<dim3(P.nlevels, P.n_frames), kQtThreads, psmem, s>>>(P, NP, count_words, split_words, min_ratio4);
    }
    ORB_LAUNCH_CHECK();
    ++*launches;
  }
  if (wide) {
    ORB_CUDA_TRY(cudaFuncSetAttribute(quadtree_select_kernel<kQtThreadsWide>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    quadtree_select_kernel<kQtThreadsWide><<<dim3(P.nlevels, P.n_frames), kQtThreadsWide, smem, s>>>(P, S, NP, only_flagged);
  } else {
    ORB_CUDA_TRY(cudaFuncSetAttribute(quadtree_select_kernel<kQtThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    quadtree_select_kernel<kQtThreads><<<dim3(P.nlevels, P.n_frames), kQtThreads, smem, s>>>(P, S, NP, only_flagged);
  }
  ORB_LAUNCH_CHECK();
  ++*launches;
  return ORB_OK;
}

}  // namespace orbfront

// developer counter: how many (frame, level) pairs needed the sorting kernel since the last call (resets)
extern "C" int orbx_debug_qt_handed_back(unsigned int* out) {
  cudaDeviceSynchronize();
  unsigned int zero = 0;
  if (cudaMemcpyFromSymbol(out, orbfront::g_qt_handed_back, sizeof(zero)) != cudaSuccess) return -1;
  cudaMemcpyToSymbol(orbfront::g_qt_handed_back, &zero, sizeof(zero));
  return 0;
}

#ifdef ORBX_QT_PROFILE2
extern "C" int orbx_debug_qt_acc(unsigned long long* out) {
  cudaDeviceSynchronize();
  static unsigned long long zero[16 * 12];
  cudaMemcpyFromSymbol(out, orbfront::g_qt_acc, sizeof(zero));
  cudaMemcpyToSymbol(orbfront::g_qt_acc, zero, sizeof(zero));
  return 0;
}
#endif
#ifdef ORBX_QT_PROFILE
extern "C" int orbx_debug_qt_prof(long long* out) {
  cudaDeviceSynchronize();
  long long zero[32] = {0};
  cudaMemcpyFromSymbol(out, orbfront::g_qt_prof, sizeof(zero));
  cudaMemcpyToSymbol(orbfront::g_qt_prof, zero, sizeof(zero));
  return 0;
}
#endif

