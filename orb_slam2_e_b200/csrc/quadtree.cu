// Deterministic quadtree keypoint selection (sm_100a): one CTA per (frame, level).
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
quadtree_select_kernel(const __grid_constant__ ExtractParams P, int S, int NP) {
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
    bitonic_sort<T>(nkey, nval, mp);
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
  bitonic_sort<T>(nkey, nval, lp);
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
  const bool wide = wide_env >= 0 ? wide_env != 0 : (long long)P.n_frames * P.nlevels <= 2 * 148;
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
  if (wide) {
    ORB_CUDA_TRY(cudaFuncSetAttribute(quadtree_select_kernel<kQtThreadsWide>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    quadtree_select_kernel<kQtThreadsWide><<<dim3(P.nlevels, P.n_frames), kQtThreadsWide, smem, s>>>(P, S, NP);
  } else {
    ORB_CUDA_TRY(cudaFuncSetAttribute(quadtree_select_kernel<kQtThreads>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    quadtree_select_kernel<kQtThreads><<<dim3(P.nlevels, P.n_frames), kQtThreads, smem, s>>>(P, S, NP);
  }
  ORB_LAUNCH_CHECK();
  ++*launches;
  return ORB_OK;
}

}  // namespace orbfront

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

