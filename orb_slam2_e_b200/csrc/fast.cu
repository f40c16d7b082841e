// FAST-9/16 detection with the per-cell iniTh -> minTh fallback (sm_100a), band-oriented.
//
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree (src/ORBextractor.cc:789-837)
// and the cv::FAST(..., nonmaxSuppression=true) calls inside it.  Semantics reproduced
// (pinned against cv2 4.13 in tests/test_oracle_prims.py):
//   * corner strength S = max over the 16 nine-long arcs of min(v-ring) and of min(ring-v), minus 1;
//     a pixel is a corner at threshold T iff S >= T; response = S (threshold independent);
//   * 3x3 strict non-max suppression among the corners of the cell's scored region only (the cell
//     ROI's outer 3-px ring has no score, so NMS never acts across cell seams);
//   * a cell with no keypoint at iniThFAST is redone at minThFAST (ORBextractor.cc:820-824).
//
// One CTA per BAND = a run of up to ~11 cells of one cell row (FastBand, built on the host).  The
// round-1 kernel (fast_cells.cu) gave every 30-px cell its own warp and paid for it: a 6-px halo on
// all four sides of every cell, per-cell set-up, and a three-stage survivor compaction that cost more
// instructions than the arithmetic it fed.  Here:
//   stage   the band's rows arrive by 1-D TMA bulk copies (cp.async.bulk + mbarrier): one copy per row,
//           issued by one warp; rows are 16-byte aligned because every level pitch is;
//   P       packed prefilter: a thread owns one 4-pixel word column and walks down its rows with the
//           seven rows it needs in a register ring (3 LDS per row instead of 11), VABSDIFF4 against the
//           eight ring positions of four opposite pairs (a nine-long arc holds one pixel of every
//           opposite pair: conservative); the survivor bits of 8 rows x 4 pixels go to one mask word;
//   X       mask words -> survivor queue (warp scan + one shared atomic per warp, bounded queue with a
//           leftover loop) -> exact strength one pixel per lane (packed u16x2 VIMNMX3 network) ->
//           score tile + corner mask;
//   N       corner mask words -> queue -> strict 3x3 NMS with the cell-seam rule -> candidates
//           appended to the (frame, level) list (one global atomic per warp batch), per-cell counts;
//   pass 2  cells without a keypoint repeat P/X/N at minThFAST with every other cell masked out.
// Candidate order in memory is arbitrary; the reference order is a function of (cell, y, x) and is
// re-derived by the quadtree kernel.
#include <algorithm>
#include <cstddef>
#include <cstdlib>

#include "extract.cuh"

namespace orbfront {

constexpr int kBT = 256;     // threads per CTA
constexpr int kBQ = 1024;    // survivor queue entries (u32: bit << 27 | mask word << 15 | row << 9 | tile byte)
constexpr int kMaxCopy = 368;  // bytes staged per tile row (tile pitch <= 416)

// exact strength of the pixel whose centre byte is at p (tile pitch tp)
__device__ __forceinline__ int fast_strength(const uint8_t* p, int tp) {
  // packed lanes: low half = 256 + (v - ring)  (dark arc),  high half = 256 + (ring - v) (bright arc)
  const int v = p[0];
  const uint32_t base = 0x01000100u + (uint32_t)v * 0xFFFF0001u;  // (256+v) | (256-v)<<16; halves stay in [1,511]
  const uint8_t* u1 = p - tp; const uint8_t* u2 = u1 - tp; const uint8_t* u3 = u2 - tp;
  const uint8_t* d1 = p + tp; const uint8_t* d2 = d1 + tp; const uint8_t* d3 = d2 + tp;
  uint32_t d[16];
#define RING(k, q, dx) d[k] = base + (uint32_t)(q)[dx] * 0x0000FFFFu;
  RING(0, d3, 0) RING(1, d3, 1) RING(2, d2, 2) RING(3, d1, 3) RING(4, p, 3) RING(5, u1, 3) RING(6, u2, 2) RING(7, u3, 1)
  RING(8, u3, 0) RING(9, u3, -1) RING(10, u2, -2) RING(11, u1, -3) RING(12, p, -3) RING(13, d1, -3) RING(14, d2, -2) RING(15, d3, -1)
#undef RING
  uint32_t m3[16], best = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) m3[k] = __vimin3_u16x2(d[k], d[(k + 1) & 15], d[(k + 2) & 15]);
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const uint32_t a = __vimin3_u16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]);
    const uint32_t b = __vimin3_u16x2(m3[k + 1], m3[(k + 4) & 15], m3[(k + 7) & 15]);
    best = __vimax3_u16x2(best, a, b);
  }
  return (int)max(best & 0xffffu, best >> 16) - 256 - 1;
}

struct BandCtl {
  FastBand band;         // the current item's band (read from here: fewer long-lived registers); 16-byte aligned, written as two quads
  uint64_t bar_tile[2];  // tile buffers (TMA row copies)
  uint64_t bar_zero;     // score tile clear (TMA copy of zeros)
  uint32_t empty;        // cells of the band without a keypoint after the first pass
  uint32_t pad;
  int qcount[2];
  int cellcnt[30];
  uint32_t lut[32];      // queue entry bits of a mask bit: (bit & 7) << 9 | bit >> 3 | bit << 27
};
static_assert(offsetof(BandCtl, band) % 16 == 0 && offsetof(BandCtl, bar_tile) % 8 == 0, "BandCtl layout");

__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}

__device__ __forceinline__ FastBand load_band(const FastBand* p) {
  FastBand B;
  const uint4* bp = reinterpret_cast<const uint4*>(p);
  uint4* bd = reinterpret_cast<uint4*>(&B);
  bd[0] = __ldg(bp);
  bd[1] = __ldg(bp + 1);
  return B;
}

// Persistent: CTA b works on items b, b + gridDim.x, ... where item = frame * n_bands + band.  While item i is
// processed, the rows of item i+1 stream into the other tile buffer, and the score tile is cleared by a
// TMA copy of zeros while the next prefilter (which does not touch it) runs.
__global__ void __launch_bounds__(kBT, 4)
fast_band_kernel(const __grid_constant__ ExtractParams P) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* const tile0 = smem;
  uint8_t* const score = tile0 + 2 * P.fast_tile_bytes;
  uint32_t* const maskw = reinterpret_cast<uint32_t*>(score + P.fast_score_bytes);  // [round][thread]: prefilter survivors
  uint32_t* const cmask = maskw + P.fast_mask_words;                               // same layout: corners (score >= T)
  uint32_t* const queue = cmask + P.fast_mask_words;
  BandCtl& C = *reinterpret_cast<BandCtl*>(queue + kBQ);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_items = P.n_bands * P.n_frames;
  int item = blockIdx.x;
  if (item >= n_items) return;
  // item -> (frame, band) by stepping: no division in the loop
  const int step_f = (int)gridDim.x / P.n_bands, step_b = (int)gridDim.x - step_f * P.n_bands;
  int f = item / P.n_bands, bi = item - f * P.n_bands;

  if (tid == 0) {
    mbar_init(&C.bar_tile[0], 1);
    mbar_init(&C.bar_tile[1], 1);
    mbar_init(&C.bar_zero, 1);
    fence_barrier_init();
  }
  for (int i = tid; i < P.fast_mask_words; i += kBT) cmask[i] = 0;  // N clears what X sets: stays zero between items
  if (tid < 32) { C.qcount[tid & 1] = 0; C.lut[tid] = (uint32_t)((tid & 7) << 9 | tid >> 3) | (uint32_t)tid << 27; }
  __syncthreads();

  // rows [row0 - 3, row0 + vh + 3) x columns [x_lo, x_lo + 16 * copy16) of band Bn -> tile buffer `buf` (warp 0)
  auto request_tile = [&](const FastBand& Bn, int fn, int buf) {
    const int ln = Bn.level, sp = level_pitch(P, ln);
    const uint8_t* src = level_ptr(P, ln, fn) + (size_t)(Bn.row0 - 3) * sp + Bn.x_lo;
    uint8_t* dst = tile0 + buf * P.fast_tile_bytes + 16;
    const uint32_t row_bytes = 16u * Bn.copy16;
    fence_proxy_async();  // the buffer's previous contents were read through the generic proxy
    if (lane == 0) mbar_expect_tx(&C.bar_tile[buf], row_bytes * (uint32_t)(Bn.vh + 6));
    __syncwarp();
    for (int r = lane; r < Bn.vh + 6; r += 32) tma_load_1d(dst + r * Bn.tp, src + (size_t)r * sp, row_bytes, &C.bar_tile[buf]);
  };
  auto request_zero = [&](const FastBand& Bn) {  // one thread
    const uint32_t bytes = (uint32_t)((Bn.vh + 2) * Bn.tp);  // a multiple of 16
    fence_proxy_async();
    mbar_expect_tx(&C.bar_zero, bytes);
    tma_load_1d(score, P.fast_zeros, bytes, &C.bar_zero);
  };
  if (warp == 0) {
    const FastBand B0 = load_band(P.fast_bands + bi);
    request_tile(B0, f, 0);
    if (lane == 0) request_zero(B0);
  }

  int qpar = 0;  // which qcount the next queue fill uses (that one is zero)
#pragma unroll 1
  for (int it = 0; item < n_items; ++it, item += gridDim.x) {
    const int buf = it & 1;
    const bool has_next = item + (int)gridDim.x < n_items;
    if (tid < 2) reinterpret_cast<uint4*>(&C.band)[tid] = __ldg(reinterpret_cast<const uint4*>(P.fast_bands + bi) + tid);
    if (warp == 0) {
      if (has_next) {  // the other buffer's reader (item it-1) is past its last barrier
        int fn = f + step_f, bn = bi + step_b;
        if (bn >= P.n_bands) { bn -= P.n_bands; ++fn; }
        request_tile(load_band(P.fast_bands + bn), fn, buf ^ 1);
      }
      while (!mbar_test(&C.bar_tile[buf], (uint32_t)(it >> 1) & 1u)) { }
    }
    if (tid < 30) C.cellcnt[tid] = 0;
    __syncthreads();
    while (!mbar_test(&C.bar_tile[buf], (uint32_t)(it >> 1) & 1u)) { }  // passes at once: every thread observes the phase itself
    const FastBand& B = C.band;

    const int l = B.level, tp = B.tp, tpw = tp >> 2, vh = B.vh;
    const LevelGeom& G = P.lv[l];
    const int wCell = G.wCell;
    const uint32_t mag_w = B.mag_w;
    const uint8_t* const tile = tile0 + buf * P.fast_tile_bytes;
    // this thread's prefilter item: word column c of row split s
    const bool has_item = tid < B.ncw * B.R;
    const int it_s = has_item ? (int)(((uint32_t)tid * B.mag_ncw) >> 16) : 0, it_c = tid - it_s * B.ncw;
    const int r0 = it_s * B.rps, r1 = has_item ? min(r0 + B.rps, vh) : 0;
    const int xbyte0 = 4 * (B.wlo + it_c);  // tile byte of the item's first pixel
    const int nrounds = B.nrounds;

    // This thread's mask words M[k * kBT + tid], k < nrounds -> bounded queue -> proc(active, entry) with all 32
    // lanes of a warp converged.  A thread unpacks its words back to back (one warp scan and one shared
    // atomic per fill); bits that do not fit the queue wait for the next fill.
    auto drain_masks = [&](uint32_t* M, const bool clear, auto&& proc) {
      int remaining = 0;
      if (has_item)
        for (int k = 0; k < nrounds; ++k) remaining += __popc(M[k * kBT + tid]);
      int k = -1;
      uint32_t mm = 0, eb = 0;
      while (true) {
        int incl = remaining;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
          const int o = __shfl_up_sync(0xffffffffu, incl, s);
          if (lane >= s) incl += o;
        }
        int wbase = 0;
        if (lane == 31 && incl) wbase = atomicAdd(&C.qcount[qpar], incl);
        wbase = __shfl_sync(0xffffffffu, wbase, 31);
        const int pos = wbase + incl - remaining;
        const int take = min(remaining, max(kBQ - pos, 0));
        for (int t = 0; t < take; ++t) {
          while (mm == 0) {
            ++k;
            mm = M[k * kBT + tid];
            if (clear) M[k * kBT + tid] = 0;
            eb = (uint32_t)(k * kBT + tid) << 15 | (uint32_t)(r0 + 8 * k) << 9 | (uint32_t)xbyte0;
          }
          const int bit = __ffs(mm) - 1;
          mm &= mm - 1;
          queue[pos + t] = eb + C.lut[bit];
        }
        remaining -= take;
        __syncthreads();
        const int cnt = min(C.qcount[qpar], kBQ);
        if (tid == 0) C.qcount[qpar ^ 1] = 0;  // nobody reads or adds to the other counter before the next barrier
        for (int b0 = warp * 32; b0 < cnt; b0 += kBT) {
          const int i = b0 + lane;
          const bool act = i < cnt;
          proc(act, act ? queue[i] : 0u);
        }
        qpar ^= 1;
        if (!__syncthreads_or(remaining != 0)) break;
      }
    };

    // Detection passes: threshold T over the pixels of the cells in `cells` (bit k = cell k of the band);
    // first iniThFAST on every cell, then minThFAST on the cells that found nothing (ORBextractor.cc:820).
    // One loop body so that the code exists once (instruction cache).
    int T = P.ini_th;
    uint32_t cells = B.nc >= 32 ? 0xffffffffu : (1u << B.nc) - 1u;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
      // ---- P: prefilter ----
      if (has_item) {
        uint32_t vm = 0;  // bit 7 of byte b: pixel b of this word column is scored in this pass
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const int xt = xbyte0 + b;
          if (xt >= B.xa_t && xt < B.xb_t) {
            const int cj = (int)(((uint32_t)(xt - B.xa_t) * mag_w) >> 20);
            if (cells >> cj & 1u) vm |= 0x80u << (8 * b);
          }
        }
        uint32_t* mw = maskw + tid;
        if (vm == 0 || r0 >= r1) {
          for (int k = 0; k < nrounds; ++k) mw[k * kBT] = 0;
        } else {
          const uint32_t K = 0x01010101u * (uint32_t)(127 - min(T, 126));
          // ring slots are indexed by (tile row - r0) mod 8 (mod 4 for the +-3 column shifts of the centre row)
          uint32_t Cc[8], A2[8], B2[8], A3[4], B3[4];
          const uint32_t* tw = reinterpret_cast<const uint32_t*>(tile) + r0 * tpw + (B.wlo + it_c);
          auto load_row = [&](int slot) {  // consumes the row at tw, advances tw
            const uint32_t lw = tw[-1], cw = tw[0], rw = tw[1];
            tw += tpw;
            Cc[slot & 7] = cw;
            A2[slot & 7] = __funnelshift_r(lw, cw, 16);  // pixels x-2
            B2[slot & 7] = __funnelshift_r(cw, rw, 16);  // pixels x+2
            A3[slot & 3] = __funnelshift_r(lw, cw, 8);   // pixels x-3
            B3[slot & 3] = __funnelshift_r(cw, rw, 24);  // pixels x+3
          };
#pragma unroll
          for (int j = 0; j < 6; ++j) load_row(j);
          for (int k = 0; k < nrounds; ++k) {
            uint32_t acc = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (r0 + 8 * k + j < r1) {
                load_row(j + 6);  // tile row of level row y + 3
                const uint32_t c = Cc[(j + 3) & 7];
                const uint32_t d0 = __vabsdiffu4(c, Cc[(j + 6) & 7]), d1 = __vabsdiffu4(c, Cc[j & 7]);          // (0,3) | (0,-3)
                const uint32_t d2 = __vabsdiffu4(c, B3[(j + 3) & 3]), d3 = __vabsdiffu4(c, A3[(j + 3) & 3]);    // (3,0) | (-3,0)
                const uint32_t d4 = __vabsdiffu4(c, B2[(j + 5) & 7]), d5 = __vabsdiffu4(c, A2[(j + 1) & 7]);    // (2,2) | (-2,-2)
                const uint32_t d6 = __vabsdiffu4(c, B2[(j + 1) & 7]), d7 = __vabsdiffu4(c, A2[(j + 5) & 7]);    // (2,-2) | (-2,2)
                // bit 7 of every byte of ((x & 0x7f..) + K) | x : x > T
                const uint32_t t0 = (((d0 | d1) & 0x7f7f7f7fu) + K) | d0 | d1;
                const uint32_t t1 = (((d2 | d3) & 0x7f7f7f7fu) + K) | d2 | d3;
                const uint32_t t2 = (((d4 | d5) & 0x7f7f7f7fu) + K) | d4 | d5;
                const uint32_t t3 = (((d6 | d7) & 0x7f7f7f7fu) + K) | d6 | d7;
                const uint32_t fl = (t0 & t1 & t2) & (t3 & vm);
                acc |= fl >> (7 - j);  // bit 8b + j = pixel b of row j of this round
              }
            }
            mw[k * kBT] = acc;
          }
        }
      }
      if (pass == 0) {  // the score tile's clear, requested before this prefilter started
        if (warp == 0)
          while (!mbar_test(&C.bar_zero, (uint32_t)it & 1u)) { }
        __syncthreads();
        while (!mbar_test(&C.bar_zero, (uint32_t)it & 1u)) { }
      }

      // ---- X: exact strength of the survivors ----
      drain_masks(maskw, false, [&](bool act, uint32_t e) {
        if (act) {
          const int row = (int)(e >> 9) & 63, xb = (int)(e & 511u);
          const int s = fast_strength(tile + (row + 3) * tp + xb, tp);
          if (s >= T) {
            score[(row + 1) * tp + xb] = (uint8_t)s;
            atomicOr(&cmask[(e >> 15) & 2047u], 1u << (e >> 27));
          }
        }
      });

      // ---- N: strict 3x3 non-max suppression inside each cell, survivors appended ----
      drain_masks(cmask, true, [&](bool act, uint32_t e) {
        bool is_max = false;
        uint32_t pay = 0;
        if (act) {
          const int row = (int)(e >> 9) & 63, xb = (int)(e & 511u);
          const uint8_t* q = score + (row + 1) * tp + xb;
          const int s = q[0];
          const int xl = xb - B.xa_t;
          const int cj = (int)(((uint32_t)xl * mag_w) >> 20);
          const int loc = xl - cj * wCell;
          // pixels of the neighbouring cell do not exist for this cell's FAST call
          const int lm = loc == 0 ? 0 : 255, rm = loc == wCell - 1 ? 0 : 255;
          const int nl = max(max((int)q[-tp - 1], (int)q[-1]), (int)q[tp - 1]) & lm;
          const int nr = max(max((int)q[-tp + 1], (int)q[1]), (int)q[tp + 1]) & rm;
          const int nc = max((int)q[-tp], (int)q[tp]);
          is_max = s > max(max(nl, nr), nc);
          if (is_max) {
            atomicAdd(&C.cellcnt[cj], 1);
            const int x = (int)B.x_lo - 16 + xb - (kEdge - 3), y = (int)B.row0 + row - (kEdge - 3);  // border-relative (minBorder = 16)
            pay = ((uint32_t)s << 24) | ((uint32_t)y << 12) | (uint32_t)x;
          }
        }
        const uint32_t m = __ballot_sync(0xffffffffu, is_max);
        if (m) {
          int slot = 0;
          if (lane == 0) slot = atomicAdd(&P.cand_count[f * P.nlevels + l], __popc(m));
          slot = __shfl_sync(0xffffffffu, slot, 0) + __popc(m & ((1u << lane) - 1u));
          if (is_max && slot < G.cand_cap) (P.cand + (size_t)f * P.cand_frame + G.cand_off)[slot] = pay;  // the quadtree kernel prepends the path code
        }
      });
      if (pass == 1 || P.min_th >= P.ini_th) break;
      if (warp == 0) {
        const uint32_t e = __ballot_sync(0xffffffffu, lane < B.nc && C.cellcnt[lane] == 0);
        if (lane == 0) C.empty = e;
      }
      __syncthreads();
      const uint32_t empty = C.empty;
      if (!empty) break;
      T = P.min_th;
      cells = empty;
    }
    // end of item: every read of this tile buffer and of the score tile is done
    __syncthreads();
    f += step_f;
    bi += step_b;
    if (bi >= P.n_bands) { bi -= P.n_bands; ++f; }
    if (tid == 0 && has_next) request_zero(load_band(P.fast_bands + bi));
  }
}

// ------------------------------------------------------------------------------------
int fast_build_bands(ExtractParams& P, std::vector<FastBand>& bands) {
  bands.clear();
  int tile_bytes = 0, score_bytes = 0, mask_words = 0;
  for (int l = 0; l < P.nlevels; ++l) {
    const LevelGeom& G = P.lv[l];
    if (G.wCell >= 64 || G.hCell >= 60) return set_error(ORB_E_INVALID, "FAST cell %dx%d exceeds 63x59", G.wCell, G.hCell);
    const int per_chunk_max = std::max(1, std::min(30, (kMaxCopy - 6 - 15) / G.wCell));
    const int n_chunks = div_up(G.nCols, per_chunk_max), per_chunk = div_up(G.nCols, n_chunks);
    for (int ci = 0; ci < G.nRows; ++ci) {
      const int row0 = kEdge + ci * G.hCell, vh = std::min(G.hCell, G.h - kEdge - row0);
      if (vh <= 0) continue;
      for (int c0 = 0; c0 < G.nCols; c0 += per_chunk) {
        const int nc = std::min(per_chunk, G.nCols - c0);
        const int xa = kEdge + c0 * G.wCell, xb = std::min(kEdge + (c0 + nc) * G.wCell, G.w - kEdge);
        if (xb <= xa) continue;
        FastBand B = {};
        B.level = (uint16_t)l; B.row0 = (uint16_t)row0; B.vh = (uint16_t)vh;
        const int x_lo = (xa - 3) & ~15;
        const int copy = (int)align_up<size_t>((size_t)(xb + 3 - x_lo), 16);
        B.x_lo = (uint16_t)x_lo; B.copy16 = (uint16_t)(copy / 16);
        B.xa_t = (uint16_t)(xa - x_lo + 16); B.xb_t = (uint16_t)(xb - x_lo + 16);
        B.nc = (uint16_t)div_up(xb - xa, G.wCell);
        B.wlo = (uint16_t)(B.xa_t / 4);
        B.ncw = (uint16_t)((B.xb_t - 1) / 4 - B.wlo + 1);
        int tp = 16 + copy;  // left padding + staged bytes
        while (tp % 64 != 32) tp += 16;  // pitch = 8 (mod 16) words: rows of a column fall into different banks
        B.tp = (uint16_t)tp;
        B.mag_w = (uint16_t)((1u << 20) / (uint32_t)G.wCell + 1u);
        B.mag_ncw = (uint16_t)(65536u / (uint32_t)B.ncw + 1u);
        if (B.ncw > kBT || B.ncw < 8 || tp > 511) return set_error(ORB_E_INVALID, "FAST band too wide (%d words, pitch %d)", B.ncw, tp);
        int R = std::max(1, std::min(kBT / (int)B.ncw, div_up(vh, 6)));  // at least ~6 rows per split: each split re-reads 6 rows
        int rps = div_up(vh, R);
        R = div_up(vh, rps);
        B.R = (uint16_t)R; B.rps = (uint16_t)rps; B.nrounds = (uint16_t)div_up(rps, 8);
        const int words = kBT * B.nrounds;  // [round][thread]
        if (B.nrounds > 8) return set_error(ORB_E_INVALID, "FAST band has %d rounds per split (limit 8)", (int)B.nrounds);
        tile_bytes = std::max(tile_bytes, (vh + 6) * tp);
        score_bytes = std::max(score_bytes, (vh + 2) * tp);
        mask_words = std::max(mask_words, words);
        bands.push_back(B);
      }
    }
  }
  P.n_bands = (int)bands.size();
  P.fast_tile_bytes = (int)align_up<size_t>((size_t)tile_bytes + 16, 128);  // the last row's right neighbour word may be read
  P.fast_score_bytes = (int)align_up<size_t>((size_t)score_bytes + 16, 128);
  P.fast_mask_words = (int)align_up<size_t>((size_t)mask_words, 4);
  return ORB_OK;
}

size_t fast_band_smem(const ExtractParams& P) {
  return 2 * (size_t)P.fast_tile_bytes + P.fast_score_bytes + 2 * sizeof(uint32_t) * (size_t)P.fast_mask_words + sizeof(uint32_t) * kBQ +
         sizeof(BandCtl) + 16;
}

// CTAs of the persistent FAST kernel that are resident at once on `device`
int fast_band_grid(const ExtractParams& P, int device) {
  int per_sm = 0, sms = 0;
  const size_t smem = fast_band_smem(P);
  if (cudaFuncSetAttribute(fast_band_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess ||
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fast_band_kernel, kBT, smem) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess)
    return 0;
  return per_sm * sms;
}

int launch_fast(const ExtractParams& P, cudaStream_t s, int* launches) {
  static const bool cells = getenv("ORBX_FAST_CELLS") && atoi(getenv("ORBX_FAST_CELLS")) != 0;  // developer knob: round-1 kernel
  if (cells) return launch_fast_cells(P, s, launches);
  if (P.n_bands <= 0 || !P.fast_bands || !P.fast_zeros) return set_error(ORB_E_STATE, "FAST band table missing");
  ORB_CUDA_TRY(cudaMemsetAsync(P.cand_count, 0, sizeof(int) * (size_t)P.n_frames * P.nlevels, s));
  const size_t smem = fast_band_smem(P);
  if (smem > 227 * 1024) return set_error(ORB_E_INVALID, "FAST: %zu bytes of shared memory needed", smem);
  ORB_CUDA_TRY(cudaFuncSetAttribute(fast_band_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long items = (long long)P.n_bands * P.n_frames;
  const int grid = (int)std::min<long long>(items, P.fast_grid > 0 ? P.fast_grid : 4 * 148);
  fast_band_kernel<<<grid, kBT, smem, s>>>(P);
  ORB_LAUNCH_CHECK();
  ++*launches;
  return ORB_OK;
}

}  // namespace orbfront
