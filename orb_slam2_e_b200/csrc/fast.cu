// FAST-9/16 detection per 30-px cell with the iniTh -> minTh fallback (sm_100a).
//
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree (src/ORBextractor.cc:789-837)
// and the cv::FAST(..., nonmaxSuppression=true) calls inside it.  Semantics reproduced
// (pinned against cv2 4.13 in tests/test_oracle_prims.py):
//   * corner strength S = max over the 16 nine-long arcs of min(v-ring) and of min(ring-v), minus 1;
//     a pixel is a corner at threshold T iff S >= T; response = S (threshold independent);
//   * 3x3 strict non-max suppression among the corners of the cell's valid region only (the
//     cell ROI's outer 3-px ring has no score, so NMS never acts across cell seams);
//   * a cell with no keypoint at iniThFAST is redone at minThFAST (ORBextractor.cc:820-824).
// One warp per cell, and -- like the reference -- one detection pass at iniThFAST followed by a
// second at minThFAST only for empty cells.  A pass:
//   1. packed prefilter, 4 pixels per lane: VABSDIFF4 of the centre word against the eight ring
//      positions of four opposite pairs; a nine-long arc holds one pixel of every opposite pair,
//      so a corner needs |centre - ring| > T on at least one side of each pair (conservative:
//      signs and the bitwise-OR "max" only let extra pixels through);
//   2. survivors are compacted into a ring queue and get the exact strength, one pixel per lane,
//      through a packed u16x2 VIMNMX3 network (dark arcs low half, bright arcs high half);
//   3. NMS on a zero-framed score tile, survivors appended (any order) to the (frame, level)
//      candidate list together with their quadtree path code.
#include <algorithm>

#include "extract.cuh"

namespace orbfront {

constexpr int kFastWarps = 4;
constexpr int kCellMax = 60;                 // wCell,hCell <= 59 (ceil(width/floor(width/30)))
constexpr int kQueue = 256;                  // ring queue entries (py << 6 | px)
// x / d for small operands without the ~20-instruction integer division: (x * (65536 / d + 1)) >> 16 is exact while x * d < 65536
__device__ __constant__ uint32_t c_rcp16[32] = {0, 65537, 32769, 21846, 16385, 13108, 10923, 9363, 8193, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4097, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115};
__device__ __forceinline__ int small_div(int x, int d) { return (int)(((uint32_t)x * c_rcp16[d]) >> 16); }
// Per-warp shared memory: tile[(maxH+6) x PITCH] | score[(maxH+2) x PITCH] | queue[kQueue] u16, later reused
// for the cell's NMS survivors | wlist (score words with a prefilter survivor).
// PITCH (bytes, multiple of 8, >= widest cell + 8) is a template parameter so the 16 ring offsets
// stay immediates; tile/score column = px + 4 (word aligned groups of 4 pixels).

// exact strength of the pixel whose centre byte is at p (tile pitch kTilePitch)
template <int kTilePitch>
__device__ __forceinline__ int fast_strength(const uint8_t* p) {
  // packed lanes: low half = 256 + (v - ring)  (dark arc),  high half = 256 + (ring - v) (bright arc)
  const int v = p[0];
  const uint32_t base = 0x01000100u + (uint32_t)v * 0xFFFF0001u;  // (256+v) | (256-v)<<16; halves stay in [1,511]
  uint32_t d[16];
#define RING(k, dx, dy) d[k] = base + (uint32_t)p[(dy) * kTilePitch + (dx)] * 0x0000FFFFu;
  RING(0, 0, 3) RING(1, 1, 3) RING(2, 2, 2) RING(3, 3, 1) RING(4, 3, 0) RING(5, 3, -1) RING(6, 2, -2) RING(7, 1, -3)
  RING(8, 0, -3) RING(9, -1, -3) RING(10, -2, -2) RING(11, -3, -1) RING(12, -3, 0) RING(13, -3, 1) RING(14, -2, 2) RING(15, -1, 3)
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

template <int kTilePitch>
__global__ void __launch_bounds__(kFastWarps * 32, 10)
fast_cells_kernel(const __grid_constant__ ExtractParams P, int max_h, int warp_bytes) {
  const int qk_entries = max(kQueue, (((kTilePitch - 8) / 2 + 1) * (max_h / 2 + 1) + 1) & ~1);
  constexpr int kTileWords = kTilePitch / 4, kScorePitch = kTilePitch;
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint8_t* tile8 = smem_raw + (threadIdx.x >> 5) * warp_bytes;
  uint8_t* score8 = tile8 + (max_h + 6) * kTilePitch;
  struct { uint32_t* tile; uint32_t* score; uint16_t* queue; } S;
  S.tile = reinterpret_cast<uint32_t*>(tile8);
  S.score = reinterpret_cast<uint32_t*>(score8);
  S.queue = reinterpret_cast<uint16_t*>(score8 + (max_h + 2) * kTilePitch);
  uint16_t* kplist = S.queue;  // NMS survivors of the cell (<= one per 2x2 block); the queue is idle by then
  uint16_t* wlist = S.queue + qk_entries;  // score words holding a corner: (max_h) x (PITCH-8)/4 entries
  const int lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int cell = blockIdx.x * kFastWarps + (threadIdx.x >> 5);
  if (cell >= P.total_cells) return;
  const int f = blockIdx.y;
  int l = 0;
  while (l + 1 < P.nlevels && cell >= P.lv[l + 1].cell_base) ++l;
  const LevelGeom& G = P.lv[l];
  const int ci = G.nCols == 1 ? cell - G.cell_base : (int)__umulhi((uint32_t)(cell - G.cell_base), G.rcp_nCols);  // / nCols
  const int cj = (cell - G.cell_base) - ci * G.nCols;
  // valid (scored) region of the cell in level coordinates
  const int vx0 = kEdge + cj * G.wCell, vy0 = kEdge + ci * G.hCell;
  const int vw = min(G.wCell, G.w - kEdge - vx0), vh = min(G.hCell, G.h - kEdge - vy0);
  if (vw <= 0 || vh <= 0) return;

  // stage columns [vx0-4, vx0+vw+4) x rows [vy0-3, vy0+vh+3): pixel (px,py) sits at tile (py+3, px+4)
  const int sp = level_pitch(P, l);
  const uint8_t* src = level_ptr(P, l, f) + (size_t)(vy0 - 3) * sp + (vx0 - 4);
  const int tw = min(vw + 8, G.w - (vx0 - 4)), th = vh + 6;
  {
    // quad-wise: every base and pitch is a multiple of 16, so all rows share the byte shift `sh` and the
    // word shift `wsh` of the tile origin inside its 16-byte quad.  A lane loads the two aligned quads
    // around four tile words, funnel-shifts the five words they span (the word shift is uniform per
    // cell: a four-way switch, static register indices) and stores them as two 8-byte halves.
    // A quad is loaded only if it starts inside the level row, so no load leaves the row whatever
    // its pitch; rows per pass = 32 / quads per row.
    const uintptr_t a0 = reinterpret_cast<uintptr_t>(src);
    const int sh8 = (int)(a0 & 3) * 8, wsh = (int)(a0 >> 2) & 3;
    const uint4* qsrc = reinterpret_cast<const uint4*>(a0 & ~(uintptr_t)15);
    const int xq0 = (vx0 - 4) - (int)(a0 & 15);  // level column of the first quad (>= 0: rows start 16-byte aligned)
    const int spq = sp >> 4, nq = (tw + 15) >> 4;  // nq <= 5
    const int rpi = small_div(32, nq);
    const int lr = small_div(lane, nq), lq = lane - lr * nq;
    if (lr < rpi) {
      const bool in0 = xq0 + 16 * lq < G.w, in1 = xq0 + 16 * lq + 16 < G.w && (sh8 | wsh) != 0;
      const bool st0 = 4 * lq < kTileWords, st1 = 4 * lq + 2 < kTileWords;  // PITCH is a multiple of 8
      const uint4* a = qsrc + (size_t)lr * spq + lq;
      uint2* d = reinterpret_cast<uint2*>(S.tile + lr * kTileWords + 4 * lq);
      const int astep = rpi * spq, dstep = rpi * (kTileWords / 2);
#pragma unroll 2
      for (int r = lr; r < th; r += rpi) {
        const uint4 q0 = in0 ? __ldg(a) : make_uint4(0, 0, 0, 0);
        const uint4 q1 = in1 ? __ldg(a + 1) : make_uint4(0, 0, 0, 0);
        uint32_t w0, w1, w2, w3, w4;
        switch (wsh) {
          case 0: w0 = q0.x; w1 = q0.y; w2 = q0.z; w3 = q0.w; w4 = q1.x; break;
          case 1: w0 = q0.y; w1 = q0.z; w2 = q0.w; w3 = q1.x; w4 = q1.y; break;
          case 2: w0 = q0.z; w1 = q0.w; w2 = q1.x; w3 = q1.y; w4 = q1.z; break;
          default: w0 = q0.w; w1 = q1.x; w2 = q1.y; w3 = q1.z; w4 = q1.w; break;
        }
        if (st0) d[0] = make_uint2(__funnelshift_r(w0, w1, sh8), __funnelshift_r(w1, w2, sh8));
        if (st1) d[1] = make_uint2(__funnelshift_r(w2, w3, sh8), __funnelshift_r(w3, w4, sh8));
        a += astep;
        d += dstep;
      }
    }
  }
  __syncwarp();

  const int ng = (vw + 3) >> 2;  // 4-pixel groups per row
  int* const count = &P.cand_count[f * P.nlevels + l];
  unsigned long long* const out = P.cand + (size_t)f * P.cand_frame + G.cand_off;

  // zero frame of the score tile (rows 0 and vh+1, words 0 and ng+1 of every row): written once per
  // cell; the interior words are cleared by each pass's prefilter loop as it visits them
  for (int i = lane; i < kScorePitch / 4; i += 32) { S.score[i] = 0; S.score[(vh + 1) * (kScorePitch / 4) + i] = 0; }
  for (int r = 1 + lane; r <= vh; r += 32) { S.score[r * (kScorePitch / 4)] = 0; S.score[r * (kScorePitch / 4) + ng + 1] = 0; }

  // prefilter lane constants: pair gp = groups 2gp, 2gp+1 of row lr (+ rpi per pass)
  const int ngp = (ng + 1) >> 1, rpi = small_div(32, ngp);  // ng <= 15: at least 4 rows per pass
  const int lr = small_div(lane, ngp), gp = lane - lr * ngp;
  const bool lane_on = lr < rpi;
  const uint32_t* const prow = &S.tile[(lr + 3) * kTileWords + 2 * gp];
  uint32_t* const srow = &S.score[(lr + 1) * (kScorePitch / 4) + 2 * gp + 1];
  auto valid_mask = [&](int g) {  // bit 7 of the bytes of group g that lie inside the cell
    const int nvalid = min(4, vw - 4 * g);
    return nvalid <= 0 ? 0u : 0x80808080u >> (32 - 8 * nvalid);
  };
  const uint32_t mask0 = valid_mask(2 * gp), mask1 = valid_mask(2 * gp + 1);
  const uint32_t ent0 = (uint32_t)(2 * gp) << 4;
  // One detection pass at threshold T; returns the number of keypoints it appended.
  auto detect = [&](const int T) -> int {
    const uint32_t K = 0x01010101u * (uint32_t)(127 - min(T, 126));
    auto gt = [&](uint32_t x) { return ((x & 0x7f7f7f7fu) + K) | x; };  // bit 7 of every byte: x > T
    int head = 0, tail = 0;  // ring queue (uniform across the warp)
    int nw = 0;              // entries of wlist
    auto drain = [&](int count_) {
      if (lane < count_) {
        const int e = S.queue[(head + lane) & (kQueue - 1)];
        const int py = e >> 6, px = e & 63;
        const int s = fast_strength<kTilePitch>(&tile8[(py + 3) * kTilePitch + px + 4]);
        if (s >= T) score8[(py + 1) * kScorePitch + px + 4] = (uint8_t)s;
      }
      head += count_;
    };
    int whead = 0;  // listed words already unpacked into the pixel queue
    auto expand = [&](int count_) {
      int e = 0, n = 0;
      if (lane < count_) { e = wlist[whead + lane]; n = __popc(e & 15); }
      int incl = n;
#pragma unroll
      for (int s = 1; s < 32; s <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, s);
        if (lane >= s) incl += o;
      }
      int pos = tail + incl - n;
      const int pix = (e >> 8) << 6 | ((e >> 4) & 15) << 2;
#pragma unroll
      for (int b = 0; b < 4; ++b)
        if (e >> b & 1) S.queue[(pos++) & (kQueue - 1)] = (uint16_t)(pix + b);
      tail += __shfl_sync(0xffffffffu, incl, 31);
      whead += count_;
      __syncwarp();
      while (tail - head >= 32) drain(32);
    };
    __syncwarp();
    // Prefilter, two horizontally adjacent 4-pixel groups per lane: lane <-> (row lr of the pass, pair gp), rows
    // advance by rpi per pass, so the loop carries two pointers instead of deriving (row, group) from a flat index.
    // The pair shares its loads (six 8-byte and four 4-byte reads instead of 22 words) and two of its twelve
    // funnel shifts ((x+2) of the left group is (x-2) of the right one).
    for (int base = 0; base < vh; base += rpi) {
      const int py = base + lr;
      uint32_t flags0 = 0, flags1 = 0;
      if (lane_on && py < vh) {
        const uint32_t* r0 = prow + base * kTileWords;  // word 2gp of the centre row: (lw, c0) (c1, rw)
        const uint2 m0 = *reinterpret_cast<const uint2*>(r0), m1 = *reinterpret_cast<const uint2*>(r0 + 2);
        const uint2 u0 = *reinterpret_cast<const uint2*>(r0 - 2 * kTileWords), u1 = *reinterpret_cast<const uint2*>(r0 - 2 * kTileWords + 2);
        const uint2 d0 = *reinterpret_cast<const uint2*>(r0 + 2 * kTileWords), d1 = *reinterpret_cast<const uint2*>(r0 + 2 * kTileWords + 2);
        const uint32_t t0 = r0[-3 * kTileWords + 1], t1 = r0[-3 * kTileWords + 2];
        const uint32_t b0 = r0[3 * kTileWords + 1], b1 = r0[3 * kTileWords + 2];
        const uint32_t c0 = m0.y, c1 = m1.x;
        const uint32_t umid = __funnelshift_r(u0.y, u1.x, 16), dmid = __funnelshift_r(d0.y, d1.x, 16);
        uint32_t fl = gt(__vabsdiffu4(c0, t0) | __vabsdiffu4(c0, b0));
        fl &= gt(__vabsdiffu4(c0, __funnelshift_r(c0, c1, 24)) | __vabsdiffu4(c0, __funnelshift_r(m0.x, c0, 8)));
        fl &= gt(__vabsdiffu4(c0, dmid) | __vabsdiffu4(c0, __funnelshift_r(u0.x, u0.y, 16)));
        fl &= gt(__vabsdiffu4(c0, umid) | __vabsdiffu4(c0, __funnelshift_r(d0.x, d0.y, 16)));
        flags0 = fl & mask0;
        fl = gt(__vabsdiffu4(c1, t1) | __vabsdiffu4(c1, b1));
        fl &= gt(__vabsdiffu4(c1, __funnelshift_r(c1, m1.y, 24)) | __vabsdiffu4(c1, __funnelshift_r(c0, c1, 8)));
        fl &= gt(__vabsdiffu4(c1, __funnelshift_r(d1.x, d1.y, 16)) | __vabsdiffu4(c1, umid));
        fl &= gt(__vabsdiffu4(c1, __funnelshift_r(u1.x, u1.y, 16)) | __vabsdiffu4(c1, dmid));
        flags1 = fl & mask1;
        uint32_t* sc = srow + base * (kScorePitch / 4);  // cleared before any of its pixels is scored
        sc[0] = 0;
        sc[1] = 0;
      }
      {  // groups with a prefilter survivor are the only score words that can hold a corner; their
         // entry keeps the four survivor bits, so pixels are unpacked once per 32 listed words
         // instead of once per 32 visited words
        const uint32_t ma = __ballot_sync(0xffffffffu, flags0 != 0), mb = __ballot_sync(0xffffffffu, flags1 != 0);
        int at = nw + __popc(ma & lt_mask) + __popc(mb & lt_mask);
        if (flags0 != 0) {
          const uint32_t nib = (((flags0 >> 7) & 0x01010101u) * 0x01020408u) >> 24;  // bit b = pixel b survives
          wlist[at++] = (uint16_t)(py << 8 | ent0 | nib);
        }
        if (flags1 != 0) {
          const uint32_t nib = (((flags1 >> 7) & 0x01010101u) * 0x01020408u) >> 24;
          wlist[at] = (uint16_t)((py << 8 | ent0 | nib) + 16);
        }
        nw += __popc(ma) + __popc(mb);
      }
      __syncwarp();
      while (nw - whead >= 32) expand(32);
    }
    if (nw > whead) expand(nw - whead);
    drain(tail - head);
    __syncwarp();

    // keep the listed words that really hold a corner (in place: the write index never passes the
    // read index, and a batch's entries are all read before its ballot)
    {
      int kept_w = 0;
      for (int base = 0; base < nw; base += 32) {
        uint32_t e = 0, w = 0;
        if (base + lane < nw) {
          e = wlist[base + lane];
          w = S.score[((e >> 8) + 1) * (kScorePitch / 4) + ((e >> 4) & 15) + 1];
        }
        const uint32_t m = __ballot_sync(0xffffffffu, w != 0);
        if (w != 0) wlist[kept_w + __popc(m & lt_mask)] = (uint16_t)e;
        kept_w += __popc(m);
      }
      nw = kept_w;
      __syncwarp();
    }
    // strict 3x3 non-max suppression (frame and sub-threshold pixels score 0): one listed score
    // word per lane
    int appended = 0;
    for (int base = 0; base < nw; base += 32) {
      uint32_t w = 0;
      int py = 0, g = 0;
      if (base + lane < nw) {
        const int e = wlist[base + lane];
        py = e >> 8; g = (e >> 4) & 15;
        w = S.score[(py + 1) * (kScorePitch / 4) + g + 1];
      }
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int s = (w >> (8 * b)) & 0xff;
        bool is_max = false;
        const int px = 4 * g + b;
        if (s) {
          const uint8_t* q = &score8[(py + 1) * kScorePitch + px + 4];
          const int n0 = max(max((int)q[-kScorePitch - 1], (int)q[-kScorePitch]), max((int)q[-kScorePitch + 1], (int)q[-1]));
          const int n1 = max(max((int)q[1], (int)q[kScorePitch - 1]), max((int)q[kScorePitch], (int)q[kScorePitch + 1]));
          is_max = s > max(n0, n1);
        }
        const uint32_t m = __ballot_sync(0xffffffffu, is_max);
        if (is_max) kplist[appended + __popc(m & lt_mask)] = (uint16_t)(py << 6 | px);
        appended += __popc(m);
      }
    }
    if (appended == 0) return 0;
    __syncwarp();
    // dense tail: one atomic per cell
    int slot = 0;
    if (lane == 0) slot = atomicAdd(count, appended);
    slot = __shfl_sync(0xffffffffu, slot, 0);
    for (int i = lane; i < appended; i += 32) {
      const uint32_t e = kplist[i];
      const int px = e & 63, py = e >> 6;
      const uint32_t score = score8[(py + 1) * kScorePitch + px + 4];
      const int x = vx0 + px - (kEdge - 3), y = vy0 + py - (kEdge - 3);  // border-relative (minBorder = 16)
      const uint32_t pay = (score << 24) | ((uint32_t)y << 12) | (uint32_t)x;
      if (slot + i < G.cand_cap) out[slot + i] = pay;  // the quadtree kernel prepends the path code
    }
    return appended;
  };

  if (detect(P.ini_th) == 0 && P.min_th < P.ini_th) detect(P.min_th);  // ORBextractor.cc:820: only for empty cells
}

template <int PITCH>
static int launch_fast_pitch(const ExtractParams& P, cudaStream_t s, int max_h) {
  const int qk_entries = std::max(kQueue, (((PITCH - 8) / 2 + 1) * (max_h / 2 + 1) + 1) & ~1);  // as in the kernel
  const int warp_bytes = (int)align_up<size_t>((max_h + 6) * PITCH + (max_h + 2) * PITCH + 2 * qk_entries + 2 * ((max_h * ((PITCH - 8) / 4) + 1) & ~1), 16);  // tiles stay 16-byte aligned
  const size_t smem = (size_t)warp_bytes * kFastWarps;  // 59 x 56 cells (a level that is one cell): 49 392 bytes, just above the 48 KB default
  if (smem > 48 * 1024) ORB_CUDA_TRY(cudaFuncSetAttribute(fast_cells_kernel<PITCH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(div_up(P.total_cells, kFastWarps), P.n_frames);
  fast_cells_kernel<PITCH><<<grid, kFastWarps * 32, smem, s>>>(P, max_h, warp_bytes);
  ORB_LAUNCH_CHECK();
  return ORB_OK;
}

int launch_fast(const ExtractParams& P, cudaStream_t s, int* launches) {
  NvtxRange nvtx_range("orbx:fast");
  int max_w = 1, max_h = 1;
  for (int l = 0; l < P.nlevels; ++l) { max_w = std::max(max_w, P.lv[l].wCell); max_h = std::max(max_h, P.lv[l].hCell); }
  if (max_w >= kCellMax || max_h >= kCellMax) return set_error(ORB_E_INVALID, "FAST cell %dx%d exceeds %d", max_w, max_h, kCellMax - 1);
  ORB_CUDA_TRY(cudaMemsetAsync(P.cand_count, 0, sizeof(int) * (size_t)P.n_frames * P.nlevels, s));
  const int need = max_w + 8;
  int rc;
  if (need <= 40) rc = launch_fast_pitch<40>(P, s, max_h);
  else if (need <= 48) rc = launch_fast_pitch<48>(P, s, max_h);
  else if (need <= 56) rc = launch_fast_pitch<56>(P, s, max_h);
  else rc = launch_fast_pitch<72>(P, s, max_h);
  if (rc) return rc;
  ++*launches;
  return ORB_OK;
}

}  // namespace orbfront
