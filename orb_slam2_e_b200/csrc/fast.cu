// FAST-9/16 detection per 30-px cell with the iniTh -> minTh fallback (sm_100a).
//
// Replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree (src/ORBextractor.cc:789-837)
// and the cv::FAST(..., nonmaxSuppression=true) calls inside it.  Semantics reproduced
// (pinned against cv2 4.13 in tests/test_oracle_prims.py):
//   * corner strength S = max over the 16 nine-long arcs of min(v-ring) and of min(ring-v), minus 1;
//     a pixel is a corner at threshold T iff S >= T; response = S (threshold independent);
//   * 3x3 strict non-max suppression inside the cell's valid region only (the cell ROI's outer
//     3-px ring has no score, so NMS never acts across cell seams);
//   * a cell with no keypoint at iniThFAST is redone at minThFAST.
// One warp per cell: stage the cell (+3 px halo) in shared memory, cheap opposite-pair test on
// every pixel, ballot-compacted survivors get the exact strength through a packed u16x2
// VIMNMX3 network (dark arcs in the low half, bright arcs in the high half), then NMS and an
// order-free append to the (frame, level) candidate list together with the quadtree path code.
#include "extract.cuh"

namespace orbfront {

constexpr int kFastWarps = 4;
constexpr int kCellMax = 60;                 // wCell,hCell <= 59 (ceil(width/floor(width/30)))
constexpr int kTilePitch = 68;               // (59+6) rounded up to a multiple of 4
constexpr int kTileRows = kCellMax + 6;
constexpr int kScorePitch = 64;

struct FastSmem {
  uint8_t tile[kTileRows * kTilePitch];
  uint8_t score[kCellMax * kScorePitch];
  uint32_t flag_ini[(kCellMax * kCellMax + 31) / 32];
  uint32_t flag_min[(kCellMax * kCellMax + 31) / 32];
  uint16_t queue[64];
};

// exact strength of the pixel at tile position p (pointer to the centre byte)
__device__ __forceinline__ int fast_strength(const uint8_t* p) {
  // packed lanes: low half = 256 + (v - ring)  (dark arc),  high half = 256 + (ring - v) (bright arc)
  const int v = p[0];
  const uint32_t base = 0x01000100u + (uint32_t)v * 0xFFFF0001u;  // (256+v) | (256-v)<<16, no borrow: both halves stay in [1,511]
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
  const int s = (int)max(best & 0xffffu, best >> 16) - 256 - 1;
  return s;
}

__global__ void __launch_bounds__(kFastWarps * 32)
fast_cells_kernel(const __grid_constant__ ExtractParams P) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  FastSmem& S = reinterpret_cast<FastSmem*>(smem_raw)[threadIdx.x >> 5];
  const int lane = threadIdx.x & 31;
  const int cell = blockIdx.x * kFastWarps + (threadIdx.x >> 5);
  if (cell >= P.total_cells) return;
  const int f = blockIdx.y;
  int l = 0;
  while (l + 1 < P.nlevels && cell >= P.lv[l + 1].cell_base) ++l;
  const LevelGeom& G = P.lv[l];
  const int ci = (cell - G.cell_base) / G.nCols, cj = (cell - G.cell_base) % G.nCols;
  // valid (scored) region of the cell in level coordinates
  const int vx0 = kEdge + cj * G.wCell, vy0 = kEdge + ci * G.hCell;
  const int vw = min(G.wCell, G.w - kEdge - vx0), vh = min(G.hCell, G.h - kEdge - vy0);
  if (vw <= 0 || vh <= 0) return;

  // stage (vw+6) x (vh+6) pixels
  const uint8_t* src = level_ptr(P, l, f) + (size_t)(vy0 - 3) * level_pitch(P, l) + (vx0 - 3);
  const int sp = level_pitch(P, l);
  const int tw = vw + 6, th = vh + 6;
  for (int r = 0; r < th; ++r) {
    const uint8_t* s = src + (size_t)r * sp;
    for (int c = lane; c < tw; c += 32) S.tile[r * kTilePitch + c] = s[c];
  }
  __syncwarp();

  // pass 1: opposite-pair rejection on all pixels, exact strength on the survivors
  const int th_min = P.min_th, th_ini = P.ini_th;
  const int npix = vw * vh;
  int qn = 0;  // queue fill (uniform across the warp)
  auto drain = [&](int count) {
    if (lane < count) {
      const int id = S.queue[lane];
      const int py = id / vw, px = id - py * vw;
      const int s = fast_strength(&S.tile[(py + 3) * kTilePitch + px + 3]);
      S.score[py * kScorePitch + px] = (uint8_t)(s >= th_min ? s : 0);
    }
  };
  for (int base = 0; base < npix; base += 32) {
    const int id = base + lane;
    bool pass = false;
    if (id < npix) {
      const int py = id / vw, px = id - py * vw;
      const uint8_t* p = &S.tile[(py + 3) * kTilePitch + px + 3];
      const int v = p[0];
      // every nine-long arc holds one pixel of each opposite pair
      bool dark = true, bright = true;
#define PAIR(dx, dy)                                                         \
  {                                                                          \
    const int a = v - (int)p[(dy) * kTilePitch + (dx)];                      \
    const int b = v - (int)p[-(dy) * kTilePitch - (dx)];                     \
    dark = dark && (a > th_min || b > th_min);                               \
    bright = bright && (a < -th_min || b < -th_min);                         \
  }
      PAIR(0, 3) PAIR(3, 0) PAIR(2, 2) PAIR(2, -2)
#undef PAIR
      pass = dark || bright;
      if (!pass) S.score[py * kScorePitch + px] = 0;
    }
    const uint32_t m = __ballot_sync(0xffffffffu, pass);
    if (pass) S.queue[qn + __popc(m & ((1u << lane) - 1))] = (uint16_t)id;
    qn += __popc(m);
    __syncwarp();
    if (qn >= 32) {
      drain(32);
      __syncwarp();
      if (lane < qn - 32) S.queue[lane] = S.queue[32 + lane];
      qn -= 32;
      __syncwarp();
    }
  }
  drain(qn);
  __syncwarp();

  // pass 2: strict 3x3 non-max suppression inside the valid region; count both thresholds
  int n_ini = 0, n_min = 0;
  const int nchunks = (npix + 31) / 32;
  for (int c = 0; c < nchunks; ++c) {
    const int id = c * 32 + lane;
    bool is_max = false;
    int s = 0;
    if (id < npix) {
      const int py = id / vw, px = id - py * vw;
      s = S.score[py * kScorePitch + px];
      if (s) {
        is_max = true;
#pragma unroll
        for (int dy = -1; dy <= 1; ++dy)
#pragma unroll
          for (int dx = -1; dx <= 1; ++dx) {
            if (dx == 0 && dy == 0) continue;
            const int qx = px + dx, qy = py + dy;
            const int o = (qx >= 0 && qx < vw && qy >= 0 && qy < vh) ? S.score[qy * kScorePitch + qx] : 0;
            is_max = is_max && s > o;
          }
      }
    }
    const uint32_t mm = __ballot_sync(0xffffffffu, is_max);
    const uint32_t mi = __ballot_sync(0xffffffffu, is_max && s >= th_ini);
    if (lane == 0) { S.flag_min[c] = mm; S.flag_ini[c] = mi; }
    n_min += __popc(mm);
    n_ini += __popc(mi);
  }
  __syncwarp();
  const bool use_ini = n_ini > 0;  // ORBextractor.cc:820: fall back only when the cell is empty
  const int total = use_ini ? n_ini : n_min;
  if (total == 0) return;
  int slot = 0;
  if (lane == 0) slot = atomicAdd(&P.cand_count[f * P.nlevels + l], total);
  slot = __shfl_sync(0xffffffffu, slot, 0);
  unsigned long long* out = P.cand + (size_t)f * P.cand_frame + G.cand_off;
  for (int c = 0; c < nchunks; ++c) {
    const uint32_t m = use_ini ? S.flag_ini[c] : S.flag_min[c];
    if (m >> lane & 1u) {
      const int id = c * 32 + lane;
      const int py = id / vw, px = id - py * vw;
      const int x = vx0 + px - (kEdge - 3), y = vy0 + py - (kEdge - 3);  // border-relative (minBorder = 16)
      const uint32_t s = S.score[py * kScorePitch + px];
      const uint32_t pay = (s << 24) | ((uint32_t)y << 12) | (uint32_t)x;
      const int at = slot + __popc(m & ((1u << lane) - 1));
      if (at < G.cand_cap) out[at] = ((unsigned long long)quadtree_path_code(x, y, G) << 32) | pay;
    }
    slot += __popc(m);
  }
}

int launch_fast(const ExtractParams& P, cudaStream_t s, int* launches) {
  const size_t smem = sizeof(FastSmem) * kFastWarps;  // < 48 KB: no opt-in attribute needed
  static_assert(sizeof(FastSmem) * kFastWarps <= 48 * 1024, "fast_cells_kernel shared memory");
  ORB_CUDA_TRY(cudaMemsetAsync(P.cand_count, 0, sizeof(int) * (size_t)P.n_frames * P.nlevels, s));
  dim3 grid(div_up(P.total_cells, kFastWarps), P.n_frames);
  fast_cells_kernel<<<grid, kFastWarps * 32, smem, s>>>(P);
  ORB_LAUNCH_CHECK();
  ++*launches;
  return ORB_OK;
}

}  // namespace orbfront
