// Image pyramid and Gaussian blur kernels (sm_100a).
//
//   resize_level_kernel : level l = bilinear (11-bit fixed point) resize of level l-1
//                         -- ORBextractor::ComputePyramid, src/ORBextractor.cc:1115-1140, with
//                         cv::resize(INTER_LINEAR, CV_8U) arithmetic.
//   blur_kernel         : 7x7 sigma=2 Gaussian, fixed-point kernel {18,34,48,56,48,34,18}/256 per
//                         axis, BORDER_REFLECT_101 -- src/ORBextractor.cc:1093-1094.
//   border_copy_kernel  : builds the 19-px reflect-101 frame on demand for orbx_get_level.
// All three are HBM-streaming byte kernels (bound: HBM bandwidth); no tensor-core work.
#include <algorithm>
#include <cstdlib>

#include "extract.cuh"

namespace orbfront {

// One warp marches down a strip of destination rows; each lane owns 4 adjacent destination
// columns, keeps their horizontal taps in registers and reuses the horizontally interpolated
// source row shared by consecutive destination rows.
// The strip length trades redundant work (each strip interpolates its first source rows again) for
// parallelism: 16 rows for batches that fill the machine anyway, 4 rows when one or a few frames
// would otherwise leave a handful of warps walking long dependent chains.
constexpr int kStripBatch = 16, kStripFew = 4;

template <int kResizeStrip>
__device__ __forceinline__ void resize_block(const ExtractParams& P, int l, int bx, int by, int f) {
  const LevelGeom& D = P.lv[l];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x0 = (bx * 32 + lane) * 4;
  const int y0 = (by * 4 + warp) * kResizeStrip;
  if (x0 >= D.w || y0 >= D.h) return;
  const uint8_t* src = level_ptr(P, l - 1, f);
  const int sp = level_pitch(P, l - 1);
  uint8_t* dst = P.pyr + (size_t)f * P.pyr_frame + D.off;

  int sx0[4], sx1[4], c0[4], c1[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int4 t = __ldg(&P.rtab[D.xtab + min(x0 + k, D.w - 1)]);
    sx0[k] = t.x; c0[k] = t.y; c1[k] = t.z; sx1[k] = t.w;
  }
  int rowA = -1, rowB = -1;
  int hA[4], hB[4];
  auto hrow = [&](int sy, int (&h)[4]) {
    const uint8_t* s = src + (size_t)sy * sp;
#pragma unroll
    for (int k = 0; k < 4; ++k) h[k] = (int)s[sx0[k]] * c0[k] + (int)s[sx1[k]] * c1[k];
  };
  const int y1 = min(y0 + kResizeStrip, D.h);
  for (int y = y0; y < y1; ++y) {
    const int4 t = __ldg(&P.rtab[D.ytab + y]);
    const int sy0 = t.x, sy1 = t.w, b0 = t.y, b1 = t.z;
    if (sy0 == rowB) {
      rowA = rowB;
#pragma unroll
      for (int k = 0; k < 4; ++k) hA[k] = hB[k];
    } else if (sy0 != rowA) {
      hrow(sy0, hA);
      rowA = sy0;
    }
    if (sy1 == rowA) {
#pragma unroll
      for (int k = 0; k < 4; ++k) hB[k] = hA[k];
      rowB = sy1;
    } else if (sy1 != rowB) {
      hrow(sy1, hB);
      rowB = sy1;
    }
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int v = (((b0 * (hA[k] >> 4)) >> 16) + ((b1 * (hB[k] >> 4)) >> 16) + 2) >> 2;
      out |= (uint32_t)(v & 0xff) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)y * D.pitch + x0) = out;  // pitch padding absorbs the tail
  }
}

// Word-wise variant for scale factors up to ~2 (every 4-pixel destination group reads source
// columns spanning at most 8 bytes): per source row a lane loads three aligned words, shifts its
// 8-byte window into place, picks the (left, right) byte pairs with two PRMTs whose selectors are
// lane constants, and interpolates with DP2A on packed (c0, c1) 16-bit weights -- 11 instructions per
// source row instead of 8 byte loads + 8 IMADs + their addressing.  Source rows are consumed in
// order and requested one row ahead.  Arithmetic identical to resize_level_kernel.
template <int kResizeStrip>
__device__ __forceinline__ void resize_words_block(const ExtractParams& P, int l, int bx, int by, int f) {
  const LevelGeom& D = P.lv[l];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x0 = (bx * 32 + lane) * 4;
  const int y0 = (by * 4 + warp) * kResizeStrip;
  if (x0 >= D.w || y0 >= D.h) return;
  const uint8_t* src = level_ptr(P, l - 1, f);
  const int sp = level_pitch(P, l - 1);
  const int src_h = P.lv[l - 1].h;
  uint8_t* dst = P.pyr + (size_t)f * P.pyr_frame + D.off + (size_t)y0 * D.pitch + x0;

  // lane constants
  uint32_t coef[4], sel01 = 0, sel23 = 0;
  int first = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int4 t = __ldg(&P.rtab[D.xtab + min(x0 + k, D.w - 1)]);
    if (k == 0) first = t.x;
    coef[k] = (uint32_t)t.y | (uint32_t)t.z << 16;
    const uint32_t pair = (uint32_t)(t.x - first) | (uint32_t)(t.w - first) << 4;  // byte indices of (left, right) in the window
    if (k < 2) sel01 |= pair << (8 * k); else sel23 |= pair << (8 * (k - 2));
  }
  const int sh = (first & 3) * 8;
  const uint32_t* wsrc = reinterpret_cast<const uint32_t*>(src) + (first >> 2);
  const int spw = sp >> 2;
  // a window word is loaded only if it starts inside the source row (the taps never select bytes
  // past the row's last pixel), so no load leaves the row whatever its pitch
  const int src_w = P.lv[l - 1].w;
  const bool need1 = (first & ~3) + 4 < src_w, need2 = (first & ~3) + 8 < src_w;

  // Horizontally interpolated source rows (already >> 4) live in two register sets that alternate
  // with the parity of the consumed row, like the word sets they are made from: no moves (a move
  // would wait for the load), and the vertical step is compiled once per parity.
  uint32_t h0[4] = {0, 0, 0, 0}, h1[4] = {0, 0, 0, 0};
  // source rows are consumed in order: the word set just consumed is refilled with the row after
  // next, so two rows stay in flight and a destination row that advances by two source rows does
  // not wait for a fresh load.  The request pointer walks down the source and stops at its last row.
  uint32_t wa0 = 0, wa1 = 0, wa2 = 0, wb0 = 0, wb1 = 0, wb2 = 0;
  const int first_row = __ldg(&P.rtab[D.ytab + y0]).x;
  int rowB = first_row - 1;  // last consumed source row
  int req = first_row;       // next source row to request
  const uint32_t* preq = wsrc + (size_t)min(req, src_h - 1) * spw;
  auto request = [&](uint32_t& a, uint32_t& b, uint32_t& c) {
    a = __ldg(preq);
    b = need1 ? __ldg(preq + 1) : 0u;
    c = need2 ? __ldg(preq + 2) : 0u;
    if (++req < src_h) preq += spw;
  };
  auto consume = [&](uint32_t w0, uint32_t w1, uint32_t w2, uint32_t (&h)[4]) {
    const uint32_t v0 = __funnelshift_r(w0, w1, sh), v1 = __funnelshift_r(w1, w2, sh);
    const uint32_t p01 = __byte_perm(v0, v1, sel01), p23 = __byte_perm(v0, v1, sel23);
    h[0] = __dp2a_lo(coef[0], p01, 0u) >> 4;
    h[1] = __dp2a_hi(coef[1], p01, 0u) >> 4;
    h[2] = __dp2a_lo(coef[2], p23, 0u) >> 4;
    h[3] = __dp2a_hi(coef[3], p23, 0u) >> 4;
  };
  const int rows = min(kResizeStrip, D.h - y0), dpitch = D.pitch;
  const int4* ytab = P.rtab + D.ytab + y0;
  bool odd = false;  // parity of the number of consumed rows: the next row goes to (wb, h1) when set
  request(wa0, wa1, wa2);
  request(wb0, wb1, wb2);
  // (b0 * a) >> 16 is the high word of (b0 << 16) * a: one IMAD.HI with the running sum as addend
  auto vertical = [&](const uint32_t (&older)[4], const uint32_t (&newer)[4], uint32_t B0, uint32_t B1, bool same) {
    uint32_t u[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t a = same ? newer[k] : older[k];
      u[k] = (__umulhi(B1, newer[k]) + (__umulhi(B0, a) + 2u)) << 6;  // result byte = bits 8..15
    }
    *reinterpret_cast<uint32_t*>(dst) = __byte_perm(__byte_perm(u[0], u[1], 0x0051), __byte_perm(u[2], u[3], 0x0051), 0x5410);
    dst += dpitch;  // pitch padding absorbs the tail
  };
  for (int i = 0; i < rows; ++i) {
    const int4 t = __ldg(ytab + i);
    const int sy1 = t.w;
#pragma unroll 1
    while (rowB < sy1) {  // one or two trips: sy1 is non-decreasing and advances by at most 2 per destination row
      if (!odd) {
        consume(wa0, wa1, wa2, h0);
        request(wa0, wa1, wa2);
      } else {
        consume(wb0, wb1, wb2, h1);
        request(wb0, wb1, wb2);
      }
      odd = !odd;
      ++rowB;
    }
    const bool same = t.x == sy1;  // clamped rows at the top / bottom border
    const uint32_t B0 = (uint32_t)t.y << 16, B1 = (uint32_t)t.z << 16;
    if (odd) vertical(h1, h0, B0, B1, same);  // the row consumed last went to h0
    else vertical(h0, h1, B0, B1, same);
  }
}

template <int kStrip>
__global__ void __launch_bounds__(128) resize_level_kernel(const __grid_constant__ ExtractParams P, int l) {
  resize_block<kStrip>(P, l, blockIdx.x, blockIdx.y, blockIdx.z);
}
template <int kStrip>
__global__ void __launch_bounds__(128) resize_level_words_kernel(const __grid_constant__ ExtractParams P, int l) {
  resize_words_block<kStrip>(P, l, blockIdx.x, blockIdx.y, blockIdx.z);
}

// Whole pyramid in ONE cooperative launch, for the tracking thread's case of one or two small frames
// whose every level fits in a single wave: the seven per-level launches are each a few microseconds
// of work behind a kernel boundary, and level l needs all of level l-1, so the levels meet at a grid
// barrier instead.
__global__ void __launch_bounds__(128) pyramid_fused_kernel(const __grid_constant__ ExtractParams P, int* bar) {
  for (int l = 1; l < P.nlevels; ++l) {
    const LevelGeom& D = P.lv[l];
    const int nbx = div_up(D.w, 128), nby = div_up(D.h, 4 * kStripFew), total = nbx * nby * P.n_frames;
    for (int b = blockIdx.x; b < total; b += gridDim.x) {
      const int f = b / (nbx * nby), r = b - f * (nbx * nby);
      if (D.resize_words) resize_words_block<kStripFew>(P, l, r % nbx, r / nbx, f);
      else resize_block<kStripFew>(P, l, r % nbx, r / nbx, f);
    }
    if (l + 1 < P.nlevels) {  // level l complete and visible before anyone reads it
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1);
        while (*(volatile int*)bar < l * (int)gridDim.x) __nanosleep(20);
        __threadfence();
      }
      __syncthreads();
    }
  }
}

int pyramid_fused_max_grid(int device) {
  int per_sm = 0, sms = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, pyramid_fused_kernel, 128, 0) != cudaSuccess ||
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess)
    return 0;
  return per_sm * sms;
}

template <int kStrip>
static int launch_levels(const ExtractParams& P, cudaStream_t s, int* launches) {
  for (int l = 1; l < P.nlevels; ++l) {
    const LevelGeom& D = P.lv[l];
    dim3 grid(div_up(D.w, 128), div_up(D.h, 4 * kStrip), P.n_frames);
    if (D.resize_words) resize_level_words_kernel<kStrip><<<grid, 128, 0, s>>>(P, l);
    else resize_level_kernel<kStrip><<<grid, 128, 0, s>>>(P, l);
    ORB_LAUNCH_CHECK();
    ++*launches;
  }
  return ORB_OK;
}

int launch_pyramid(const ExtractParams& P, cudaStream_t s, int* launches) {
  if (P.nlevels < 2) return ORB_OK;
  const LevelGeom& D = P.lv[1];
  // warps the largest level offers with long strips; below about one warp per scheduler slot the
  // short strips win (measured: 640x480 x1 76 -> 49 us, 3840x2160 x1 121 -> 88 us)
  const long warps_batch = (long)div_up(D.w, 128) * div_up(D.h, kStripBatch) * P.n_frames;
  if (warps_batch >= 148 * 32) return launch_levels<kStripBatch>(P, s, launches);
  const int blocks_few = div_up(D.w, 128) * div_up(D.h, 4 * kStripFew) * P.n_frames;
  if (P.nlevels > 2 && P.pyr_bar && blocks_few <= P.pyr_max_grid && !getenv("ORBX_NO_FUSED_PYRAMID")) {
    int* bar = P.pyr_bar;
    ORB_CUDA_TRY(cudaMemsetAsync(bar, 0, sizeof(int), s));
    void* args[] = {(void*)&P, (void*)&bar};
    ORB_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)pyramid_fused_kernel, dim3(blocks_few), dim3(128), args, 0, s));
    count_launch();
    ++*launches;
    return ORB_OK;
  }
  return launch_levels<kStripFew>(P, s, launches);
}

// ------------------------------------------------------------------------------------
__device__ __forceinline__ int reflect101(int p, int n) {
  if (p < 0) p = -p;
  if (p >= n) p = 2 * n - 2 - p;
  return p;  // valid for |overshoot| < n, which holds for 3-px taps on levels >= 7 px
}

// Blur: one warp per 128-column x 36-row tile, each lane owns 4 adjacent columns (one 32-bit
// word) and marches down the rows.  Horizontal 7 taps = two DP4A per output on byte windows
// funnel-shifted out of three aligned words; the last 7 horizontal rows live in registers (ring),
// vertical 7 taps use the kernel's symmetry and PRMT packs byte 2 of the four sums.  The 42 input
// rows run as 2 x 21: 21 is the common period of the 7-row ring and the 3-deep load pipeline, so
// register indices stay static while the loop body (about 20 KB of SASS) stays inside the 32 KB
// instruction cache -- the fully unrolled form stalled on instruction fetch.  No shared memory.
constexpr int kBlurTW = 128, kBlurTH = 36, kBlurWarps = 4;
constexpr int kBlurRows = kBlurTH + 6, kBlurBody = 21;
static_assert(kBlurRows % kBlurBody == 0, "row loop must be whole bodies");

__global__ void __launch_bounds__(kBlurWarps * 32, 8)
blur_kernel(const __grid_constant__ ExtractParams P, const int* __restrict__ tile_base /* nlevels+1 */, int skip_empty_levels) {
  const int lane = threadIdx.x & 31;
  const int tile = blockIdx.x * kBlurWarps + (threadIdx.x >> 5);
  if (tile >= tile_base[P.nlevels]) return;
  int l = 0;
  while (l + 1 < P.nlevels && tile >= tile_base[l + 1]) ++l;
  const LevelGeom& G = P.lv[l];
  const int f = blockIdx.z;
  // the reference skips levels without keypoints (:1089); only possible when the quadtree has already run
  if (skip_empty_levels && P.kept_count[f * P.nlevels + l] == 0) return;
  const int t = tile - tile_base[l];
  const int tiles_x = div_up(G.w, kBlurTW);
  const int tx0 = (t % tiles_x) * kBlurTW;
  const int x0 = tx0 + 4 * lane, y0 = (t / tiles_x) * kBlurTH;
  const uint8_t* src = level_ptr(P, l, f);
  const int sp = level_pitch(P, l);
  uint8_t* dst = P.blur + (size_t)f * P.blur_frame + G.off;
  // Row edges without a divergent path: every lane with a pixel in the row runs the word-wise loop on
  // its 12-byte window [x0-4, x0+8); where the window leaves the row, BORDER_REFLECT_101 is applied
  // by three PRMTs whose selectors are lane constants (identity for interior lanes).
  const int e = G.w - x0;  // pixels of this lane's window that exist from x0 on (>= 8: no right edge)
  const bool active = e > 0;
  // tiles away from the top/bottom border walk a running row pointer; the others reflect per row
  const bool y_inside = y0 >= 3 && y0 + kBlurTH + 3 <= G.h;
  constexpr uint32_t K1 = 18u | 34u << 8 | 48u << 16 | 56u << 24, K2 = 48u | 34u << 8 | 18u << 16;
  // window byte i (0..11) <-> x0-4+i; byte i >= e+4 mirrors to 2(e+3)-i, byte i < 4 of lane x0 = 0 to 8-i.
  // a' = PRMT(b, c, selA); b' = PRMT(a, b, selB); c' = PRMT(lo, hi, selC) with (lo, hi) = (a, b) if e <= 4 else (b, c)
  uint32_t selA = 0x3210u, selB = 0x7654u, selC = 0x7654u;  // identity: a' = a is taken separately
  const bool left = x0 == 0, c_from_ab = e <= 4;
  if (e < 8) {
    selB = 0; selC = 0;
    for (int j = 0; j < 4; ++j) {
      const int ib = 4 + j, ic = 8 + j;
      const int sb = ib >= e + 4 ? 2 * (e + 3) - ib : ib;  // source window byte, inside (a, b)
      const int sc = max(ic >= e + 4 ? 2 * (e + 3) - ic : ic, 0);  // inside (a, b) when e <= 4, else inside (b, c); bytes past e+6 are never used
      selB |= (uint32_t)sb << (4 * j);
      selC |= (uint32_t)(c_from_ab ? sc : sc - 4) << (4 * j);
    }
  }
  if (left) selA = 0x1234u;  // x = -4..-1 <- x = 4, 3, 2, 1 out of (b, c)

  if (active) {
    uint32_t ring[7][4] = {}, hn[4];
    constexpr int kAhead = 3;  // the three words of row yy+3 are requested while row yy is filtered
    uint32_t pa[kAhead], pb[kAhead], pc[kAhead];
    const uint8_t* rowp = src + (size_t)(y_inside ? y0 - 3 : 0) * sp + x0;  // next row to fetch (y_inside)
    auto fetch = [&](int yy, uint32_t& a, uint32_t& b, uint32_t& c) {
      const uint8_t* p = rowp;
      if (y_inside) {
        rowp += sp;
      } else {
        const int y = y0 + yy - 3;
        p = src + (size_t)reflect101(min(y, G.h + 2), G.h) * sp + x0;  // rows past h+2 are never used
      }
      const uint32_t* w = reinterpret_cast<const uint32_t*>(p);
      a = left ? 0u : __ldg(w - 1);
      b = __ldg(w);
      c = e > 4 ? __ldg(w + 1) : 0u;  // never read a word that starts past the row
    };
#pragma unroll
    for (int k = 0; k < kAhead; ++k) fetch(k, pa[k], pb[k], pc[k]);
    uint8_t* outp = dst + (size_t)y0 * G.pitch + x0;  // next output row
    int rows_left = min(kBlurTH, G.h - y0);
#pragma unroll 1
    for (int blk = 0; blk < kBlurRows / kBlurBody; ++blk) {
#pragma unroll
      for (int r = 0; r < kBlurBody; ++r) {
        const int yy = blk * kBlurBody + r;
        {
          const uint32_t a0 = pa[r % kAhead], b0 = pb[r % kAhead], c0 = pc[r % kAhead];
          if (yy + kAhead < kBlurRows) fetch(yy + kAhead, pa[r % kAhead], pb[r % kAhead], pc[r % kAhead]);
          const uint32_t a = left ? __byte_perm(b0, c0, selA) : a0;
          const uint32_t b = __byte_perm(a0, b0, selB);
          const uint32_t c = __byte_perm(c_from_ab ? a0 : b0, c_from_ab ? b0 : c0, selC);
          // the ring keeps (row yy-1, row yy) as u16 pairs, so the vertical taps are three DP2A and one IMAD
          hn[0] = __dp4a(__funnelshift_r(a, b, 8), K1, __dp4a(__funnelshift_r(b, c, 8), K2, 0u));
          hn[1] = __dp4a(__funnelshift_r(a, b, 16), K1, __dp4a(__funnelshift_r(b, c, 16), K2, 0u));
          hn[2] = __dp4a(__funnelshift_r(a, b, 24), K1, __dp4a(__funnelshift_r(b, c, 24), K2, 0u));
          hn[3] = __dp4a(b, K1, __dp4a(c, K2, 0u));
#pragma unroll
          for (int k = 0; k < 4; ++k) ring[r % 7][k] = __byte_perm(ring[(r + 6) % 7][k], hn[k], 0x5432);  // horizontal sums are <= 255 * 256
        }
        if ((r >= 6 || blk > 0) && rows_left > 0) {  // the ring holds the 7 horizontal rows of the next output row
          uint32_t acc[4];
#pragma unroll
          for (int k = 0; k < 4; ++k)  // rows yy-6..yy: pairs written at yy-5, yy-3, yy-1 and the row just filtered
            acc[k] = __dp2a_lo(ring[(r + 2) % 7][k], K1, __dp2a_hi(ring[(r + 4) % 7][k], K1,
                     __dp2a_lo(ring[(r + 6) % 7][k], K2, 18u * hn[k] + 32768u)));
          // result bytes are bits 16..23 of each sum; pitch padding absorbs the bytes past the row's end
          *reinterpret_cast<uint32_t*>(outp) =
              __byte_perm(__byte_perm(acc[0], acc[1], 0x0062), __byte_perm(acc[2], acc[3], 0x0062), 0x5410);
          outp += G.pitch;
          --rows_left;
        }
      }
    }
  }
}

int launch_blur(const ExtractParams& P, cudaStream_t s, int* launches, bool after_quadtree) {
  int total = 0;
  for (int l = 0; l < P.nlevels; ++l) total += div_up(P.lv[l].w, kBlurTW) * div_up(P.lv[l].h, kBlurTH);
  blur_kernel<<<dim3(div_up(total, kBlurWarps), 1, P.n_frames), kBlurWarps * 32, 0, s>>>(P, P.blur_tiles, after_quadtree ? 1 : 0);
  ORB_LAUNCH_CHECK();
  ++*launches;
  return ORB_OK;
}

// ------------------------------------------------------------------------------------
__global__ void border_copy_kernel(const uint8_t* __restrict__ src, int w, int h, size_t pitch, uint8_t* __restrict__ dst) {
  const int W = w + 2 * kEdge, H = h + 2 * kEdge;
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= W || y >= H) return;
  int sx = x - kEdge, sy = y - kEdge;
  while (sx < 0 || sx >= w) sx = sx < 0 ? -sx : 2 * w - 2 - sx;
  while (sy < 0 || sy >= h) sy = sy < 0 ? -sy : 2 * h - 2 - sy;
  dst[(size_t)y * W + x] = src[(size_t)sy * pitch + sx];
}

int launch_border_copy(const uint8_t* src, int w, int h, size_t pitch, uint8_t* dst, cudaStream_t s) {
  dim3 grid(div_up(w + 2 * kEdge, 256), h + 2 * kEdge);
  border_copy_kernel<<<grid, 256, 0, s>>>(src, w, h, pitch, dst);
  ORB_LAUNCH_CHECK();
  return ORB_OK;
}

}  // namespace orbfront
