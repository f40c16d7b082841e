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
// parallelism: 32 or 16 rows for batches that fill the machine anyway, 4 rows when one or a few frames
// would otherwise leave a handful of warps walking long dependent chains.
constexpr int kStripLong = 32, kStripBatch = 16, kStripFew = 4;

// Lane mapping (round 2): a warp owns four sub-tiles of 32 destination columns x kResizeStrip rows (8 lanes each),
// consecutive in the level's row-major sub-tile order -- not one 128-column strip: level widths are rarely
// multiples of 128, and the idle lanes of a row's last strip were 16 % of the executed thread slots.
__host__ __device__ inline int resize_warp_tiles(int w, int h, int strip) { return div_up(div_up(w, 32) * div_up(h, strip), 4); }

template <int kResizeStrip>
__device__ __forceinline__ bool resize_origin(const LevelGeom& D, int warp_tile, int* x0, int* y0) {
  const int lane = threadIdx.x & 31;
  const int t = warp_tile * 4 + (lane >> 3), tiles_x = div_up(D.w, 32);
  const int ty = t / tiles_x, tx = t - ty * tiles_x;
  *x0 = tx * 32 + 4 * (lane & 7);
  *y0 = ty * kResizeStrip;
  return *x0 < D.w && *y0 < D.h;
}

template <int kResizeStrip>
__device__ __forceinline__ void resize_block(const ExtractParams& P, int l, int warp_tile, int f) {
  const LevelGeom& D = P.lv[l];
  int x0, y0;
  if (!resize_origin<kResizeStrip>(D, warp_tile, &x0, &y0)) return;
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
// lane constants, and interpolates with DP2A on packed (c0, c1) 16-bit weights.  Arithmetic identical to
// resize_block.
template <int kResizeStrip>
__device__ __forceinline__ void resize_words_block(const ExtractParams& P, int l, int warp_tile, int f) {
  const LevelGeom& D = P.lv[l];
  int x0, y0;
  if (!resize_origin<kResizeStrip>(D, warp_tile, &x0, &y0)) return;
  const uint8_t* src = level_ptr(P, l - 1, f);
  const int sp = level_pitch(P, l - 1);
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
  const uint8_t* wsrc = src + (first & ~3);
  asm("" : "+l"(wsrc));  // one 64-bit lane pointer: otherwise the uniform base is re-added to every row address (2 more adds per source row)
  // a window word is loaded only if it starts inside the source row (the taps never select bytes
  // past the row's last pixel), so no load leaves the row whatever its pitch
  const int src_w = P.lv[l - 1].w;
  const bool need1 = (first & ~3) + 4 < src_w, need2 = (first & ~3) + 8 < src_w;
  // The loop runs over SOURCE rows: consecutive destination rows share a source row five times out of six at
  // scale 1.2 (row i's lower row is row i+1's upper row), so every source row is loaded and interpolated
  // horizontally once, kept for one iteration, and a destination row is emitted whenever its upper row is the
  // previous source row (the table guarantees strictly increasing upper rows and lower = upper + 1, clamped at
  // the last row -- checked on the host, extractor.cu).  Four source rows of loads are in flight.
  struct Win { uint32_t w0, w1, w2; };
  const int rows = min(kResizeStrip, D.h - y0);
  const ptrdiff_t dpitch = D.pitch;
  const int4* yt = P.rtab + D.ytab + y0;
  const int s_end = __ldg(yt + rows - 1).x + 1;  // last source row consumed (clamped when loaded)
  const int row_max = min(s_end, P.lv[l - 1].h - 1);
  int4 t_cur = __ldg(yt), t_nxt = __ldg(yt + 1);  // the table keeps one entry past the last level's rows (extractor.cu)
  yt += 2;
  asm("" : "+l"(yt));  // a lane pointer that walks the table: otherwise its address is re-derived from the parameters at every row
  int nw = 1 + (int)need1 + (int)need2;  // window words that start inside the source row
  asm("" : "+r"(nw));                    // compared with immediates in the loop (the two predicates were spilled and rebuilt per row)
  auto request = [&](int row, Win& w) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(wsrc + (size_t)((uint32_t)min(row, row_max) * (unsigned long long)(uint32_t)sp));
    w.w0 = __ldg(p); w.w1 = nw > 1 ? __ldg(p + 1) : 0u; w.w2 = nw > 2 ? __ldg(p + 2) : 0u;
  };
  int srow = t_cur.x;  // source row consumed next
  Win wa, wb, wc, wd;
  request(srow, wa);
  request(srow + 1, wb);
  request(srow + 2, wc);
  request(srow + 3, wd);
  uint32_t h0[4] = {0, 0, 0, 0}, h1[4] = {0, 0, 0, 0};
  // one source row: interpolate it horizontally (already >> 4) into hc, re-fill its window slot with row + 3, and emit
  // the destination row whose upper row is the previous source row (hp)
  auto step = [&](Win& w, const uint32_t (&hp)[4], uint32_t (&hc)[4]) {
    const uint32_t v0 = __funnelshift_r(w.w0, w.w1, sh), v1 = __funnelshift_r(w.w1, w.w2, sh);
    const uint32_t p01 = __byte_perm(v0, v1, sel01), p23 = __byte_perm(v0, v1, sel23);
    request(srow + 4, w);
    hc[0] = __dp2a_lo(coef[0], p01, 0u) >> 4;
    hc[1] = __dp2a_hi(coef[1], p01, 0u) >> 4;
    hc[2] = __dp2a_lo(coef[2], p23, 0u) >> 4;
    hc[3] = __dp2a_hi(coef[3], p23, 0u) >> 4;
    if (t_cur.x == srow - 1) {  // never true for the strip's first source row
      // (b * a) >> 16 is the high word of (b << 16) * a: one IMAD.HI with the running sum as addend
      const uint32_t B0 = (uint32_t)t_cur.y << 16, B1 = (uint32_t)t_cur.z << 16;
      uint32_t u[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) u[k] = (__umulhi(B1, hc[k]) + (__umulhi(B0, hp[k]) + 2u)) << 6;  // result byte = bits 8..15
      *reinterpret_cast<uint32_t*>(dst) = __byte_perm(__byte_perm(u[0], u[1], 0x0051), __byte_perm(u[2], u[3], 0x0051), 0x5410);
      dst += dpitch;  // pitch padding absorbs the tail
      t_cur = t_nxt;
      t_nxt = __ldg(yt);
      ++yt;
    }
    return ++srow > s_end;
  };
  for (;;) {  // 4 = window ring (4) and interpolated-row ping-pong (2): no register moves between iterations
    if (step(wa, h0, h1)) break;
    if (step(wb, h1, h0)) break;
    if (step(wc, h0, h1)) break;
    if (step(wd, h1, h0)) break;
  }
}

template <int kStrip>
__global__ void __launch_bounds__(128) resize_level_kernel(const __grid_constant__ ExtractParams P, int l) {
  resize_block<kStrip>(P, l, blockIdx.x * 4 + (threadIdx.x >> 5), blockIdx.z);
}
template <int kStrip>
__global__ void __launch_bounds__(128) resize_level_words_kernel(const __grid_constant__ ExtractParams P, int l) {
  resize_words_block<kStrip>(P, l, blockIdx.x * 4 + (threadIdx.x >> 5), blockIdx.z);
}

// Whole pyramid in ONE cooperative launch, for the tracking thread's case of one or two small frames
// whose every level fits in a single wave: the seven per-level launches are each a few microseconds
// of work behind a kernel boundary, and level l needs all of level l-1, so the levels meet at a grid
// barrier instead.
__global__ void __launch_bounds__(128) pyramid_fused_kernel(const __grid_constant__ ExtractParams P, int* bar) {
  for (int l = 1; l < P.nlevels; ++l) {
    const LevelGeom& D = P.lv[l];
    const int nb = div_up(resize_warp_tiles(D.w, D.h, kStripFew), 4), total = nb * P.n_frames;
    for (int b = blockIdx.x; b < total; b += gridDim.x) {
      const int f = b / nb, r = b - f * nb;
      if (D.resize_words) resize_words_block<kStripFew>(P, l, r * 4 + (threadIdx.x >> 5), f);
      else resize_block<kStripFew>(P, l, r * 4 + (threadIdx.x >> 5), f);
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
    dim3 grid(div_up(resize_warp_tiles(D.w, D.h, kStrip), 4), 1, P.n_frames);
    if (D.resize_words) resize_level_words_kernel<kStrip><<<grid, 128, 0, s>>>(P, l);
    else resize_level_kernel<kStrip><<<grid, 128, 0, s>>>(P, l);
    ORB_LAUNCH_CHECK();
    ++*launches;
  }
  return ORB_OK;
}

int launch_pyramid(const ExtractParams& P, cudaStream_t s, int* launches) {
  NvtxRange nvtx_range("orbx:pyramid");
  if (P.nlevels < 2) return ORB_OK;
  const LevelGeom& D = P.lv[1];
  // warps the largest level offers with long strips; below about one warp per scheduler slot the
  // short strips win (measured: 640x480 x1 76 -> 49 us, 3840x2160 x1 121 -> 88 us)
  // (32-row strips where even they fill the machine: the per-strip set-up -- tap table, first window loads -- is paid half as
  // often: 1024 low-texture frames 0.804 -> 0.762 ms, textured 0.621 -> 0.614 ms)
  if ((long)resize_warp_tiles(D.w, D.h, kStripLong) * P.n_frames >= 148 * 32) return launch_levels<kStripLong>(P, s, launches);
  const long warps_batch = (long)resize_warp_tiles(D.w, D.h, kStripBatch) * P.n_frames;
  if (warps_batch >= 148 * 32) return launch_levels<kStripBatch>(P, s, launches);
  const int blocks_few = div_up(resize_warp_tiles(D.w, D.h, kStripFew), 4) * P.n_frames;
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

// Blur: each lane owns 4 adjacent columns (one 32-bit word) of a kBlurTH-row tile and marches down the rows.  Horizontal 7 taps = two DP4A per output on byte windows
// funnel-shifted out of three aligned words; the last 7 horizontal rows live in registers as u16 pairs
// (row-1, row), so the vertical taps are three DP2A and one IMAD; PRMT packs byte 2 of the four sums.
// Round 2: the row loop has no branch.  The source row pointer walks the BORDER_REFLECT_101 sequence by
// itself (its step flips sign above row 0 and from row h-1 on), stores past the tile's last row are
// predicated off, and the body is 14 rows = the common period of the 7-row ring and the 2-deep load
// pipeline (about 10 KB of SASS).  The round-1 loop spent 43 % of its instructions on the three
// warp-uniform branches per row and their address arithmetic.  No shared memory.
// Round 2, lane mapping: a warp owns FOUR sub-tiles of 32 columns (8 lanes each), consecutive in the level's
// row-major sub-tile order, instead of one 128-column tile: level widths are rarely multiples of 128 (533, 444,
// 370, 308, 257 ... for a 640-wide frame), and whole idle lanes in the last tile of a row were 12 % of the
// executed thread slots; with 32-column granularity 3 % are.  Nothing in the row loop crosses lanes.
constexpr int kBlurSubW = 32, kBlurWarps = 4, kBlurBody = 14;
constexpr int kBlurTH = 112, kBlurTHFew = 14;  // rows per tile: batches / one or a few frames (more, shorter warps)

__host__ __device__ inline int blur_warp_tiles(int w, int h, int th) { return div_up(div_up(w, kBlurSubW) * div_up(h, th), 4); }

__global__ void __launch_bounds__(kBlurWarps * 32, 8)
blur_kernel(const __grid_constant__ ExtractParams P, const int* __restrict__ tile_base /* nlevels+1 */, int th, int skip_empty_levels) {
  const int lane = threadIdx.x & 31;
  const int tile = blockIdx.x * kBlurWarps + (threadIdx.x >> 5);
  if (tile >= tile_base[P.nlevels]) return;
  int l = 0;
  while (l + 1 < P.nlevels && tile >= tile_base[l + 1]) ++l;
  const LevelGeom& G = P.lv[l];
  const int f = blockIdx.z;
  // the reference skips levels without keypoints (:1089); only possible when the quadtree has already run
  if (skip_empty_levels && P.kept_count[f * P.nlevels + l] == 0) return;
  const int t = (tile - tile_base[l]) * 4 + (lane >> 3);  // this lane's sub-tile
  const int tiles_x = div_up(G.w, kBlurSubW);
  const int ty = t / tiles_x, tx = t - ty * tiles_x;
  const int x0 = tx * kBlurSubW + 4 * (lane & 7), y0 = ty * th;
  const int e = G.w - x0;  // pixels of this lane's window that exist from x0 on (>= 8: no right edge)
  if (e <= 0 || y0 >= G.h) return;
  const int sp = level_pitch(P, l);
  // Row edges without a divergent path: every lane runs the word-wise loop on its 12-byte window
  // [x0-4, x0+8); where the window leaves the row, BORDER_REFLECT_101 is applied by three PRMTs whose
  // selectors are lane constants (identity for interior lanes).
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
  const bool ld_a = !left, ld_c = e > 4;  // never read a word that starts outside the row

  const int rows_out = min(th, G.h - y0);
  // source row of input step yy is reflect101(y0 - 3 + yy): the pointer starts there and steps by
  // +-sp, downwards (in memory: backwards) while the un-reflected row is above row 0 or at / below row h-1
  int yin = y0 - 3;
  const uint8_t* rowp = level_ptr(P, l, f) + (size_t)reflect101(min(yin, G.h + 2), G.h) * sp + x0;
  const unsigned turn = (unsigned)(G.h - 1);
  uint32_t pa[2], pb[2], pc[2];
  auto fetch = [&](uint32_t& a, uint32_t& b, uint32_t& c) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(rowp);
    a = ld_a ? __ldg(w - 1) : 0u;
    b = __ldg(w);
    c = ld_c ? __ldg(w + 1) : 0u;
    rowp += (unsigned)yin >= turn ? -(ptrdiff_t)sp : (ptrdiff_t)sp;
    ++yin;
  };
  fetch(pa[0], pb[0], pc[0]);
  fetch(pa[1], pb[1], pc[1]);
  uint32_t ring[7][4] = {}, hn[4];
  uint8_t* outp = P.blur + (size_t)f * P.blur_frame + G.off + (size_t)y0 * G.pitch + x0;  // next output row
  int out_left = rows_out;
  // one input row: horizontal taps into the ring; kOut: the ring then holds the 7 rows of the next output row
  auto step = [&](const int r, const bool kOut) {  // r: input step modulo 14
    const uint32_t a0 = pa[r & 1], b0 = pb[r & 1], c0 = pc[r & 1];
    fetch(pa[r & 1], pb[r & 1], pc[r & 1]);  // row yy + 2 (a few rows past the tile are fetched and never used)
    const uint32_t a = left ? __byte_perm(b0, c0, selA) : a0;
    const uint32_t b = __byte_perm(a0, b0, selB);
    const uint32_t c = __byte_perm(c_from_ab ? a0 : b0, c_from_ab ? b0 : c0, selC);
    hn[0] = __dp4a(__funnelshift_r(a, b, 8), K1, __dp4a(__funnelshift_r(b, c, 8), K2, 0u));
    hn[1] = __dp4a(__funnelshift_r(a, b, 16), K1, __dp4a(__funnelshift_r(b, c, 16), K2, 0u));
    hn[2] = __dp4a(__funnelshift_r(a, b, 24), K1, __dp4a(__funnelshift_r(b, c, 24), K2, 0u));
    hn[3] = __dp4a(b, K1, __dp4a(c, K2, 0u));
#pragma unroll
    for (int k = 0; k < 4; ++k) ring[r % 7][k] = __byte_perm(ring[(r + 6) % 7][k], hn[k], 0x5432);  // (row yy-1, row yy); horizontal sums are <= 255 * 256
    if (kOut) {
      uint32_t acc[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)  // rows yy-6..yy: pairs written at yy-5, yy-3, yy-1 and the row just filtered
        acc[k] = __dp2a_lo(ring[(r + 2) % 7][k], K1, __dp2a_hi(ring[(r + 4) % 7][k], K1,
                 __dp2a_lo(ring[(r + 6) % 7][k], K2, 18u * hn[k] + 32768u)));
      // result bytes are bits 16..23 of each sum; pitch padding absorbs the bytes past the row's end
      const uint32_t v = __byte_perm(__byte_perm(acc[0], acc[1], 0x0062), __byte_perm(acc[2], acc[3], 0x0062), 0x5410);
      if (out_left > 0) *reinterpret_cast<uint32_t*>(outp) = v;
      outp += G.pitch;
      --out_left;
    }
  };
#pragma unroll
  for (int r = 0; r < 6; ++r) step(r, false);
#pragma unroll 1
  while (out_left > 0) {
#pragma unroll
    for (int j = 0; j < kBlurBody; ++j) step(6 + j, true);  // (6 + j) % 14 keeps ring slot and pipeline slot static
  }
}

int launch_blur(const ExtractParams& P, cudaStream_t s, int* launches, bool after_quadtree) {
  NvtxRange nvtx_range("orbx:blur");
  // blur_tiles (extractor.cu) is built for kBlurTH-row tiles; a few frames use kBlurTHFew-row tiles, whose
  // table is the second half of the same array
  const bool few = (long)P.n_frames * blur_warp_tiles(P.lv[0].w, P.lv[0].h, kBlurTH) < 2 * 148;
  const int th = few ? kBlurTHFew : kBlurTH;
  int total = 0;
  for (int l = 0; l < P.nlevels; ++l) total += blur_warp_tiles(P.lv[l].w, P.lv[l].h, th);
  blur_kernel<<<dim3(div_up(total, kBlurWarps), 1, P.n_frames), kBlurWarps * 32, 0, s>>>(P, P.blur_tiles + (few ? P.nlevels + 1 : 0), th,
                                                                                       after_quadtree ? 1 : 0);
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

// Every level of one frame with its 19-px BORDER_REFLECT_101 frame, back to back in `dst` (level l at
// off[l], row pitch w_l + 38): the mvImagePyramid a Frame reads (Frame.cc:534,624-641) in ONE launch, so that
// the single-frame drop-in path needs one device-to-host copy instead of one kernel + copy + sync per level.
struct BorderOffsets { unsigned long long off[kMaxLevels]; };
__global__ void border_all_kernel(const __grid_constant__ ExtractParams P, int f, const __grid_constant__ BorderOffsets A, uint8_t* __restrict__ dst) {
  const int l = blockIdx.y;
  const LevelGeom& G = P.lv[l];
  const int W = G.w + 2 * kEdge, H = G.h + 2 * kEdge;
  const uint8_t* src = level_ptr(P, l, f);
  const int sp = level_pitch(P, l);
  uint8_t* out = dst + A.off[l];
  for (int y = blockIdx.x; y < H; y += gridDim.x) {
    int sy = y - kEdge;
    sy = sy < 0 ? -sy : (sy >= G.h ? 2 * G.h - 2 - sy : sy);
    const uint8_t* row = src + (size_t)sy * sp;
    for (int x = threadIdx.x; x < W; x += blockDim.x) {
      int sx = x - kEdge;
      sx = sx < 0 ? -sx : (sx >= G.w ? 2 * G.w - 2 - sx : sx);
      out[(size_t)y * W + x] = row[sx];
    }
  }
}

int launch_border_all(const ExtractParams& P, int f, const size_t* off, uint8_t* dst, cudaStream_t s) {
  BorderOffsets A;
  for (int l = 0; l < kMaxLevels; ++l) A.off[l] = l < P.nlevels ? off[l] : 0;
  border_all_kernel<<<dim3(std::min(P.lv[0].h + 2 * kEdge, 148), P.nlevels), 256, 0, s>>>(P, f, A, dst);
  ORB_LAUNCH_CHECK();
  return ORB_OK;
}

int launch_border_copy(const uint8_t* src, int w, int h, size_t pitch, uint8_t* dst, cudaStream_t s) {
  dim3 grid(div_up(w + 2 * kEdge, 256), h + 2 * kEdge);
  border_copy_kernel<<<grid, 256, 0, s>>>(src, w, h, pitch, dst);
  ORB_LAUNCH_CHECK();
  return ORB_OK;
}

}  // namespace orbfront
