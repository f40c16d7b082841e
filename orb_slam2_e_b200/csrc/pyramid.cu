// Image pyramid and Gaussian blur kernels (sm_100a).
//
//   resize_level_kernel : level l = bilinear (11-bit fixed point) resize of level l-1
//                         -- ORBextractor::ComputePyramid, src/ORBextractor.cc:1115-1140, with
//                         cv::resize(INTER_LINEAR, CV_8U) arithmetic.
//   blur_kernel         : 7x7 sigma=2 Gaussian, fixed-point kernel {18,34,48,56,48,34,18}/256 per
//                         axis, BORDER_REFLECT_101 -- src/ORBextractor.cc:1093-1094.
//   border_copy_kernel  : builds the 19-px reflect-101 frame on demand for orbx_get_level.
// All three are HBM-streaming byte kernels (bound: HBM bandwidth); no tensor-core work.
#include "extract.cuh"

namespace orbfront {

// One warp marches down a strip of destination rows; each lane owns 4 adjacent destination
// columns, keeps their horizontal taps in registers and reuses the horizontally interpolated
// source row shared by consecutive destination rows.
constexpr int kResizeStrip = 16;

__global__ void __launch_bounds__(128)
resize_level_kernel(const __grid_constant__ ExtractParams P, int l) {
  const LevelGeom& D = P.lv[l];
  const int f = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x0 = (blockIdx.x * 32 + lane) * 4;
  const int y0 = (blockIdx.y * 4 + warp) * kResizeStrip;
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

int launch_pyramid(const ExtractParams& P, cudaStream_t s, int* launches) {
  for (int l = 1; l < P.nlevels; ++l) {
    const LevelGeom& D = P.lv[l];
    dim3 grid(div_up(D.w, 128), div_up(D.h, 4 * kResizeStrip), P.n_frames);
    resize_level_kernel<<<grid, 128, 0, s>>>(P, l);
    ORB_LAUNCH_CHECK();
    ++*launches;
  }
  return ORB_OK;
}

// ------------------------------------------------------------------------------------
__device__ __forceinline__ int reflect101(int p, int n) {
  if (p < 0) p = -p;
  if (p >= n) p = 2 * n - 2 - p;
  return p;  // valid for |overshoot| < n, which holds for 3-px taps on levels >= 7 px
}

constexpr int kBlurTW = 64, kBlurTH = 32;

// grid: (tiles of all levels, 1, frames).  Tiles are numbered level by level.
__global__ void __launch_bounds__(256)
blur_kernel(const __grid_constant__ ExtractParams P, const int* __restrict__ tile_base /* nlevels+1 */) {
  __shared__ uint8_t raw[kBlurTH + 6][kBlurTW + 8];
  __shared__ uint16_t hor[kBlurTH + 6][kBlurTW];
  int l = 0;
  while (l + 1 < P.nlevels && (int)blockIdx.x >= tile_base[l + 1]) ++l;
  const LevelGeom& G = P.lv[l];
  const int f = blockIdx.z;
  if (P.kept_count[f * P.nlevels + l] == 0) return;  // reference skips levels without keypoints (:1089)
  const int t = blockIdx.x - tile_base[l];
  const int tiles_x = div_up(G.w, kBlurTW);
  const int tx = (t % tiles_x) * kBlurTW, ty = (t / tiles_x) * kBlurTH;
  const uint8_t* src = level_ptr(P, l, f);
  const int sp = level_pitch(P, l);
  for (int i = threadIdx.x; i < (kBlurTH + 6) * (kBlurTW + 6); i += 256) {
    const int r = i / (kBlurTW + 6), c = i % (kBlurTW + 6);
    const int sy = reflect101(ty + r - 3, G.h), sx = reflect101(tx + c - 3, G.w);
    raw[r][c] = src[(size_t)sy * sp + sx];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < (kBlurTH + 6) * kBlurTW; i += 256) {
    const int r = i / kBlurTW, c = i % kBlurTW;
    const uint8_t* p = &raw[r][c];
    hor[r][c] = (uint16_t)(18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3]);
  }
  __syncthreads();
  uint8_t* dst = P.blur + (size_t)f * P.blur_frame + G.off;
  for (int i = threadIdx.x; i < kBlurTH * kBlurTW / 4; i += 256) {
    const int r = i / (kBlurTW / 4), c = (i % (kBlurTW / 4)) * 4;
    if (ty + r >= G.h || tx + c >= G.w) continue;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned acc = 18u * (hor[r][c + k] + hor[r + 6][c + k]) + 34u * (hor[r + 1][c + k] + hor[r + 5][c + k]) +
                           48u * (hor[r + 2][c + k] + hor[r + 4][c + k]) + 56u * hor[r + 3][c + k];
      out |= ((acc + 32768u) >> 16) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dst + (size_t)(ty + r) * G.pitch + tx + c) = out;
  }
}

int launch_blur(const ExtractParams& P, cudaStream_t s, int* launches) {
  int total = 0;
  for (int l = 0; l < P.nlevels; ++l) total += div_up(P.lv[l].w, kBlurTW) * div_up(P.lv[l].h, kBlurTH);
  blur_kernel<<<dim3(total, 1, P.n_frames), 256, 0, s>>>(P, P.blur_tiles);
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
