// The rest of Frame construction after extraction, one launch per frame (sm_100a):
//   Frame::UndistortKeyPoints      src/Frame.cc:419-449  (cv::undistortPoints(mat, mat, mK, mDistCoef, Mat(), mK))
//   Frame::ComputeImageBounds      src/Frame.cc:451-485  (the same call on the four image corners)
//   Frame::AssignFeaturesToGrid    src/Frame.cc:245-260  (+ PosInGrid, :397-408)
//   Frame::ComputeStereoFromRGBD   src/Frame.cc:703-727
// cv::undistortPoints as OpenCV 4.x computes it without a termination criterion argument: exactly five
// fixed-point iterations in double on the normalised point, then re-projection through P = K and a
// round to float.  The build runs with -fmad=false, so every double operation rounds on its own as the
// host code's does; division is IEEE.  One CTA per frame: the per-keypoint part is a stride loop, the
// cell lists (CSR, keypoint order inside a cell as push_back leaves it) come from a shared-memory
// histogram, a block scan, an atomic fill and a rank sort of every (short) cell list.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "common.cuh"

namespace orbfront {

namespace {

constexpr int kFrameThreads = 1024;
constexpr int kMaxCells = 24576;  // two int arrays of cells + 1 in shared memory (<= 192 KB)

struct CamD {
  double fx, fy, cx, cy, ifx, ify;
  double k[5];  // k1 k2 p1 p2 k3
  int distorted;
  float bf;
};

CamD make_cam(const orbx_camera& c) {
  CamD d;
  d.fx = c.fx; d.fy = c.fy; d.cx = c.cx; d.cy = c.cy;  // mK is CV_32F, widened to double by cvConvert
  d.ifx = 1. / d.fx; d.ify = 1. / d.fy;
  d.k[0] = c.k1; d.k[1] = c.k2; d.k[2] = c.p1; d.k[3] = c.p2; d.k[4] = c.k3;
  d.distorted = c.k1 != 0.0f;  // Frame.cc:421, :453: mDistCoef.at<float>(0) == 0.0 skips the call
  d.bf = c.bf;
  return d;
}

__device__ __forceinline__ void undistort_point(const CamD& c, float px, float py, float* ox, float* oy) {
  double x = px, y = py;
  x = (x - c.cx) * c.ifx;
  y = (y - c.cy) * c.ify;
  const double x0 = x, y0 = y;
  const double k1 = c.k[0], k2 = c.k[1], p1 = c.k[2], p2 = c.k[3], k3 = c.k[4];
#pragma unroll 1
  for (int j = 0; j < 5; ++j) {
    const double r2 = x * x + y * y;
    // the rational (k4..k6) and thin-prism (s1..s4) coefficients are zero for ORB-SLAM2's 4/5-coefficient
    // models: OpenCV's numerator 1 + ((k6*r2 + k5)*r2 + k4)*r2 is exactly 1 and its s-terms add +0
    const double icdist = 1. / (1 + ((k3 * r2 + k2) * r2 + k1) * r2);
    if (icdist < 0) {
      x = x0; y = y0;  // (u - cx)*ifx, (v - cy)*ify
      break;
    }
    const double deltaX = 2 * p1 * x * y + p2 * (r2 + 2 * x * x);
    const double deltaY = p1 * (r2 + 2 * y * y) + 2 * p2 * x * y;
    x = (x0 - deltaX) * icdist;
    y = (y0 - deltaY) * icdist;
  }
  // RR = P * I with P = K: rows (fx 0 cx), (0 fy cy), (0 0 1)
  const double xx = c.fx * x + 0. * y + c.cx;
  const double yy = 0. * x + c.fy * y + c.cy;
  const double ww = 1. / (0. * x + 0. * y + 1.);
  *ox = (float)(xx * ww);
  *oy = (float)(yy * ww);
}

__global__ void __launch_bounds__(kFrameThreads) undistort_kernel(const float2* __restrict__ in, int n, const CamD cam,
                                                                  float2* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float2 p = in[i];
  float2 q = p;
  if (cam.distorted) undistort_point(cam, p.x, p.y, &q.x, &q.y);
  out[i] = q;
}

struct FinishArgs {
  const orbx_keypoint* keys;
  int n;
  CamD cam;
  float min_x, min_y, width_inv, height_inv;
  int cols, rows;
  const float* depth;  // device, or nullptr
  int depth_w, depth_h;
  size_t depth_pitch;  // in floats
  float2* xy_un;
  float* u_right;
  float* depth_out;
  int* cell;        // n scratch: cell of keypoint i or -1
  uint32_t* tmp;    // n scratch: unordered fill
  int* cell_off;    // cols*rows + 1
  uint32_t* cell_idx;
};

__global__ void __launch_bounds__(kFrameThreads) frame_finish_kernel(const FinishArgs A) {
  extern __shared__ int sm[];
  const int ncells = A.cols * A.rows;
  int* cnt = sm;                   // ncells + 1: histogram -> exclusive offsets
  int* cursor = sm + ncells + 1;   // ncells: fill cursors
  __shared__ int warp_sum[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int c = tid; c <= ncells; c += kFrameThreads) cnt[c] = 0;
  __syncthreads();

  for (int i = tid; i < A.n; i += kFrameThreads) {
    const orbx_keypoint kp = A.keys[i];
    float xu = kp.x, yu = kp.y;
    if (A.cam.distorted) undistort_point(A.cam, kp.x, kp.y, &xu, &yu);
    A.xy_un[i] = make_float2(xu, yu);
    // PosInGrid: round() half away from zero on the float product, then the range test
    const float gx = roundf((xu - A.min_x) * A.width_inv), gy = roundf((yu - A.min_y) * A.height_inv);
    int c = -1;
    if (gx >= 0.f && gx < (float)A.cols && gy >= 0.f && gy < (float)A.rows) c = (int)gx * A.rows + (int)gy;
    A.cell[i] = c;
    if (c >= 0) atomicAdd(&cnt[c], 1);
    if (A.depth) {  // ComputeStereoFromRGBD: imDepth.at<float>(v, u) truncates the float coordinates
      float ur = -1.f, dz = -1.f;
      const int u = (int)kp.x, v = (int)kp.y;
      if (u >= 0 && u < A.depth_w && v >= 0 && v < A.depth_h) {
        const float d = A.depth[(size_t)v * A.depth_pitch + u];
        if (d > 0) { dz = d; ur = xu - A.cam.bf / d; }
      }
      A.u_right[i] = ur;
      A.depth_out[i] = dz;
    }
  }
  __syncthreads();

  // exclusive scan of cnt[0..ncells): every thread owns a run of `per` consecutive cells
  const int per = (ncells + kFrameThreads - 1) / kFrameThreads;
  const int c0 = min(tid * per, ncells), c1 = min(c0 + per, ncells);
  int mine = 0;
  for (int c = c0; c < c1; ++c) mine += cnt[c];
  int incl = mine;
#pragma unroll
  for (int s = 1; s < 32; s <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, s);
    if (lane >= s) incl += o;
  }
  if (lane == 31) warp_sum[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = warp_sum[lane], wi = w;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, wi, s);
      if (lane >= s) wi += o;
    }
    warp_sum[lane] = wi - w;
  }
  __syncthreads();
  int run = warp_sum[warp] + incl - mine;
  for (int c = c0; c < c1; ++c) {
    const int k = cnt[c];
    cnt[c] = run;
    cursor[c] = run;
    A.cell_off[c] = run;
    run += k;
  }
  __syncthreads();
  for (int i = tid; i < A.n; i += kFrameThreads) {
    const int c = A.cell[i];
    if (c >= 0) A.tmp[atomicAdd(&cursor[c], 1)] = (uint32_t)i;
  }
  __syncthreads();
  if (tid == 0) {
    const int total = ncells > 0 ? cursor[ncells - 1] : 0;
    cnt[ncells] = total;
    A.cell_off[ncells] = total;
  }
  __syncthreads();
  // rank sort inside every cell list: ascending keypoint index = push_back order
  const int total = cnt[ncells];
  for (int p = tid; p < total; p += kFrameThreads) {
    const uint32_t me = A.tmp[p];
    const int c = A.cell[me];
    const int b = cnt[c], e = cursor[c];
    int rank = 0;
    for (int j = b; j < e; ++j) rank += A.tmp[j] < me;
    A.cell_idx[b + rank] = me;
  }
}

}  // namespace
}  // namespace orbfront

using namespace orbfront;

extern "C" {

int orbx_undistort_points(int device, const float* xy, int n, const orbx_camera* cam, float* xy_out) {
  if (n < 0 || !cam || (n > 0 && (!xy || !xy_out))) return set_error(ORB_E_INVALID, "orbx_undistort_points: bad argument");
  int rc = use_device(device);
  if (rc) return rc;
  if (n == 0) return ORB_OK;
  if (!(cam->fx != 0.f) || !(cam->fy != 0.f)) return set_error(ORB_E_INVALID, "orbx_undistort_points: zero focal length");
  Layout L;
  const size_t o_in = L.place((size_t)n * 8), o_out = L.place((size_t)n * 8);
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  memcpy(S->host + o_in, xy, (size_t)n * 8);
  ORB_CUDA_TRY(cudaMemcpyAsync(S->dev + o_in, S->host + o_in, (size_t)n * 8, cudaMemcpyHostToDevice, S->stream));
  undistort_kernel<<<div_up(n, kFrameThreads), kFrameThreads, 0, S->stream>>>((const float2*)(S->dev + o_in), n, make_cam(*cam),
                                                                               (float2*)(S->dev + o_out));
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(S->host + o_out, S->dev + o_out, (size_t)n * 8, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  memcpy(xy_out, S->host + o_out, (size_t)n * 8);
  return ORB_OK;
}

int orbx_frame_finish(int device, const orbx_keypoint* keys, int n, const orbx_camera* cam, const orbx_frame_bounds* grid,
                      const float* depth, int depth_w, int depth_h, size_t depth_stride, int depth_memory, float* xy_un,
                      float* u_right, float* depth_out, int32_t* cell_off, uint32_t* cell_idx, int* n_in_grid) {
  if (n < 0 || !cam || !grid || !cell_off || (n > 0 && (!keys || !xy_un || !cell_idx)))
    return set_error(ORB_E_INVALID, "orbx_frame_finish: bad argument");
  if (grid->cols < 1 || grid->rows < 1 || (long long)grid->cols * grid->rows > kMaxCells)
    return set_error(ORB_E_INVALID, "orbx_frame_finish: grid %d x %d (1..%d cells supported)", grid->cols, grid->rows, kMaxCells);
  if (depth && (!u_right || !depth_out || depth_w <= 0 || depth_h <= 0 || depth_stride < (size_t)depth_w * 4 || depth_stride % 4 ||
                (depth_memory != 0 && depth_memory != 1)))
    return set_error(ORB_E_INVALID, "orbx_frame_finish: bad depth image");
  int rc = use_device(device);
  if (rc) return rc;
  if (cam->k1 != 0.f && (!(cam->fx != 0.f) || !(cam->fy != 0.f))) return set_error(ORB_E_INVALID, "orbx_frame_finish: zero focal length");
  const int ncells = grid->cols * grid->rows;
  const bool up_depth = depth && depth_memory == 0;
  Layout L;
  const size_t o_keys = L.place((size_t)n * sizeof(orbx_keypoint));
  const size_t o_depth = up_depth ? L.place(depth_stride * (size_t)depth_h) : 0;
  const size_t o_up_end = L.bytes;
  const size_t o_cell = L.place((size_t)n * 4), o_tmp = L.place((size_t)n * 4);
  const size_t o_down = L.bytes;
  const size_t o_xy = L.place((size_t)n * 8), o_ur = L.place((size_t)n * 4), o_dz = L.place((size_t)n * 4);
  const size_t o_off = L.place((size_t)(ncells + 1) * 4), o_idx = L.place((size_t)n * 4);
  CallScratch* S = nullptr;
  if ((rc = call_scratch(device, L.bytes, &S))) return rc;
  if (n) memcpy(S->host + o_keys, keys, (size_t)n * sizeof(orbx_keypoint));
  if (up_depth) memcpy(S->host + o_depth, depth, depth_stride * (size_t)depth_h);
  ORB_CUDA_TRY(cudaMemcpyAsync(S->dev, S->host, o_up_end, cudaMemcpyHostToDevice, S->stream));
  FinishArgs A;
  A.keys = (const orbx_keypoint*)(S->dev + o_keys);
  A.n = n;
  A.cam = make_cam(*cam);
  A.min_x = grid->min_x; A.min_y = grid->min_y; A.width_inv = grid->width_inv; A.height_inv = grid->height_inv;
  A.cols = grid->cols; A.rows = grid->rows;
  A.depth = !depth ? nullptr : up_depth ? (const float*)(S->dev + o_depth) : depth;
  A.depth_w = depth_w; A.depth_h = depth_h; A.depth_pitch = depth_stride / 4;
  A.xy_un = (float2*)(S->dev + o_xy);
  A.u_right = (float*)(S->dev + o_ur);
  A.depth_out = (float*)(S->dev + o_dz);
  A.cell = (int*)(S->dev + o_cell);
  A.tmp = (uint32_t*)(S->dev + o_tmp);
  A.cell_off = (int*)(S->dev + o_off);
  A.cell_idx = (uint32_t*)(S->dev + o_idx);
  const size_t smem = (size_t)(2 * ncells + 1) * sizeof(int);
  static thread_local size_t smem_set[16] = {};
  if (smem > 48 * 1024 && smem > smem_set[device]) {
    ORB_CUDA_TRY(cudaFuncSetAttribute(frame_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set[device] = smem;
  }
  frame_finish_kernel<<<1, kFrameThreads, smem, S->stream>>>(A);
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(S->host + o_down, S->dev + o_down, L.bytes - o_down, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  if (n) memcpy(xy_un, S->host + o_xy, (size_t)n * 8);
  if (depth && n) {
    memcpy(u_right, S->host + o_ur, (size_t)n * 4);
    memcpy(depth_out, S->host + o_dz, (size_t)n * 4);
  }
  memcpy(cell_off, S->host + o_off, (size_t)(ncells + 1) * 4);
  const int total = cell_off[ncells];
  if (total) memcpy(cell_idx, S->host + o_idx, (size_t)total * 4);
  if (n_in_grid) *n_in_grid = total;
  return ORB_OK;
}

}  // extern "C"
