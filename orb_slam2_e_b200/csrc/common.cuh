// Shared host/device helpers for liborbfront.so (error reporting, launch accounting,
// mbarrier / TMA bulk-copy wrappers for sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>  // header-only: ranges are no-ops unless a profiler injects the NVTX library

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/orbfront.h"

namespace orbfront {

int set_error(int code, const char* fmt, ...);
extern std::atomic<long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }

#define ORB_CUDA_TRY(expr)                                                                    \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess)                                                                    \
      return ::orbfront::set_error(ORB_E_CUDA, "%s failed: %s (%s:%d)", #expr,                \
                                   cudaGetErrorString(_e), __FILE__, __LINE__);               \
  } while (0)

#define ORB_LAUNCH_CHECK()                                                                    \
  do {                                                                                        \
    ::orbfront::count_launch();                                                               \
    cudaError_t _e = cudaGetLastError();                                                      \
    if (_e != cudaSuccess)                                                                    \
      return ::orbfront::set_error(ORB_E_CUDA, "kernel launch failed: %s (%s:%d)",            \
                                   cudaGetErrorString(_e), __FILE__, __LINE__);               \
  } while (0)

// Selects the device; ORB_E_NO_DEVICE when there is none (no CPU fallback anywhere).
int use_device(int device);

// NVTX range around one stage / entry point on the calling host thread (tracing); kernels launched
// inside show up under it in a timeline
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

// Per-thread, per-device scratch of the small host-buffer entry points (SearchByBoW, stereo search,
// pairwise distances): a grow-only device buffer, a page-locked host mirror of the same size and a
// stream, so a call is pack -> one H2D -> kernels -> one D2H -> one sync with no cudaMalloc/cudaFree
// (which cost more than the work itself).  Re-entrant across threads, as the three SLAM threads need.
struct CallScratch {
  uint8_t* dev = nullptr;
  uint8_t* host = nullptr;  // pinned
  size_t cap = 0;
  cudaStream_t stream = nullptr;
};
// Returns the calling thread's scratch for `device` with at least `bytes` in both buffers.
int call_scratch(int device, size_t bytes, CallScratch** out);
// Offsets of 256-byte aligned pieces inside one scratch block.
struct Layout {
  size_t bytes = 0;
  size_t place(size_t n) {
    const size_t at = bytes;
    bytes = (bytes + n + 255) & ~(size_t)255;
    return at;
  }
};

// Device-resident pyramid of one frame of an extractor handle (defined in extractor.cu), for kernels
// outside the extractor that read mvImagePyramid in place (stereo refinement).
struct PyramidView {
  enum { kLevels = 16 };
  const uint8_t* px[kLevels];
  int pitch[kLevels], w[kLevels], h[kLevels];
  float scale[kLevels], inv_scale[kLevels];
  int n_levels, device;
};
int pyramid_view(orbx_handle* h, int frame, PyramidView* out);  // waits for the handle's stream
// Device-resident results of an extractor handle's last batch, for kernels that consume them in place
// (batched stereo matching).  frame f: keypoints kp + f * out_cap, descriptors desc + 32 * f * out_cap, count[f].
struct BatchView {
  const orbx_keypoint* kp; const uint8_t* desc; const int* count;
  int out_cap, n_frames, device;
  size_t level0_frame_stride, level_frame_stride;  // bytes between consecutive frames of level 0 / of levels >= 1
  cudaStream_t stream;                             // the handle's stream (work queued there is ordered after the extraction)
};
int batch_view(orbx_handle* h, BatchView* out);  // does not wait

#ifdef __CUDACC__
#define ORB_HD __host__ __device__
#else
#define ORB_HD
#endif
template <typename T>
ORB_HD inline T div_up(T a, T b) { return (a + b - 1) / b; }
template <typename T>
ORB_HD inline T align_up(T a, T b) { return div_up(a, b) * b; }

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D TMA bulk copy global -> shared, completion counted in bytes on an mbarrier.
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
#endif

}  // namespace orbfront
