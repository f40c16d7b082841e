// Probe 2: programming-guide style TMA 2-D tile load (libcu++ wrappers) vs raw PTX variants.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
#include <cstdio>
#include <cstdint>
#include <cstring>
#include <vector>
namespace cde = cuda::device::experimental;
using barrier = cuda::barrier<cuda::thread_scope_block>;
constexpr int BW = 48, BH = 43;
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
#if VARIANT == 0
__global__ void k(const __grid_constant__ CUtensorMap tm, int x, int y, uint8_t* out) {
  __shared__ alignas(128) uint8_t sm[BH][BW];
  __shared__ barrier bar;
  if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
  __syncthreads();
  barrier::arrival_token token;
  if (threadIdx.x == 0) {
    cde::cp_async_bulk_tensor_2d_global_to_shared(&sm, &tm, x, y, bar);
    token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(sm));
  } else token = bar.arrive();
  bar.wait(std::move(token));
  for (int i = threadIdx.x; i < BW * BH; i += 32) out[i] = (&sm[0][0])[i];
}
#else
__global__ void k(const __grid_constant__ CUtensorMap tm, int x, int y, uint8_t* out, const CUtensorMap* gtm) {
  __shared__ alignas(128) uint8_t sm[BH * BW];
  __shared__ alignas(8) uint64_t bar;
  const CUtensorMap* desc = VARIANT == 2 ? gtm : &tm;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&bar)), "r"(BW * BH) : "memory");
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(s32(sm)), "l"(reinterpret_cast<uint64_t>(desc)), "r"(x), "r"(y), "r"(s32(&bar)) : "memory");
  }
  __syncwarp();
  asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(s32(&bar)) : "memory");
  for (int i = threadIdx.x; i < BW * BH; i += 32) out[i] = sm[i];
}
#endif
int main() {
  const int w = 640, h = 480, pitch = 640;
  std::vector<uint8_t> img((size_t)pitch * h);
  for (size_t i = 0; i < img.size(); ++i) img[i] = (uint8_t)((i * 2654435761u) >> 13);
  uint8_t *d, *o; cudaMalloc(&d, img.size()); cudaMalloc(&o, 8192);
  cudaMemcpy(d, img.data(), img.size(), cudaMemcpyHostToDevice);
  void* fn = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
  typedef CUresult (*enc_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  CUtensorMap tm; memset(&tm, 0, sizeof(tm));
  cuuint64_t dims[2] = {w, h}; cuuint64_t str[1] = {pitch}; cuuint32_t box[2] = {BW, BH}; cuuint32_t es[2] = {1, 1};
  CUresult r = ((enc_t)fn)(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("variant %d x %d: encode %d; ", VARIANT, XC, (int)r);
  const int x = XC, y = 450;
#if VARIANT == 0
  k<<<1, 32>>>(tm, x, y, o);
#else
  CUtensorMap* g; cudaMalloc(&g, sizeof(tm)); cudaMemcpy(g, &tm, sizeof(tm), cudaMemcpyHostToDevice);
  k<<<1, 32>>>(tm, x, y, o, g);
#endif
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s; ", cudaGetErrorString(e));
  if (e == cudaSuccess) {
    std::vector<uint8_t> got(BW * BH); cudaMemcpy(got.data(), o, got.size(), cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r2 = 0; r2 < BH; ++r2) for (int c = 0; c < BW; ++c) {
      const int xx = x + c, yy = y + r2;
      const uint8_t want = (xx < w && yy < h) ? img[(size_t)yy * pitch + xx] : 0;
      bad += got[r2 * BW + c] != want;
    }
    printf("mismatches %d", bad);
  }
  printf("\n");
  return 0;
}
