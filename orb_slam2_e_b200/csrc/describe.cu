// Orientation (intensity centroid) + steered rBRIEF descriptor + output assembly (sm_100a).
// One warp per selected keypoint.
//
//   IC_Angle              src/ORBextractor.cc:77-104   integer moments over the radius-15 disc of the
//                                                      UN-blurred level, cv::fastAtan2 in degrees
//   computeOrbDescriptor  src/ORBextractor.cc:107-147  256 tests on the BLURRED level at pattern points
//                                                      rotated by the keypoint angle (cvRound'ed)
//   operator() tail       src/ORBextractor.cc:1101-1111 pt *= mvScaleFactor[level], level-ascending output
// The moments are exact integers (warp-shuffle reduction); fastAtan2 and the rotation use
// individually rounded float ops (__f*_rn, no FMA contraction) to match the host arithmetic.
#include "extract.cuh"

namespace orbfront {

__device__ const int8_t g_pattern[1024] = {
#include "orb_pattern.inc"
};

// circular patch row half-widths (umax), ORBextractor.cc:452-469 evaluated for HALF_PATCH_SIZE = 15
__device__ __constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
  constexpr float scale = (float)(180.0 / 3.14159265358979323846);
  constexpr float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale,
                  p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
  constexpr float eps = 2.2204460492503131e-16f;  // (float)DBL_EPSILON
  const float ax = fabsf(x), ay = fabsf(y);
  float a, c, c2;
  if (ax >= ay) {
    c = __fdiv_rn(ay, __fadd_rn(ax, eps));
    c2 = __fmul_rn(c, c);
    a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = __fdiv_rn(ax, __fadd_rn(ay, eps));
    c2 = __fmul_rn(c, c);
    a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0) a = __fsub_rn(180.f, a);
  if (y < 0) a = __fsub_rn(360.f, a);
  return a;
}

// sin and cos of x in [0, 2*pi] in double (quadrant reduction with a two-part pi/2 and the classic
// degree-13/14 kernels, error < 1 ulp of double): rounded to float these equal the correctly rounded
// sinf/cosf the host library returns, except in vanishingly rare double-rounding cases.  About half
// the instructions of the general-range library sincos.
__device__ __forceinline__ void sincos_2pi(double x, double* sn, double* cs) {
  const double q = rint(x * 6.36619772367581382433e-01);
  double r = __fma_rn(-q, 1.57079632673412561417e+00, x);
  r = __fma_rn(-q, 6.07710050650619224932e-11, r);
  const double z = r * r;
  double ps = __fma_rn(z, 1.58969099521155010221e-10, -2.50507602534068634195e-08);
  ps = __fma_rn(z, ps, 2.75573137070700676789e-06);
  ps = __fma_rn(z, ps, -1.98412698298579493134e-04);
  ps = __fma_rn(z, ps, 8.33333333332248946124e-03);
  ps = __fma_rn(z, ps, -1.66666666666666324348e-01);
  const double sr = __fma_rn(z * r, ps, r);
  double pc = __fma_rn(z, -1.13596475577881948265e-11, 2.08757232129817482790e-09);
  pc = __fma_rn(z, pc, -2.75573143513906633035e-07);
  pc = __fma_rn(z, pc, 2.48015872894767294178e-05);
  pc = __fma_rn(z, pc, -1.38888888888741095749e-03);
  pc = __fma_rn(z, pc, 4.16666666666666019037e-02);
  const double cr = __fma_rn(z * z, pc, __fma_rn(z, -0.5, 1.0));
  const int iq = (int)q;
  const double s0 = (iq & 1) ? cr : sr, c0 = (iq & 1) ? sr : cr;
  *sn = (iq & 2) ? -s0 : s0;
  *cs = ((iq + 1) & 2) ? -c0 : c0;
}

constexpr int kDescWarps = 8;
constexpr int kPatchR = 18;                 // rotated pattern coordinates stay within +-18 (|p| <= 18.38, cvRound)
constexpr int kPatchRows = 2 * kPatchR + 1; // 37
constexpr int kPatchWords = 10;             // 37 bytes at any 4-byte misalignment fit 10 aligned words
constexpr int kPatchPitch = 4 * kPatchWords;

// cvRound for |v| < 2^22 without the XU pipe: adding 1.5 * 2^23 (12582912.0f, bits 0x4B400000) leaves
// round-half-even(v) in the mantissa; the kernel folds the constant into the patch offset.

constexpr int kKpPerWarp = 4;  // keypoint slots per warp: amortises the per-CTA pattern table fill

// grid: (ceil(kept slots per frame / (8 warps x 4)), frames)
__global__ void __launch_bounds__(kDescWarps * 32)
orient_describe_kernel(const __grid_constant__ ExtractParams P, int slots_per_frame) {
  // pattern as floats, transposed so that lane (= descriptor byte) is the fastest index:
  // patf[(t * 2 + e) * 32 + lane] = point e of test t of byte `lane`
  __shared__ float2 patf[16 * 32];
  __shared__ __align__(16) uint32_t patch_s[kDescWarps][(kPatchRows + 2) * kPatchWords];  // 39 rows: 13 x 3
  for (int i = threadIdx.x; i < 512; i += kDescWarps * 32) {
    const int ln = i & 31, te = i >> 5;
    patf[i] = make_float2((float)g_pattern[32 * ln + 2 * te], (float)g_pattern[32 * ln + 2 * te + 1]);
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int f = blockIdx.y;
  // per-level keypoint counts of the frame -> exclusive prefix (output row of a level's first keypoint)
  int cnt = lane < P.nlevels ? P.kept_count[f * P.nlevels + lane] : 0, incl = cnt;
#pragma unroll
  for (int s = 1; s < kMaxLevels; s <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, s);
    if (lane >= s) incl += o;
  }
  const int total = __shfl_sync(0xffffffffu, incl, kMaxLevels - 1);
  if (blockIdx.x == 0 && threadIdx.x == 0) P.out_count[f] = min(total, P.out_cap);

  // lane constants: orientation column u = lane - 15 with its disc half-height; staging slot
  // (3 patch rows x 10 words per warp load)
  const int u = lane - kHalfPatch;
  const int vlim = lane < 31 ? c_umax[u < 0 ? -u : u] : -1;  // the disc is symmetric: |u| <= umax[|v|] <=> |v| <= umax[|u|]
  const int srow = lane / kPatchWords, scol = lane - srow * kPatchWords;  // lanes 30, 31 idle while staging

  const int g0 = (blockIdx.x * kDescWarps + warp) * kKpPerWarp;
  int l = 0;
#pragma unroll 1
  for (int g = g0; g < min(g0 + kKpPerWarp, slots_per_frame); ++g) {
    while (l + 1 < P.nlevels && g >= P.lv[l + 1].kept_off) ++l;  // slots ascend, so does their level
    const LevelGeom& G = P.lv[l];
    const int k = g - G.kept_off;
    const int n_l = __shfl_sync(0xffffffffu, cnt, l), before = __shfl_sync(0xffffffffu, incl, l) - n_l;
    if (k >= n_l) continue;
    const int at = before + k;
    if (at >= P.out_cap) continue;

    const uint32_t pay = P.kept[(size_t)f * P.kept_frame + G.kept_off + k];
    const int x = (int)(pay & 0xfff) + (kEdge - 3), y = (int)((pay >> 12) & 0xfff) + (kEdge - 3);  // level coords
    const int score = (int)(pay >> 24);

    // --- stage the 37x37 blurred patch around the keypoint as aligned words (coalesced, 3 rows per
    // warp load) -- the 512 pattern samples then come from shared memory instead of 16 scattered
    // 32-sector global gathers.  Keypoints lie in [19, w-19) x [19, h-19), so the patch is inside the
    // image; the last word may reach into the row padding (pitch >= w + 4), never past it.  Rows
    // 37 and 38 of the last load are never sampled and are clamped to the image.
    const int px0 = x - kPatchR, shift = px0 & 3;
    __syncwarp();  // previous keypoint's samples are done
    if (lane < 3 * kPatchWords) {
      const uint8_t* blur_l = P.blur + (size_t)f * P.blur_frame + G.off;
      const uint32_t* bsrc = reinterpret_cast<const uint32_t*>(blur_l + (size_t)(y - kPatchR + srow) * G.pitch + (px0 - shift)) + scol;
      const int step = 3 * (G.pitch >> 2);
      uint32_t* dstp = patch_s[warp] + lane;
#pragma unroll
      for (int it = 0; it < (kPatchRows + 2) / 3; ++it) {
        // the last load covers patch rows 36..38; rows 37 and 38 are never sampled and may lie below the image
        if (it < (kPatchRows + 2) / 3 - 1 || srow == 0) dstp[it * 3 * kPatchWords] = __ldg(bsrc + it * step);
      }
    }

    // --- orientation: lane <-> column u; m10 = u * (column sum), m01 = sum of v * pixel
    const int sp = level_pitch(P, l);
    int colsum = 0, m01 = 0;
    {
      // rows +v and -v together: one bound test, colsum += a + b, m01 += v * (a - b)
      const uint8_t* cp = level_ptr(P, l, f) + (size_t)y * sp + (x + u);
      if (vlim >= 0) colsum = cp[0];
      const uint8_t* pp = cp;  // rows y + v and y - v by running pointers: one 64-bit add per load
      const uint8_t* pm = cp;
#pragma unroll
      for (int v = 1; v <= kHalfPatch; ++v) {
        pp += sp;
        pm -= sp;
        asm("" : "+l"(pp), "+l"(pm));  // keep them as pointers: the compiler otherwise re-derives base + offset (two more adds per load)
        if (v <= vlim) {
          const int a = *pp, b = *pm;
          colsum += a + b;
          m01 += v * (a - b);
        }
      }
    }
    int m10 = u * colsum;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
      m10 += __shfl_xor_sync(0xffffffffu, m10, s);
      m01 += __shfl_xor_sync(0xffffffffu, m01, s);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10);

    // --- steered rBRIEF on the blurred patch: lane <-> descriptor byte
    float ca = 0.f, sa = 0.f;
    if (lane == 0) {
      constexpr float factorPI = (float)(3.14159265358979323846 / 180.0);
      const float rad = __fmul_rn(angle, factorPI);
      double ds, dc;
      sincos_2pi((double)rad, &ds, &dc);
      ca = (float)dc;
      sa = (float)ds;
    }
    ca = __shfl_sync(0xffffffffu, ca, 0);
    sa = __shfl_sync(0xffffffffu, sa, 0);
    __syncwarp();  // patch visible to every lane
    // cvRound of the rotated coordinates is the add-1.5*2^23 trick; the constant left in the float's
    // bit pattern (0x4B400000 per coordinate) is taken out of the patch offset once per keypoint
    // instead of once per coordinate: offset = ybits * pitch + xbits + kbase (mod 2^32)
    const uint8_t* patch_bytes = reinterpret_cast<const uint8_t*>(&patch_s[0][0]);
    const uint32_t kbase = (uint32_t)(warp * (int)sizeof(patch_s[0]) + kPatchR * kPatchPitch + kPatchR + shift) -
                           0x4B400000u * (uint32_t)(kPatchPitch + 1);
    uint32_t byte = 0;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      int val[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const float2 p = patf[(t * 2 + e) * 32 + lane];
        const uint32_t yb = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(p.x, sa), __fmul_rn(p.y, ca)), 12582912.0f));
        const uint32_t xb = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(p.x, ca), __fmul_rn(p.y, sa)), 12582912.0f));
        val[e] = patch_bytes[yb * (uint32_t)kPatchPitch + kbase + xb];
      }
      byte |= (uint32_t)(val[0] < val[1]) << t;
    }
    P.out_desc[((size_t)f * P.out_cap + at) * 32 + lane] = (uint8_t)byte;
    if (lane == 0) {
      orbx_keypoint kp;
      kp.x = (float)x; kp.y = (float)y;
      if (l != 0) { kp.x = __fmul_rn(kp.x, G.scale); kp.y = __fmul_rn(kp.y, G.scale); }
      kp.size = G.patch;
      kp.angle = angle;
      kp.response = (float)score;
      kp.octave = l;
      kp.class_id = -1;
      P.out_kp[(size_t)f * P.out_cap + at] = kp;
    }
  }
}

int launch_describe(const ExtractParams& P, cudaStream_t s, int* launches) {
  NvtxRange nvtx_range("orbx:describe");
  const LevelGeom& last = P.lv[P.nlevels - 1];
  const int slots = last.kept_off + last.kept_cap;
  orient_describe_kernel<<<dim3(div_up(slots, kDescWarps * kKpPerWarp), P.n_frames), kDescWarps * 32, 0, s>>>(P, slots);
  ORB_LAUNCH_CHECK();
  ++*launches;
  return ORB_OK;
}

}  // namespace orbfront
