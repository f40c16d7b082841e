// Shared declarations of the ORB extraction pipeline (pyramid -> FAST -> quadtree select ->
// orientation -> blur -> rBRIEF) for sm_100a.  Reference path: src/ORBextractor.cc:1051-1113.
//
// Data layout in HBM (per handle, frame-major so one frame's whole pyramid is contiguous and
// stays L2-resident while its stages run):
//   pyr     [frame][level >= 1][h_l][pitch_l]  u8   resized levels (level 0 aliases the input
//                                                    or a pitch-aligned upload buffer)
//   blur    [frame][level][h_l][pitch_l]        u8   7x7 sigma-2 fixed-point Gaussian of each level
//   cand    [frame][level][cand_cap_l]          u64  FAST candidates: score << 24 | y << 12 | x in the low
//                                                    half, any order; the quadtree kernel puts the path
//                                                    code in the high half
//   kept    [frame][level][kept_cap_l]          u32  selected keypoints in the reference's list order
//   out_kp  [frame][out_cap] orbx_keypoint, out_desc [frame][out_cap][32] u8, counts
// No 19-px border is materialised on the hot path: FAST, the orientation disc and the rBRIEF
// pattern never leave the image; only the blur's +-3 taps do, and they reflect in-kernel.
// orbx_get_level(with_border) builds the BORDER_REFLECT_101 frame on demand.
#pragma once
#include "common.cuh"

namespace orbfront {

constexpr int kEdge = 19;        // EDGE_THRESHOLD, ORBextractor.cc:74
constexpr int kHalfPatch = 15;   // HALF_PATCH_SIZE
constexpr int kMaxLevels = 16;
constexpr int kMaxDepth = 14;    // quadtree path digits (2 bits each) + 4 root bits = 32
constexpr int kRootBits = 4;
constexpr int kMaxBox = 4096;    // widest quadtree box supported (image width <= 4128)

struct LevelGeom {
  int w, h, pitch;          // level image; pitch in bytes (multiple of 64)
  size_t off;               // byte offset of this level inside one frame's pyr/blur block
  // FAST cell grid (ORBextractor.cc:771-787)
  int nCols, nRows, wCell, hCell, cell_base;
  uint32_t rcp_nCols;       // 2^32 / nCols + 1: x / nCols == __umulhi(x, rcp_nCols) while x * nCols < 2^32
  // quadtree (ORBextractor.cc:539-563)
  int boxW, boxH, nIni, quota, dmax;
  int qt_d0;                // depth of the finest grid of the sort-free quadtree kernel (0: this level always takes the sorting kernel)
  int qt_counts_global;     // its per-cell key counts live in the level's global scratch instead of shared memory (large quotas)
  float hX;
  // capacities / offsets (entries) inside one frame's cand / kept blocks
  int cand_cap, cand_off, kept_cap, kept_off;
  float scale;              // mvScaleFactor[level]
  float patch;              // (int)(PATCH_SIZE*scale), as float (KeyPoint::size)
  int xtab, ytab;           // offsets into the resize coefficient table
  int qx, qy;               // offsets of this level's x / y path-code tables in qcode
  int resize_words;         // every 4-px destination group reads a source window of <= 8 bytes: word-wise resize kernel
};

struct ExtractParams {
  LevelGeom lv[kMaxLevels];
  int nlevels, n_frames;
  int ini_th, min_th;
  // level 0 source (input alias or upload buffer)
  const uint8_t* l0;
  size_t l0_pitch, l0_frame;
  uint8_t* pyr;   size_t pyr_frame;    // levels >= 1
  uint8_t* blur;  size_t blur_frame;   // all levels
  unsigned long long* cand;  size_t cand_frame;  // entries per frame
  int* cand_count;                     // [frame][level]
  uint32_t* kept;  size_t kept_frame;
  int* kept_count;                     // [frame][level]
  uint8_t* qt_b;                       // global scratch for oversized (frame,level): boundary depths
  uint32_t* qt_seq;                    //   and creation sequence numbers, cand layout
  unsigned long long* qt_tmp;          //   and the radix sort's second key buffer, cand layout
  const int4* rtab;                    // resize taps: (src index, c0, c1, clamped src index + 1)
  const int* blur_tiles;               // [nlevels+1] first blur tile of each level
  const uint32_t* qcode;               // quadtree path-code tables (per level: boxW+1 x entries, boxH+1 y entries)
  orbx_keypoint* out_kp;  uint8_t* out_desc;  int out_cap;
  int* out_count;                      // [frame]
  int total_cells;                     // sum over levels of nCols*nRows
  int* pyr_bar;                        // grid-barrier counter of the fused pyramid kernel (null: per-level launches)
  int pyr_max_grid;                    // CTAs of that kernel that can be resident at once
};

__device__ __forceinline__ const uint8_t* level_ptr(const ExtractParams& P, int l, int f) {
  return l == 0 ? P.l0 + (size_t)f * P.l0_frame : P.pyr + (size_t)f * P.pyr_frame + P.lv[l].off;
}
__device__ __forceinline__ int level_pitch(const ExtractParams& P, int l) {
  return l == 0 ? (int)P.l0_pitch : P.lv[l].pitch;
}

// Quadtree path code of a candidate at border-relative (x,y): root strip, then one 2-bit quadrant
// digit (by * 2 + bx) per split with the reference's ceil-half boundaries (ExtractorNode::DivideNode,
// ORBextractor.cc:481-526; root assignment :545-569).  The x and y halvings are independent, so the
// host tabulates them per level (extractor.cu: build_path_tables) with the digits already spread to
// their bit positions: code = qcode[qx + x] | qcode[qy + y].

// kernel launchers (each returns an ORB_* code)
int launch_pyramid(const ExtractParams& P, cudaStream_t s, int* launches);
int pyramid_fused_max_grid(int device);
int launch_fast(const ExtractParams& P, cudaStream_t s, int* launches);
int launch_quadtree(const ExtractParams& P, cudaStream_t s, int* launches);
// after_quadtree: kept_count is valid, levels without keypoints are skipped like the reference does
int launch_blur(const ExtractParams& P, cudaStream_t s, int* launches, bool after_quadtree = true);
int launch_describe(const ExtractParams& P, cudaStream_t s, int* launches);
int launch_border_copy(const uint8_t* src, int w, int h, size_t pitch, uint8_t* dst, cudaStream_t s);
// all levels of frame f with their 19-px reflect-101 border, level l at dst + off[l] (row pitch w_l + 38)
int launch_border_all(const ExtractParams& P, int f, const size_t* off, uint8_t* dst, cudaStream_t s);

}  // namespace orbfront
