/* orbfront.h -- C ABI of the B200-native ORB front end (liborbfront.so).
 *
 * This is the drop-in boundary for the reference's data-parallel front end.  The reference
 * exposes it as two C++ classes inside libORB_SLAM2_E.so (there is no FFI in the reference);
 * each entry point below names the reference interface it replaces.  The source-compatible
 * C++ classes (include/ORBextractor.h, include/ORBmatcher.h) are thin wrappers over this ABI.
 *
 * Conventions: plain C types only; every function returns ORB_OK (0) or a negative ORB_E_*
 * code and records a thread-local message readable through orb_last_error().  No function
 * falls back to a CPU implementation: without a CUDA device they fail with ORB_E_NO_DEVICE.
 * Handles are not thread-safe but independent (one CUDA stream each -- a handle also owns a few
 * internal streams that fork from and join back into it, so ordering against the handle's stream
 * orders against all of a call's work), matching the reference,
 * which never shares an ORBextractor between threads (Frame.cc:78-81) and constructs a stack
 * ORBmatcher per call site (Tracking.cc:1750).
 */
#ifndef ORBFRONT_H_
#define ORBFRONT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  ORB_OK = 0,
  ORB_E_INVALID = -1,    /* bad argument */
  ORB_E_NO_DEVICE = -2,  /* no usable CUDA device */
  ORB_E_CUDA = -3,       /* CUDA runtime error (message has the detail) */
  ORB_E_CAPACITY = -4,   /* caller buffer too small; *n_out holds the needed count */
  ORB_E_STATE = -5       /* call order violated (e.g. reading results before extracting) */
};

const char* orb_last_error(void);
/* Number of CUDA devices visible (0 when none); never fails. */
int orb_device_count(void);
/* Kernel launches issued by this library in the calling process so far (for bench accounting). */
long long orb_launch_count(void);

/* ------------------------------------------------------------------ extractor ---------- */

typedef struct orbx_handle orbx_handle;

/* ORBextractor::ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST,
 * int minThFAST) -- reference: include/ORBextractor.h:53-54, src/ORBextractor.cc:410-470.
 * device/max_* size the device buffers once; with max_* > 0 a larger frame or batch is rejected with
 * ORB_E_INVALID by every extract call (no reallocation on the hot path); max_* = 0 lets the buffers grow on demand. */
typedef struct {
  int nfeatures;
  float scale_factor;
  int nlevels;
  int ini_th_fast;
  int min_th_fast;
  int device;     /* CUDA ordinal */
  int max_width;  /* largest frame accepted */
  int max_height;
  int max_batch;  /* frames per orbx_extract_batch call */
} orbx_config;

/* Same field order and size (28 bytes) as cv::KeyPoint, so a std::vector<cv::KeyPoint>'s
 * storage can be passed directly. */
typedef struct {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orbx_keypoint;

int orbx_create(const orbx_config* cfg, orbx_handle** out);
int orbx_destroy(orbx_handle* h);

/* Scale tables the reference's inline getters return (ORBextractor.h:64-84) and the
 * per-level quota (mnFeaturesPerLevel).  which: 0 scale, 1 inverse scale, 2 sigma^2,
 * 3 inverse sigma^2. */
int orbx_get_scale_table(const orbx_handle* h, int which, float* out /* nlevels */);
int orbx_get_level_quota(const orbx_handle* h, int* out /* nlevels */);

/* ORBextractor::operator()(image, mask, keypoints, descriptors) -- reference:
 * include/ORBextractor.h:60-62, src/ORBextractor.cc:1051-1113.  img is a HOST CV_8UC1 buffer
 * with `stride` bytes per row; the mask argument of the reference is ignored there
 * (ORBextractor.h:59) and has no counterpart here.  Writes up to cap keypoints / cap*32
 * descriptor bytes in the reference's output order and sets *n_out.  An empty image
 * (w or h <= 0, or img NULL) returns ORB_OK with *n_out = 0, as :1054-1055 does. */
int orbx_extract(orbx_handle* h, const uint8_t* img, int w, int ht, size_t stride,
                 orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out);

/* Batched form (frames are independent): n frames of identical size, frame f at
 * imgs + f*frame_stride.  memory = 0: imgs is host memory (pinned or pageable), copied inside
 * the call.  memory = 1: imgs is device memory on the handle's device.  Results stay on the
 * device until read.  Lifetime with memory = 1: frames whose stride and frame_stride are multiples of 16 are
 * read IN PLACE as pyramid level 0, also after the call returns (orbx_get_level(level 0), orbx_get_pyramid_host,
 * orbx_stereo_matches[_batch]); the buffer must stay valid and unmodified until the next extract call on the
 * handle or its destruction.  Other layouts are copied into the handle. */
int orbx_extract_batch(orbx_handle* h, const uint8_t* imgs, int n, int w, int ht, size_t stride,
                       size_t frame_stride, int memory);
/* Blocks until the batch is done; counts[f] = keypoints of frame f. */
int orbx_batch_counts(orbx_handle* h, int* counts /* n */);
/* Copies frame f's results (reference output order) to host buffers. */
int orbx_batch_read(orbx_handle* h, int frame, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out);
/* Copies every frame's results in one transfer: frame f occupies rows [f*cap_per_frame, ...). */
int orbx_batch_read_all(orbx_handle* h, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* counts);
/* Asynchronous form of orbx_batch_read_all (no reference counterpart: a Frame keeps its keypoints and descriptors on the
 * host, Frame.cc:262-268, and a stereo Frame then matches them, :527-700 -- with the matching on the device
 * (orbx_stereo_matches_batch) the download need not wait for it, nor it for the download).  _begin queues the
 * copies behind the extraction on the handle's download stream and returns; other calls on the handle run beside the
 * transfer; _end waits for it.  The host buffers must stay valid until _end and should be page-locked (pageable memory
 * makes _begin itself wait).  The next extract call on the handle waits for an unfinished read. */
int orbx_batch_read_begin(orbx_handle* h, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* counts);
int orbx_batch_read_end(orbx_handle* h);
/* Rows per frame that orbx_batch_read_all / orbx_extract_batch_host need for w x ht frames
 * (sum over levels of quota + 4 * quadtree roots + 3; the reference's DistributeOctTree may exceed
 * a level's quota by up to 3). */
int orbx_frame_capacity(orbx_handle* h, int w, int ht, int* cap);
/* Host-to-host batch in one call: frames are uploaded, extracted and their results downloaded in
 * chunks of chunk_frames (<= 0: a sixteenth of the batch, at least 32) on copy streams either side of four rotating
 * compute streams, so PCIe traffic in both directions overlaps the kernels and one chunk's tails overlap the next chunks.  Pinned host buffers give full overlap; pageable ones still work.  Output
 * layout as orbx_batch_read_all (frame f at row f*cap_per_frame; counts[f] keypoints).  Blocks
 * until every result is in host memory. */
int orbx_extract_batch_host(orbx_handle* h, const uint8_t* imgs, int n, int w, int ht, size_t stride,
                            size_t frame_stride, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame,
                            int* counts, int chunk_frames);
/* Wait for the handle's stream. */
int orbx_sync(orbx_handle* h);

/* mvImagePyramid[level] (public member read by Frame.cc:534,624-641): size and pixels of
 * the level computed by the last extract call, for `frame` of the batch.  with_border != 0
 * copies the (h+38)x(w+38) buffer including the 19-px BORDER_REFLECT_101 frame
 * (src/ORBextractor.cc:1115-1140). */
int orbx_level_size(const orbx_handle* h, int level, int* w, int* ht);
int orbx_get_level(orbx_handle* h, int frame, int level, uint8_t* dst, size_t dst_stride, int with_border);
/* Single-frame drop-in path (ORBextractor::operator() called by Frame.cc:262-268 on one image).  With
 * staging enabled, every single-frame orbx_extract_batch also builds all bordered levels with ONE kernel
 * and copies them with ONE device-to-host transfer into page-locked memory owned by the handle, on a side
 * stream beside FAST / quadtree / blur / descriptors (the pyramid is the first stage to finish).
 * orbx_get_pyramid_host waits for that copy and returns, per level, the address of the level's pixel (0,0)
 * inside its (h+38) x (w+38) bordered buffer and that buffer's row stride -- what the reference's
 * mvImagePyramid[level] cv::Mat ROI points at (src/ORBextractor.cc:1115-1140).  The memory is rewritten by
 * the next extract call on the handle and freed by orbx_destroy: a Frame reads mvImagePyramid only while it is
 * constructed (Frame.cc:534,624-641), i.e. before its extractor is called again. */
int orbx_set_pyramid_staging(orbx_handle* h, int enabled);
int orbx_get_pyramid_host(orbx_handle* h, int level, const uint8_t** pixel00, int* w, int* ht, size_t* stride);
/* Frame 0's results with one synchronisation: count, keypoints and descriptors are copied to page-locked
 * staging in one go and the first *n_out rows handed to the caller (orbx_batch_counts + orbx_batch_read
 * synchronise twice). */
int orbx_read_frame0(orbx_handle* h, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out);

/* The 7x7 sigma-2 Gaussian of `level` (src/ORBextractor.cc:1093-1094), w x h pixels without border: the image
 * computeOrbDescriptor samples.  For parity tests of the blur stage itself. */
int orbx_get_blur_level(orbx_handle* h, int frame, int level, uint8_t* dst, size_t dst_stride);

/* Stage outputs for parity tests: FAST candidates of (frame, level) as (x,y,score) int
 * triples in border-relative coordinates (vToDistributeKeys, src/ORBextractor.cc:789-837);
 * order is unspecified (the reference's order is a function of (x,y)). */
int orbx_get_candidates(orbx_handle* h, int frame, int level, int32_t* xys, int cap, int* n_out);

/* Elapsed device milliseconds of the last orbx_extract_batch (CUDA events on the handle's
 * stream, all kernels of the batch), and the number of kernel launches it issued. */
int orbx_last_batch_ms(orbx_handle* h, float* ms, int* launches);

/* The handle's CUDA stream (a cudaStream_t), so a caller can order its own work / events
 * against the extractor's kernels. */
int orbx_get_stream(orbx_handle* h, void** cuda_stream);

/* Per-stage device timing.  When enabled, every batch records CUDA events between its stages on
 * the handle's stream; orbx_last_stage_ms then returns the milliseconds of
 * [0] pyramid (resize) [1] FAST cells [2] quadtree select [3] Gaussian blur [4] orientation+rBRIEF. */
enum { ORBX_N_STAGES = 5 };
int orbx_set_stage_timing(orbx_handle* h, int enabled);
int orbx_last_stage_ms(orbx_handle* h, float* ms /* ORBX_N_STAGES */);

/* ------------------------------------------------------------------ matcher ------------ */

enum { ORBM_TH_HIGH = 95, ORBM_TH_LOW = 45, ORBM_HISTO_LENGTH = 30 }; /* ORBmatcher.cc:37-40 */

/* static int ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&) -- reference:
 * src/ORBmatcher.cc:1848-1864.  Single pair on the host (a kernel launch for 32 bytes would be
 * absurd); bulk distances go through the device entry points below. */
int orbm_hamming(const uint8_t* a /* 32 B */, const uint8_t* b /* 32 B */);

/* Row-wise distances of n pairs on the device: out[i] = Hamming(a[i], b[i]).  Host buffers. */
int orbm_hamming_pairs(int device, const uint8_t* a, const uint8_t* b, int n, int32_t* out);

/* Result of the top-2 scan of src/ORBmatcher.cc:417-426 for one query: best1 = smallest
 * distance (lowest row index on ties), best2 = second smallest as a multiset, idx = global row
 * of best1 (-1 when the database is empty; then best1 = best2 = 256). */
typedef struct {
  int32_t best1;
  int32_t best2;
  int64_t idx;
} orbm_top2;

/* A shard of the relocalization descriptor database resident on one GPU:
 * rows are 32-byte descriptors (cv::Mat N x 32 CV_8U, row-major).  row_base is the global
 * index of the shard's first row; shards are contiguous ranges in increasing global order. */
typedef struct orbm_db orbm_db;
int orbm_db_create(int device, int64_t row_base, orbm_db** out);
int orbm_db_append(orbm_db* db, const uint8_t* rows, int64_t n, int memory /* 0 host, 1 device */);
int orbm_db_size(const orbm_db* db, int64_t* n);
int orbm_db_destroy(orbm_db* db);

/* Brute-force top-2 of nq queries against the shard.  memory = 0: q and out are host buffers
 * (copies inside the call); memory = 1: both are device buffers on db's device and the call is
 * asynchronous on the db's stream (orbm_db_sync to wait). */
int orbm_top2_search(orbm_db* db, const uint8_t* q, int nq, orbm_top2* out, int memory);

/* The same search with no allocation anywhere in the call, for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  It has no counterpart in the reference.
 * rows (n_rows x 32 bytes, 16-byte aligned), q (nq x 32
 * bytes) and out (nq results) are DEVICE buffers on `device` that the caller owns; row_base is added to the
 * returned indices; the two kernels are launched on cuda_stream (NULL = the default stream) and the call
 * returns without waiting.  The per-chunk partial results go to a static device array of 304 x 1024 8-byte
 * entries (2 432 KiB, held by every GPU that loads the library until exit): a call needs
 * ceil(nq / 1024) * 1024 x (row chunks) entries, which on a 148-SM GPU admits any nq <= 311 296 against up to
 * 2^20 x floor(296 / ceil(nq / 1024)) rows; larger calls are rejected with ORB_E_INVALID and a message that
 * states both numbers.  Every call on a device shares that array (it keeps the last call's contents; nothing
 * in it has to start at zero, each entry that is read is written earlier in the same call), so two calls that
 * overlap on the GPU -- different streams, or different threads -- race: serialise them.  Results are those
 * of orbm_top2_search on the same rows and queries. */
int orbm_top2_search_noalloc(int device, const uint8_t* rows, int64_t n_rows, int64_t row_base, const uint8_t* q,
                             int nq, orbm_top2* out, void* cuda_stream);
int orbm_db_sync(orbm_db* db);
int orbm_db_last_ms(orbm_db* db, float* ms, int* launches);
int orbm_db_get_stream(orbm_db* db, void** cuda_stream);

/* Exact merge of per-shard results (shard-major: shards[s*nq + q], shards ordered by row_base)
 * into the result of one sequential scan, then the acceptance rule of ORBmatcher.cc:429-431:
 * accept[q] = best1 <= th && (float)best1 < ratio*(float)best2.  Host arrays (n_shards*nq*16 B). */
int orbm_merge_top2(const orbm_top2* shards, int n_shards, int nq, float ratio, int th,
                    orbm_top2* out, uint8_t* accept /* may be NULL */);

/* The sharded search as ONE call (BASELINE.json config 4): every member scans its shard, one ncclAllGather of
 * nq x 16 B per rank exchanges the per-shard results, a device kernel merges them (orbm_merge_top2's rule: the lowest
 * shard wins ties, i.e. the lowest global row, as ORBmatcher.cc:417-426 scanning rows in order would) and applies
 * ORBmatcher.cc:429-431, one transfer returns nq x 17 B -- all on the shard's stream, no host synchronisation in between.
 * Ranks must be ordered by row_base.  NCCL is bound at run time (dlopen libnccl.so.2); ORB_E_STATE if it is absent.
 *   multi-process (one rank per GPU, e.g. under torchrun or MPI): rank 0 calls orbm_group_unique_id and hands the 128
 *     bytes to the others by whatever means the application has; every rank calls orbm_group_create on its own shard and
 *     later orbm_group_search with the same queries; every rank receives the merged answer;
 *   single process, several GPUs: orbm_group_create_local over one shard per device (ncclCommInitAll). */
typedef struct orbm_group orbm_group;
int orbm_group_unique_id(uint8_t* id128 /* 128 bytes out */);
int orbm_group_create(orbm_db* shard, int rank, int n_ranks, const uint8_t* id128, orbm_group** out);
int orbm_group_create_local(orbm_db* const* shards, int n_shards, orbm_group** out);
int orbm_group_search(orbm_group* g, const uint8_t* q /* host, nq x 32 */, int nq, float ratio, int th,
                      orbm_top2* out /* host, nq */, uint8_t* accept /* host, nq; may be NULL */);
/* device milliseconds of the last search on the first local shard's stream: its scan, and everything up to the answer's copy */
int orbm_group_last_ms(orbm_group* g, float* scan_ms, float* total_ms);
int orbm_group_destroy(orbm_group* g);

/* A DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) flattened to CSR: node ids
 * ascending, off[n_nodes+1], feature indices in insertion order. */
typedef struct {
  const uint32_t* node;
  const int32_t* off;
  const uint32_t* idx;
  int32_t n_nodes;
} orbm_featvec;

/* int ORBmatcher::SearchByBoW(KeyFrame*, Frame&, vector<MapPoint*>&) -- src/ORBmatcher.cc:360-489
 * (kf_kf = 0) and SearchByBoW(KeyFrame*, KeyFrame*, vector<MapPoint*>&) -- :723-856 (kf_kf = 1).
 * valid1[i] / valid2[i]: the feature has a non-bad MapPoint (valid2 may be NULL for kf_kf = 0).
 * out has n2 entries for kf_kf = 0 (index on side 1 matched to frame feature, or -1) and n1
 * entries for kf_kf = 1 (index on side 2, or -1).  *n_matches = return value of the reference.
 * All pointers are host memory. */
int orbm_search_bow(int device, const uint8_t* desc1, int n1, const uint8_t* desc2, int n2,
                    const orbm_featvec* fv1, const orbm_featvec* fv2, const uint8_t* valid1,
                    const uint8_t* valid2, const float* angle1, const float* angle2, float nnratio,
                    int check_orientation, int kf_kf, int32_t* out, int* n_matches);

/* Relocalization and loop detection call SearchByBoW once per candidate keyframe against the same Frame / KeyFrame
 * (src/Tracking.cc:1768-1795, src/LoopClosing.cc:265).  The batched form takes the K candidates at once: the shared
 * side 2 is uploaded once, the K searches run as one launch (grid.y = K) with their own claim masks, and the K result
 * vectors come back in one transfer.  Results equal K orbm_search_bow calls.  side1[p] = candidate p (descriptors,
 * FeatureVector, MapPoint validity, keypoint angles); side2 = the Frame (valid may be NULL for kf_kf = 0);
 * out[p] has side2->n entries for kf_kf = 0 and side1[p].n entries for kf_kf = 1; n_matches[p] = the reference's
 * return value for candidate p.  All pointers are host memory. */
typedef struct orbm_bow_side {
  const uint8_t* desc;      /* n x 32 */
  int32_t n;
  const orbm_featvec* fv;
  const uint8_t* valid;     /* n flags: feature has a non-bad MapPoint */
  const float* angle;       /* n keypoint angles (degrees); may be NULL when check_orientation = 0 */
} orbm_bow_side;
int orbm_search_bow_batch(int device, const orbm_bow_side* side1, int k, const orbm_bow_side* side2, float nnratio,
                          int check_orientation, int kf_kf, int32_t* const* out, int* n_matches);

/* Descriptor part of Frame::ComputeStereoMatches -- src/Frame.cc:527-610: per left keypoint the
 * best right keypoint on its row band (octave within +-1, uR in [uL-max_d, uL-min_d]) with
 * initial bound TH_HIGH.  best_dist = -1 for keypoints the reference skips before the scan.
 * Host memory. */
int orbm_stereo_search(int device, const orbx_keypoint* left, const uint8_t* ldesc, int nl,
                       const orbx_keypoint* right, const uint8_t* rdesc, int nr,
                       const float* scale_factors, int n_levels, int n_rows, float min_d, float max_d,
                       int32_t* best_idx, int32_t* best_dist);

/* ---- Windowed projection searches -------------------------------------------
 * One primitive behind the five ORBmatcher::SearchByProjection overloads (src/ORBmatcher.cc:46-132,
 * :134-220, :491-604, :1529-1671, :1673-1800), both Fuse overloads (:1026-1176, :1178-1301) and SearchBySim3
 * (:1303-1527): for every projected query, gather the keypoints of
 * a frame inside a square window through the frame's cell grid (Frame::GetFeaturesInArea,
 * src/Frame.cc:342-395; KeyFrame::GetFeaturesInArea, src/KeyFrame.cc:613-652), keep the best and
 * second-best Hamming distance in the reference's scan order, apply its acceptance rule, and
 * -- as the reference does by writing mvpMapPoints inside the loop -- let an accepted match hide
 * its keypoint from every LATER query.  The caller projects (pose, camera and MapPoint state stay
 * on its side) and passes flat queries in the reference's loop order. */
typedef struct orbm_grid_frame { /* the searched Frame / KeyFrame; all host memory */
  const float* x;          /* n: mvKeysUn[i].pt.x */
  const float* y;          /* n: mvKeysUn[i].pt.y */
  const int32_t* octave;   /* n: mvKeysUn[i].octave */
  const float* u_right;    /* n: mvuRight[i], or NULL (monocular: treated as -1) */
  const uint8_t* desc;     /* n x 32: mDescriptors */
  int32_t n;
  float min_x, min_y;           /* mnMinX, mnMinY */
  float width_inv, height_inv;  /* mfGridElementWidthInv, mfGridElementHeightInv */
  int32_t cols, rows;           /* FRAME_GRID_COLS x FRAME_GRID_ROWS (64 x 48) or mnGridCols x mnGridRows */
  const int32_t* cell_off;      /* cols*rows + 1: cell (ix, iy) is entries [off[ix*rows+iy], off[ix*rows+iy+1]) */
  const uint32_t* cell_idx;     /* mGrid[ix][iy] flattened; a keypoint may appear in at most one cell */
} orbm_grid_frame;

typedef struct orbm_window_query {
  float x, y, r;        /* GetFeaturesInArea(x, y, r, min_level, max_level) */
  float u_right;        /* predicted right-image coordinate of the query */
  float er_max;         /* a keypoint with mvuRight > 0 is skipped when fabs(u_right - mvuRight) > er_max;
                           negative: no stereo check */
  int32_t min_level, max_level; /* as GetFeaturesInArea: filter iff min_level > 0 || max_level >= 0 */
  int32_t blocks;       /* nonzero: once accepted, the matched keypoint is hidden from later queries
                           (the reference's `mvpMapPoints[idx] && Observations() > 0` test) */
} orbm_window_query;

typedef struct orbm_window_result {
  int32_t idx;          /* matched keypoint (index into the frame), or -1 when not accepted */
  int32_t best_idx;     /* best keypoint of the scan even if rejected, -1 if the window was empty */
  int32_t best_dist, best_dist2;  /* 256 when absent */
  int32_t best_level, best_level2; /* -1 when absent */
} orbm_window_result;

typedef struct orbm_window_params {
  int32_t th;      /* accept iff best_dist <= th */
  float ratio;     /* > 0: also reject when best_level == best_level2 && best_dist > ratio * best_dist2
                      (src/ORBmatcher.cc:119-123); <= 0: distance test only (:1622, :1759, :595) */
  int32_t claim;   /* nonzero: an accepted query with `blocks` hides its keypoint from later queries;
                      0: matches never hide keypoints (:180-181, :1148, :1276: separate outputs) */
  /* Reprojection gate of ORBmatcher::Fuse (src/ORBmatcher.cc:1118-1146), off when inv_level_sigma2
   * is NULL: a candidate with mvuRight >= 0 is skipped when
   * ((x-kx)^2 + (y-ky)^2 + (u_right-kr)^2) * inv_level_sigma2[octave] > chi2_stereo, any other when
   * ((x-kx)^2 + (y-ky)^2) * inv_level_sigma2[octave] > chi2_mono (float products, double compare). */
  const float* inv_level_sigma2; /* n_levels: mvInvLevelSigma2 */
  int32_t n_levels;
  double chi2_mono, chi2_stereo; /* 5.99, 7.8 */
} orbm_window_params;

/* claimed: n flags (keypoints hidden from every query, e.g. mvpMapPoints already set) or NULL.
 * out has nq entries; *n_accepted counts idx >= 0; *rounds (may be NULL) reports how many parallel
 * passes the order-exact resolution took. */
int orbm_window_search(int device, const orbm_grid_frame* frame, const orbm_window_query* queries,
                       const uint8_t* query_desc, int nq, const uint8_t* claimed,
                       const orbm_window_params* params, orbm_window_result* out, int* n_accepted,
                       int* rounds);

/* Candidate lists of windowed queries, for callers whose acceptance rule needs more than the best
 * two (ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:606-721, re-assigns a keypoint to a
 * later query with a smaller distance): for every query the keypoints GetFeaturesInArea returns, in
 * its order, each with its Hamming distance to the query descriptor.  Query q owns entries
 * [offset[q], offset[q] + count[q]) of cand_idx / cand_dist (ranges are disjoint, not ordered by q).
 * cap: capacity of cand_idx / cand_dist; ORB_E_CAPACITY with *n_total = entries needed when too
 * small.  u_right / er_max / blocks of the queries are ignored. */
int orbm_window_candidates(int device, const orbm_grid_frame* frame, const orbm_window_query* queries,
                           const uint8_t* query_desc, int nq, int32_t* offset, int32_t* count,
                           int32_t* cand_idx, int32_t* cand_dist, int cap, int* n_total);

/* ORBmatcher::SearchForTriangulation -- src/ORBmatcher.cc:858-1024 (LocalMapping::CreateNewMapPoints):
 * keypoints of two keyframes that have no MapPoint yet, matched inside common vocabulary nodes
 * under the epipolar constraint of F12.  For every usable side-1 keypoint of a common node the
 * side-2 keypoint with the smallest distance <= TH_LOW (the LAST one on ties, as `dist > bestDist`
 * skips only strictly worse candidates) among those that pass the epipole test (:921-927, both
 * keypoints monocular) and CheckDistEpipolarLine (:341-358); then the rotation-histogram filter.
 * The reference never sets vbMatched2, so side-1 keypoints are independent.  Host memory. */
typedef struct orbm_tri_keyframe {
  const uint8_t* desc;     /* n x 32 */
  const float* x;          /* n: mvKeysUn[i].pt.x */
  const float* y;
  const int32_t* octave;   /* n */
  const float* angle;      /* n (read when check_orientation) */
  const uint8_t* usable;   /* n: GetMapPoint(i) == NULL && (!bOnlyStereo || mvuRight[i] >= 0) */
  const uint8_t* stereo;   /* n: mvuRight[i] >= 0 */
  int32_t n;
  const orbm_featvec* fv;  /* mFeatVec */
} orbm_tri_keyframe;
/* F12: 9 floats row-major; ex, ey: epipole in image 2 (:868-872); scale_factors2 / level_sigma2_2:
 * pKF2->mvScaleFactors / mvLevelSigma2.  matches12 has kf1->n entries (index in kf2 or -1);
 * *n_matches = the reference's return value. */
int orbm_search_triangulation(int device, const orbm_tri_keyframe* kf1, const orbm_tri_keyframe* kf2,
                              const float* F12, float ex, float ey, const float* scale_factors2,
                              const float* level_sigma2_2, int n_levels, int check_orientation,
                              int32_t* matches12, int* n_matches);

/* Frame::ComputeStereoMatches, whole -- src/Frame.cc:527-700: the row-band descriptor search
 * (orbm_stereo_search), the 11x11 sliding-window refinement on the pyramid level of the left
 * keypoint (:613-653), the parabola fit (:655-684) and the median-based rejection (:687-700).
 * The pyramids are read where the two extractors left them on the device (frame_left / frame_right
 * of their last batches; the two may be the same handle holding both images as one batch of two),
 * so a stereo frame needs no mvImagePyramid download.  Keypoints and descriptors are the outputs
 * of those extractions (host memory).  mb, mbf: Frame::mb, Frame::mbf.  u_right / depth: mvuRight /
 * mvDepth, -1 where unmatched. */
int orbx_stereo_matches(orbx_handle* left, int frame_left, orbx_handle* right, int frame_right,
                        const orbx_keypoint* keys_left, const uint8_t* desc_left, int n_left,
                        const orbx_keypoint* keys_right, const uint8_t* desc_right, int n_right,
                        float mb, float mbf, float* u_right, float* depth);
/* The same for a batch of pairs held by ONE handle: pair p = frames (2p, 2p+1) of its last batch (left, right).
 * Keypoints, descriptors and counts are read where the extraction left them on the device, three launches cover
 * every pair (row search, refinement, median rejection on the device), one transfer brings mvuRight / mvDepth
 * back: u_right / depth hold cap_per_frame floats per pair (>= orbx_frame_capacity), row i = left keypoint i in
 * output order, -1 where unmatched and past the left image's keypoint count. */
int orbx_stereo_matches_batch(orbx_handle* h, int n_pairs, float mb, float mbf, float* u_right, float* depth,
                              int cap_per_frame);

/* ---- Rest of Frame construction ---------------------------------------------
 * What the Frame constructors do with the extractor's output before tracking sees the frame
 * (src/Frame.cc:85-120, :148-178, :207-240): UndistortKeyPoints, ComputeStereoFromRGBD (RGB-D only) and
 * AssignFeaturesToGrid, plus -- once per camera -- ComputeImageBounds. */
typedef struct orbx_camera {
  float fx, fy, cx, cy;      /* mK (CV_32F 3x3): K(0,0), K(1,1), K(0,2), K(1,2) */
  float k1, k2, p1, p2, k3;  /* mDistCoef (CV_32F): k3 = 0 when the settings file has none (src/Tracking.cc:66-77).
                                k1 == 0 means "no distortion": keypoints are copied (src/Frame.cc:421-425) */
  float bf;                  /* mbf (only read with a depth image) */
} orbx_camera;
typedef struct orbx_frame_bounds {
  float min_x, min_y;           /* Frame::mnMinX, mnMinY */
  float width_inv, height_inv;  /* mfGridElementWidthInv, mfGridElementHeightInv (src/Frame.cc:106-107) */
  int32_t cols, rows;           /* FRAME_GRID_COLS x FRAME_GRID_ROWS (64 x 48, include/Frame.h:39-40) */
} orbx_frame_bounds;
/* cv::undistortPoints(src, dst, K, distCoef, Mat(), K) on n (x, y) float pairs, as
 * Frame::UndistortKeyPoints (src/Frame.cc:419-449) and Frame::ComputeImageBounds (:451-485, the four
 * corners) call it: five fixed-point iterations in double, result rounded to float.  With k1 == 0
 * the points are copied.  Host memory; xy_out may alias xy. */
int orbx_undistort_points(int device, const float* xy /* n x 2 */, int n, const orbx_camera* cam,
                          float* xy_out /* n x 2 */);
/* One launch for the rest of a Frame: xy_un[i] = mvKeysUn[i].pt (src/Frame.cc:419-449); the cell lists
 * of AssignFeaturesToGrid (:245-260 with PosInGrid, :397-408) in the CSR layout orbm_grid_frame takes
 * (cell (ix, iy) = entries [cell_off[ix*rows+iy], cell_off[ix*rows+iy+1]) of cell_idx, keypoint order
 * inside a cell; *n_in_grid = keypoints that fall inside the grid); and, when depth != NULL,
 * ComputeStereoFromRGBD (:703-727): depth_out[i] = mvDepth[i] = imDepth.at<float>(y, x) when positive,
 * u_right[i] = mvuRight[i] = xy_un[i].x - bf / depth, else -1 for both.  depth is a CV_32F image
 * (depth_stride in bytes) in host (depth_memory 0) or device (1) memory; everything else is host memory.
 * cell_idx needs n entries, cell_off cols*rows + 1. */
int orbx_frame_finish(int device, const orbx_keypoint* keys, int n, const orbx_camera* cam,
                      const orbx_frame_bounds* grid, const float* depth, int depth_w, int depth_h,
                      size_t depth_stride, int depth_memory, float* xy_un /* n x 2 */, float* u_right,
                      float* depth_out, int32_t* cell_off, uint32_t* cell_idx, int* n_in_grid);

/* ---- Vocabulary tree --------------------------------------------------------
 * DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> as ORB-SLAM2 uses it (include/ORBVocabulary.h):
 * the k-ary tree stays resident on the device and descriptors are pushed down it in bulk
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1217-1259, called per feature from :1127-1196 by
 * Frame::ComputeBoW, src/Frame.cc:410-417, and KeyFrame::ComputeBoW, src/KeyFrame.cc:70-79). */
typedef struct orbm_vocab orbm_vocab;
/* Nodes 0..n_nodes-1, node 0 the root; parent[i] < i; children keep ascending node order (the order
 * loadFromTextFile appends them, :1377-1391).  is_word[i]: node i carries a word; word ids count
 * those nodes in node order (:1407-1414).  desc: n_nodes x 32; weight: per node (Node::weight). */
int orbm_vocab_create(int device, int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_word,
                      const uint8_t* desc, const double* weight, orbm_vocab** out);
/* Reads the text format of TemplatedVocabulary::loadFromTextFile (:1337-1421; "k L scoring weighting",
 * then one "parent leaf d0..d31 weight" line per node).  Unlike the reference, an empty last line does
 * not become a node. */
int orbm_vocab_load_text(int device, const char* path, orbm_vocab** out);
int orbm_vocab_info(const orbm_vocab* v, int* k, int* L, int* n_nodes, int* n_words);
int orbm_vocab_destroy(orbm_vocab* v);
/* transform(feature, word_id, weight, &nid, levelsup) for n descriptors (host memory): word id and
 * weight of the leaf each descriptor reaches (first child with the smallest distance at every
 * level), and the node it passed at level L - levelsup (0 when that is the root or above). */
int orbm_vocab_transform(orbm_vocab* v, const uint8_t* desc, int n, int levelsup, int32_t* word_id,
                         int32_t* node_id, double* weight);

/* MapPoint::ComputeDistinctiveDescriptors -- src/MapPoint.cc:305-370, batched over map points
 * (LocalMapping::ProcessNewKeyFrame / CreateNewMapPoints / SearchInNeighbors refresh hundreds per
 * keyframe): point p owns descriptors [offset[p], offset[p+1]) of desc (the rows of its good
 * observations, in the reference's std::map iteration order).  best[p] = index inside that range
 * of the descriptor with the least median distance to the others -- median = element
 * (int)(0.5*(N-1)) of the sorted row of the N x N distance matrix, diagonal included, first row on
 * ties -- or -1 for an empty range.  At most 1024 descriptors per point.  Host memory. */
int orbm_distinctive_descriptors(int device, const uint8_t* desc, const int32_t* offset, int n_points,
                                 int32_t* best);

/* Descriptor database interchange: a flat little-endian file, 16-byte header
 * {"ORBD", u32 version = 1, u64 n_rows} followed by n_rows x 32 descriptor bytes; a shard's rows can
 * be appended straight from / saved straight to it.  orbm_db_load appends rows [first, first+count)
 * of the file (count < 0: to the end) to db. */
int orbm_db_save(orbm_db* db, const char* path);
int orbm_db_load(orbm_db* db, const char* path, int64_t first, int64_t count);

/* Measures the integer POPC issue rate of the device (dependent-free chains on every SM), the
 * denominator of the matching roofline.  Returns 32-bit POPC per second. */
int orbm_popc_peak(int device, double* popc_per_s);

#ifdef __cplusplus
}
#endif
#endif /* ORBFRONT_H_ */
