// Host side of the extractor: handle, geometry/tables (the reference constructor and the
// per-level constants of ComputePyramid / ComputeKeyPointsOctTree), device buffers, stream
// orchestration and the orbx_* C ABI (include/orbfront.h).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "extract.cuh"

using namespace orbfront;

namespace {

inline int cv_round(double v) { return (int)lrint(v); }
inline int cv_round(float v) { return (int)lrintf(v); }
inline int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

template <typename T>
struct Dev {  // grow-only device allocation
  T* p = nullptr;
  size_t cap = 0;
  int reserve(size_t n) {
    if (n <= cap) return ORB_OK;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc(&p, n * sizeof(T));
    if (e != cudaSuccess) return set_error(ORB_E_CUDA, "cudaMalloc(%zu bytes): %s", n * sizeof(T), cudaGetErrorString(e));
    cap = n;
    return ORB_OK;
  }
  ~Dev() { if (p) cudaFree(p); }
};

}  // namespace

struct orbx_handle {
  orbx_config cfg;
  // reference constructor state (ORBextractor.cc:410-446)
  std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
  std::vector<int> quota;
  // geometry for the current frame size
  int w = 0, h = 0;
  ExtractParams P;
  size_t l0_pitch = 0, l0_frame = 0;
  int slots = 0;
  // device state
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // per compute lane: the stream the blur runs on beside the quadtree, and its fork / join events
  std::vector<cudaStream_t> s_blur;
  std::vector<cudaEvent_t> ev_fork, ev_join;
  // copy streams + per-chunk events of orbx_extract_batch_host (upload | extract | download pipeline)
  cudaStream_t s_up = nullptr, s_down = nullptr, s_alt = nullptr;  // s_alt: second compute stream
  std::vector<cudaStream_t> s_more;  // further compute streams (ORBX_COMPUTE_STREAMS > 2)
  std::vector<cudaEvent_t> ev_up, ev_done;
  int* pinned_counts = nullptr;  // page-locked staging for per-frame counts (a pageable destination would block the host per chunk)
  int pinned_counts_cap = 0;
  cudaEvent_t stage_ev[ORBX_N_STAGES + 1] = {};
  bool stage_timing = false, stage_valid = false;
  Dev<uint8_t> l0, pyr, blur, qt_b, out_desc;
  Dev<unsigned long long> cand, qt_tmp;
  Dev<uint32_t> kept, qt_seq;
  Dev<int> cand_count, kept_count, out_count, pyr_bar;
  int pyr_max_grid = -1;
  Dev<int4> rtab;
  Dev<uint32_t> qcode;
  Dev<int> blur_tiles;
  Dev<orbx_keypoint> out_kp;
  Dev<uint8_t> border_tmp;
  // single-frame drop-in path: all bordered levels staged to page-locked memory on a side stream
  bool pyr_staging = false, pyr_pending = false;
  bool read_pending = false;  // orbx_batch_read_begin queued copies on s_down that orbx_batch_read_end has not waited for
  cudaEvent_t ev_read = nullptr;
  Dev<uint8_t> pyr_stage_dev;
  uint8_t* pyr_stage_host = nullptr;
  size_t pyr_stage_cap = 0, pyr_stage_off[kMaxLevels] = {};
  cudaStream_t s_pyr = nullptr;
  cudaEvent_t ev_pyr_ready = nullptr, ev_pyr_done = nullptr;
  uint8_t* out_stage_host = nullptr;  // count + keypoints + descriptors of one frame (orbx_read_frame0)
  size_t out_stage_cap = 0;
  int n_frames = 0;       // frames of the last batch (0: nothing extracted yet)
  int last_launches = 0;
  std::vector<int> host_counts;
  bool counts_valid = false;
};

namespace {

// Constructor arithmetic of the reference, expression types included (scaleFactor is stored as
// double, ORBextractor.h:99).
void init_tables(orbx_handle* H) {
  const orbx_config& c = H->cfg;
  const int nl = c.nlevels;
  const double sf = (double)c.scale_factor;
  H->scale.assign(nl, 1.f); H->sigma2.assign(nl, 1.f); H->inv_scale.assign(nl, 1.f); H->inv_sigma2.assign(nl, 1.f);
  for (int i = 1; i < nl; ++i) {
    H->scale[i] = (float)(H->scale[i - 1] * sf);
    H->sigma2[i] = H->scale[i] * H->scale[i];
  }
  for (int i = 0; i < nl; ++i) { H->inv_scale[i] = 1.0f / H->scale[i]; H->inv_sigma2[i] = 1.0f / H->sigma2[i]; }
  H->quota.assign(nl, 0);
  const float factor = (float)(1.0f / sf);
  float want = c.nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nl));
  int sum = 0;
  for (int l = 0; l < nl - 1; ++l) {
    H->quota[l] = cv_round(want);
    sum += H->quota[l];
    want *= factor;
  }
  H->quota[nl - 1] = std::max(c.nfeatures - sum, 0);
}

// cv::resize(INTER_LINEAR) taps of one axis (11-bit weights); x axis clamps the fraction with
// the index, y axis keeps the fraction and clamps rows (see oracle/cvprims.hpp, pinned to cv2).
void axis_taps(int src_n, int dst_n, bool clamp_frac, int4* out) {
  const double scale = (double)src_n / dst_n;
  for (int d = 0; d < dst_n; ++d) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= s;
    if (clamp_frac) {
      if (s < 0) { f = 0; s = 0; }
      if (s >= src_n - 1) { f = 0; s = src_n - 1; }
    }
    const int c0 = (short)cv_round((1.f - f) * 2048.f), c1 = (short)cv_round(f * 2048.f);
    const int s0 = std::min(std::max(s, 0), src_n - 1), s1 = std::min(std::max(s + 1, 0), src_n - 1);
    out[d] = make_int4(s0, c0, c1, s1);
  }
}

int configure(orbx_handle* H, int w, int h, int n_frames, bool own_l0) {
  const orbx_config& c = H->cfg;
  const int nl = c.nlevels;
  ExtractParams& P = H->P;
  const bool geom_changed = (w != H->w || h != H->h);
  if (geom_changed) {
    // every early return below leaves P half-built: forget the old size first, so that the next call -- whatever its
    // size -- rebuilds the geometry instead of running the kernels on it
    H->w = H->h = 0;
    H->n_frames = 0;
    memset(&P, 0, sizeof(P));
    P.nlevels = nl;
    P.ini_th = c.ini_th_fast;
    P.min_th = c.min_th_fast;
    size_t off = 0;
    int cell_base = 0, cand_off = 0, kept_off = 0, tab = 0, tiles = 0, tiles_few = 0;
    std::vector<int> blur_tiles(2 * (nl + 1), 0);  // first warp tile of each level: 112-row tiles, then 14-row tiles (pyramid.cu kBlurTH / kBlurTHFew)
    for (int l = 0; l < nl; ++l) {
      LevelGeom& G = P.lv[l];
      G.w = cv_round((float)w * H->inv_scale[l]);
      G.h = cv_round((float)h * H->inv_scale[l]);
      const int minB = kEdge - 3, maxBX = G.w - kEdge + 3, maxBY = G.h - kEdge + 3;
      const float width = (float)(maxBX - minB), height = (float)(maxBY - minB);
      if (width < 30.f || height < 30.f)
        return set_error(ORB_E_INVALID, "level %d is %dx%d: too small for the 30-px FAST cell grid (reference divides by zero)", l, G.w, G.h);
      G.pitch = (int)align_up<size_t>((size_t)G.w + 4, 64);
      G.off = off;
      off += align_up<size_t>((size_t)G.pitch * G.h, 256);
      G.nCols = (int)(width / 30.f);
      G.nRows = (int)(height / 30.f);
      G.wCell = (int)ceilf(width / G.nCols);
      G.hCell = (int)ceilf(height / G.nRows);
      G.cell_base = cell_base;
      G.rcp_nCols = (uint32_t)((1ull << 32) / (unsigned)G.nCols + 1ull);
      cell_base += G.nCols * G.nRows;
      G.boxW = maxBX - minB;
      G.boxH = maxBY - minB;
      G.nIni = (int)roundf((float)G.boxW / (float)G.boxH);
      if (G.nIni < 1 || G.nIni > (1 << kRootBits))
        return set_error(ORB_E_INVALID, "level %d aspect ratio gives %d quadtree roots (supported 1..%d)", l, G.nIni, 1 << kRootBits);
      if (G.boxW > kMaxBox || G.boxH > kMaxBox)
        return set_error(ORB_E_INVALID, "frame %dx%d exceeds the supported %d-px detection box", w, h, kMaxBox);
      G.hX = (float)G.boxW / G.nIni;
      G.quota = H->quota[l];
      int span = std::max((int)ceilf(G.hX) + 1, G.boxH), lg = 0;
      while ((1 << lg) < span) ++lg;
      G.dmax = std::min(lg + 2, kMaxDepth);
      const int vw = G.w - 2 * kEdge, vh = G.h - 2 * kEdge;
      G.cand_cap = (int)align_up<size_t>((size_t)((vw + 1) / 2 + G.nCols) * ((vh + 1) / 2 + G.nRows), 4);  // NMS survivors: <= one per 2x2 block of a cell
      G.cand_off = cand_off;
      cand_off += G.cand_cap;
      {  // sort-free quadtree (quadtree.cu): finest grid of about 6 x quota cells; the loop must be able to end one depth above it
        int rb = 0;
        while ((1 << rb) < G.nIni) ++rb;
        const long long R = 1ll << rb, N = std::max(G.quota, 1);
        int d0 = 2;
        while (d0 < G.dmax && (R << (2 * d0)) < 6 * N) ++d0;
        const long long cells = R << (2 * d0), total = cells * 4 / 3 + 8;
        // scratch of the level: sequence numbers (u32 per pyramid cell) then, for large grids, the u16 counts; best keys (u64 per finest cell)
        const bool fits = d0 >= 2 && d0 <= G.dmax && (R << (2 * (d0 - 1))) * 2 >= 3 * N && total * 6 + 64 <= (long long)G.cand_cap * 4 && cells <= G.cand_cap &&
                          cells <= (1 << 18);
        G.qt_d0 = fits ? d0 : 0;
        G.qt_counts_global = fits && total > 12288 ? 1 : 0;  // up to 24 KB of shared-memory counters
      }
      G.kept_cap = std::max(G.quota, 0) + 4 * G.nIni + 3;
      G.kept_off = kept_off;
      kept_off += G.kept_cap;
      G.scale = H->scale[l];
      G.patch = (float)(int)(31 * H->scale[l]);  // const int scaledPatchSize = PATCH_SIZE*mvScaleFactor[level]
      G.xtab = tab; tab += G.w;
      G.ytab = tab; tab += G.h;
      blur_tiles[l] = tiles;
      tiles += div_up(div_up(G.w, 32) * div_up(G.h, 112), 4);  // pyramid.cu blur_warp_tiles(w, h, kBlurTH)
      blur_tiles[nl + 1 + l] = tiles_few;
      tiles_few += div_up(div_up(G.w, 32) * div_up(G.h, 14), 4);
    }
    blur_tiles[nl] = tiles;
    blur_tiles[2 * nl + 1] = tiles_few;
    P.total_cells = cell_base;
    P.pyr_frame = P.blur_frame = off;
    P.cand_frame = (size_t)cand_off;
    P.kept_frame = (size_t)kept_off;
    P.out_cap = kept_off;
    H->slots = kept_off;
    // resize taps (levels >= 1 read level l-1)
    std::vector<int4> taps((size_t)tab + 2, make_int4(0, 0, 0, 0));  // + 2: the resize row loop reads its table one row ahead
    for (int l = 1; l < nl; ++l) {
      axis_taps(P.lv[l - 1].w, P.lv[l].w, true, &taps[P.lv[l].xtab]);
      axis_taps(P.lv[l - 1].h, P.lv[l].h, false, &taps[P.lv[l].ytab]);
      // word-wise kernel precondition: window of a 4-px group (first left tap .. last right tap) <= 8 bytes,
      // and consecutive destination rows advance the source row by at most 2
      bool ok = true;
      const int4* xt = &taps[P.lv[l].xtab];
      for (int x0 = 0; x0 < P.lv[l].w && ok; x0 += 4) {
        const int xe = std::min(x0 + 3, P.lv[l].w - 1);
        for (int x = x0; x <= xe; ++x) ok = ok && xt[x].x >= xt[x0].x && xt[x].w >= xt[x0].x && xt[x].w - xt[x0].x < 8 && xt[x].x - xt[x0].x < 8;
      }
      const int4* yt = &taps[P.lv[l].ytab];
      // rows: lower row = upper row + 1 (clamped at the last row), upper rows strictly increasing (the kernel walks the source rows once)
      for (int y = 0; y < P.lv[l].h && ok; ++y) ok = ok && yt[y].x >= 0 && yt[y].w == std::min(yt[y].x + 1, P.lv[l - 1].h - 1) && (y == 0 || yt[y].x > yt[y - 1].x);
      P.lv[l].resize_words = ok ? 1 : 0;
    }
    int rc;
    // quadtree path-code tables (extract.cuh): per level the digits of the x halvings (bit 2k of
    // digit k) with the root strip on top, and of the y halvings (bit 2k+1)
    std::vector<uint32_t> qcode;
    for (int l = 0; l < nl; ++l) {
      LevelGeom& G = P.lv[l];
      G.qx = (int)qcode.size();
      for (int x = 0; x <= G.boxW; ++x) {
        const int root = (int)((float)x / G.hX);  // ORBextractor.cc:569, float division
        int lo = (int)(G.hX * (float)root), hi = (int)(G.hX * (float)(root + 1));
        uint32_t code = (uint32_t)root;
        for (int k = 0; k < G.dmax; ++k) {
          const int mid = lo + ((hi - lo + 1) >> 1);
          const int b = x >= mid;
          if (b) lo = mid; else hi = mid;
          code = (code << 2) | (uint32_t)b;
        }
        qcode.push_back(code);
      }
      G.qy = (int)qcode.size();
      for (int y = 0; y <= G.boxH; ++y) {
        int lo = 0, hi = G.boxH;
        uint32_t code = 0;
        for (int k = 0; k < G.dmax; ++k) {
          const int mid = lo + ((hi - lo + 1) >> 1);
          const int b = y >= mid;
          if (b) lo = mid; else hi = mid;
          code = (code << 2) | (uint32_t)(b * 2);
        }
        qcode.push_back(code);
      }
    }
    if ((rc = H->rtab.reserve(taps.size())) || (rc = H->blur_tiles.reserve(2 * (nl + 1))) || (rc = H->qcode.reserve(qcode.size()))) return rc;
    ORB_CUDA_TRY(cudaMemcpyAsync(H->qcode.p, qcode.data(), qcode.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, H->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(H->rtab.p, taps.data(), taps.size() * sizeof(int4), cudaMemcpyHostToDevice, H->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(H->blur_tiles.p, blur_tiles.data(), blur_tiles.size() * sizeof(int), cudaMemcpyHostToDevice, H->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));  // host vectors go out of scope
    H->w = w; H->h = h;
    H->l0_pitch = align_up<size_t>((size_t)w + 4, 64);
    H->l0_frame = H->l0_pitch * (size_t)h;  // pitch is a multiple of 64, so frames stay 64-byte aligned
  }
  // buffers (grow-only)
  int rc;
  const size_t nf = (size_t)n_frames;
  if ((own_l0 && (rc = H->l0.reserve(nf * H->l0_frame + 256))) || (rc = H->pyr.reserve(nf * P.pyr_frame + 256)) ||
      (rc = H->blur.reserve(nf * P.blur_frame + 256)) || (rc = H->cand.reserve(nf * P.cand_frame)) ||
      (rc = H->qt_b.reserve(nf * (P.cand_frame + kMaxLevels) + 16)) || (rc = H->qt_seq.reserve(nf * P.cand_frame)) || (rc = H->qt_tmp.reserve(nf * P.cand_frame)) ||
      (rc = H->kept.reserve(nf * P.kept_frame)) || (rc = H->cand_count.reserve(nf * nl)) ||
      (rc = H->kept_count.reserve(nf * nl)) || (rc = H->out_count.reserve(nf)) || (rc = H->pyr_bar.reserve(1)) ||
      (rc = H->out_kp.reserve(nf * P.out_cap)) || (rc = H->out_desc.reserve(nf * P.out_cap * 32)))
    return rc;
  P.n_frames = n_frames;
  P.pyr = H->pyr.p; P.blur = H->blur.p; P.cand = H->cand.p; P.cand_count = H->cand_count.p;
  P.kept = H->kept.p; P.kept_count = H->kept_count.p; P.qt_b = H->qt_b.p; P.qt_seq = H->qt_seq.p; P.qt_tmp = H->qt_tmp.p;
  P.rtab = H->rtab.p; P.blur_tiles = H->blur_tiles.p; P.qcode = H->qcode.p;
  P.out_kp = H->out_kp.p; P.out_desc = H->out_desc.p; P.out_count = H->out_count.p;
  if (H->pyr_max_grid < 0) H->pyr_max_grid = pyramid_fused_max_grid(H->cfg.device);
  P.pyr_bar = H->pyr_bar.p; P.pyr_max_grid = H->pyr_max_grid;
  return ORB_OK;
}

}  // namespace

// ====================================================================================
extern "C" {

// max_width / max_height / max_batch of orbx_create, when given, are limits: everything was sized for them up front,
// and a larger frame or batch would reallocate on the hot path.
static int check_limits(const orbx_handle* H, int n, int w, int ht) {
  const orbx_config& c = H->cfg;
  if (c.max_width > 0 && c.max_height > 0 && (w > c.max_width || ht > c.max_height))
    return set_error(ORB_E_INVALID, "frame %dx%d exceeds the %dx%d the handle was created for", w, ht, c.max_width, c.max_height);
  if (c.max_width > 0 && c.max_height > 0 && c.max_batch > 0 && n > c.max_batch)
    return set_error(ORB_E_INVALID, "batch of %d frames exceeds max_batch %d of orbx_create", n, c.max_batch);
  return ORB_OK;
}

// The parameter block of frames [f0, f0 + nf) of the configured batch.
static ExtractParams frame_range(const ExtractParams& P, int f0, int nf) {
  ExtractParams Q = P;
  const size_t f = (size_t)f0;
  Q.n_frames = nf;
  Q.l0 += f * P.l0_frame;
  Q.pyr += f * P.pyr_frame;
  Q.blur += f * P.blur_frame;
  Q.cand += f * P.cand_frame;
  Q.cand_count += f * P.nlevels;
  Q.kept += f * P.kept_frame;
  Q.kept_count += f * P.nlevels;
  Q.qt_b += f * (P.cand_frame + kMaxLevels);
  Q.qt_seq += f * P.cand_frame;
  Q.qt_tmp += f * P.cand_frame;
  Q.out_kp += f * P.out_cap;
  Q.out_desc += f * (size_t)P.out_cap * 32;
  Q.out_count += f;
  Q.pyr_bar = nullptr;  // chunks of one handle run concurrently on two streams: no shared barrier counter
  return Q;
}

int orbx_create(const orbx_config* cfg, orbx_handle** out) {
  if (!cfg || !out) return set_error(ORB_E_INVALID, "orbx_create: null argument");
  if (cfg->nlevels < 1 || cfg->nlevels > kMaxLevels || cfg->nfeatures < 0 || !(cfg->scale_factor > 1.0f) ||
      cfg->min_th_fast < 1 || cfg->ini_th_fast < cfg->min_th_fast || cfg->ini_th_fast > 254)
    return set_error(ORB_E_INVALID,
                     "orbx_create: need 1 <= nlevels <= %d, scaleFactor > 1, 1 <= minThFAST <= iniThFAST <= 254", kMaxLevels);
  int rc = use_device(cfg->device);
  if (rc) return rc;
  orbx_handle* H = new orbx_handle();
  H->cfg = *cfg;
  init_tables(H);
  cudaError_t e = cudaStreamCreateWithFlags(&H->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreate(&H->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&H->ev1);
  if (e != cudaSuccess) {
    delete H;
    return set_error(ORB_E_CUDA, "orbx_create: %s", cudaGetErrorString(e));
  }
  if (cfg->max_width > 0 && cfg->max_height > 0) {  // size everything up front
    rc = configure(H, cfg->max_width, cfg->max_height, std::max(cfg->max_batch, 1), true);
    if (rc) { orbx_destroy(H); return rc; }
  }
  *out = H;
  return ORB_OK;
}

int orbx_destroy(orbx_handle* H) {
  if (!H) return ORB_OK;
  cudaSetDevice(H->cfg.device);
  if (H->stream) cudaStreamSynchronize(H->stream);
  if (H->ev0) cudaEventDestroy(H->ev0);
  if (H->ev1) cudaEventDestroy(H->ev1);
  if (H->s_up) { cudaStreamSynchronize(H->s_up); cudaStreamDestroy(H->s_up); }
  if (H->s_down) { cudaStreamSynchronize(H->s_down); cudaStreamDestroy(H->s_down); }
  if (H->s_alt) { cudaStreamSynchronize(H->s_alt); cudaStreamDestroy(H->s_alt); }
  for (cudaStream_t t : H->s_more) { cudaStreamSynchronize(t); cudaStreamDestroy(t); }
  for (cudaStream_t t : H->s_blur) { cudaStreamSynchronize(t); cudaStreamDestroy(t); }
  for (cudaEvent_t e : H->ev_fork) cudaEventDestroy(e);
  for (cudaEvent_t e : H->ev_join) cudaEventDestroy(e);
  if (H->pinned_counts) cudaFreeHost(H->pinned_counts);
  if (H->s_pyr) { cudaStreamSynchronize(H->s_pyr); cudaStreamDestroy(H->s_pyr); }
  if (H->ev_pyr_ready) cudaEventDestroy(H->ev_pyr_ready);
  if (H->ev_pyr_done) cudaEventDestroy(H->ev_pyr_done);
  if (H->ev_read) cudaEventDestroy(H->ev_read);
  if (H->pyr_stage_host) cudaFreeHost(H->pyr_stage_host);
  if (H->out_stage_host) cudaFreeHost(H->out_stage_host);
  for (cudaEvent_t e : H->ev_up) cudaEventDestroy(e);
  for (cudaEvent_t e : H->ev_done) cudaEventDestroy(e);
  for (cudaEvent_t e : H->stage_ev)
    if (e) cudaEventDestroy(e);
  if (H->stream) cudaStreamDestroy(H->stream);
  delete H;
  return ORB_OK;
}

int orbx_get_scale_table(const orbx_handle* H, int which, float* out) {
  if (!H || !out || which < 0 || which > 3) return set_error(ORB_E_INVALID, "orbx_get_scale_table: bad argument");
  const std::vector<float>& v = which == 0 ? H->scale : which == 1 ? H->inv_scale : which == 2 ? H->sigma2 : H->inv_sigma2;
  memcpy(out, v.data(), v.size() * sizeof(float));
  return ORB_OK;
}

int orbx_get_level_quota(const orbx_handle* H, int* out) {
  if (!H || !out) return set_error(ORB_E_INVALID, "orbx_get_level_quota: bad argument");
  memcpy(out, H->quota.data(), H->quota.size() * sizeof(int));
  return ORB_OK;
}

// The five stages of one batch (or chunk) on compute stream `cs`.  The blur needs only the pyramid,
// and the quadtree (one CTA per (frame, level), barrier- and latency-bound) leaves half the issue
// slots idle: the blur runs beside it on the lane's second stream and joins before the descriptors
// (1024 frames: 5.59 -> 5.52 ms, 64 frames: 0.545 -> 0.508 ms).  ORBX_FORK_BLUR=0 serialises.
static int launch_blur_after_quadtree(const ExtractParams& P, cudaStream_t s, int* launches) { return launch_blur(P, s, launches, true); }

// After the pyramid stage of a single frame: bordered levels -> page-locked staging on the side stream.
static int stage_pyramid(orbx_handle* H, const ExtractParams& Q, cudaStream_t cs, int* launches) {
  if (!H->pyr_staging || Q.n_frames != 1) return ORB_OK;
  size_t total = 0;
  for (int l = 0; l < Q.nlevels; ++l) {
    H->pyr_stage_off[l] = total;
    total += align_up<size_t>((size_t)(Q.lv[l].w + 2 * kEdge) * (Q.lv[l].h + 2 * kEdge), 256);
  }
  int rc;
  if ((rc = H->pyr_stage_dev.reserve(total))) return rc;
  if (total > H->pyr_stage_cap) {
    if (H->pyr_stage_host) cudaFreeHost(H->pyr_stage_host);
    H->pyr_stage_host = nullptr; H->pyr_stage_cap = 0;
    ORB_CUDA_TRY(cudaMallocHost(&H->pyr_stage_host, total));
    H->pyr_stage_cap = total;
  }
  if (!H->s_pyr) {
    ORB_CUDA_TRY(cudaStreamCreateWithFlags(&H->s_pyr, cudaStreamNonBlocking));
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&H->ev_pyr_ready, cudaEventDisableTiming));
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&H->ev_pyr_done, cudaEventDisableTiming));
  }
  ORB_CUDA_TRY(cudaEventRecord(H->ev_pyr_ready, cs));
  ORB_CUDA_TRY(cudaStreamWaitEvent(H->s_pyr, H->ev_pyr_ready, 0));
  if ((rc = launch_border_all(Q, 0, H->pyr_stage_off, H->pyr_stage_dev.p, H->s_pyr))) return rc;
  ++*launches;
  ORB_CUDA_TRY(cudaMemcpyAsync(H->pyr_stage_host, H->pyr_stage_dev.p, total, cudaMemcpyDeviceToHost, H->s_pyr));
  ORB_CUDA_TRY(cudaEventRecord(H->ev_pyr_done, H->s_pyr));
  H->pyr_pending = true;
  return ORB_OK;
}

static int run_stages(orbx_handle* H, const ExtractParams& Q, cudaStream_t cs, int lane, int* launches) {
  static const bool fork_blur = !(getenv("ORBX_FORK_BLUR") && atoi(getenv("ORBX_FORK_BLUR")) == 0);
  int rc;
  if (!fork_blur) {
    if ((rc = launch_pyramid(Q, cs, launches)) || (rc = stage_pyramid(H, Q, cs, launches)) || (rc = launch_fast(Q, cs, launches)) ||
        (rc = launch_quadtree(Q, cs, launches)) || (rc = launch_blur(Q, cs, launches)) || (rc = launch_describe(Q, cs, launches)))
      return rc;
    return ORB_OK;
  }
  while ((int)H->s_blur.size() <= lane) {
    cudaStream_t t = nullptr;
    cudaEvent_t a = nullptr, b = nullptr;
    ORB_CUDA_TRY(cudaStreamCreateWithFlags(&t, cudaStreamNonBlocking));
    H->s_blur.push_back(t);
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
    H->ev_fork.push_back(a);
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
    H->ev_join.push_back(b);
  }
  if ((rc = launch_pyramid(Q, cs, launches)) || (rc = stage_pyramid(H, Q, cs, launches)) || (rc = launch_fast(Q, cs, launches))) return rc;
  ORB_CUDA_TRY(cudaEventRecord(H->ev_fork[lane], cs));
  ORB_CUDA_TRY(cudaStreamWaitEvent(H->s_blur[lane], H->ev_fork[lane], 0));
  if ((rc = launch_blur(Q, H->s_blur[lane], launches, /*after_quadtree=*/false))) return rc;
  ORB_CUDA_TRY(cudaEventRecord(H->ev_join[lane], H->s_blur[lane]));
  if ((rc = launch_quadtree(Q, cs, launches))) return rc;
  ORB_CUDA_TRY(cudaStreamWaitEvent(cs, H->ev_join[lane], 0));
  return launch_describe(Q, cs, launches);
}

int orbx_extract_batch(orbx_handle* H, const uint8_t* imgs, int n, int w, int ht, size_t stride, size_t frame_stride,
                       int memory) {
  NvtxRange nvtx_range("orbx_extract_batch");
  if (!H || n < 0 || (memory != 0 && memory != 1)) return set_error(ORB_E_INVALID, "orbx_extract_batch: bad argument");
  H->counts_valid = false;
  if (n == 0 || !imgs || w <= 0 || ht <= 0) {  // empty image: nothing to do (ORBextractor.cc:1054-1055)
    H->n_frames = 0;
    return ORB_OK;
  }
  if (stride < (size_t)w || (n > 1 && frame_stride < stride * (size_t)(ht - 1) + w))
    return set_error(ORB_E_INVALID, "orbx_extract_batch: stride smaller than a row / frame");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  if (H->pyr_pending) {  // the previous frame's staging copy reads the buffers this call rewrites
    ORB_CUDA_TRY(cudaStreamWaitEvent(H->stream, H->ev_pyr_done, 0));
    H->pyr_pending = false;
  }
  if (H->read_pending) {  // so does an unfinished orbx_batch_read_begin
    ORB_CUDA_TRY(cudaStreamSynchronize(H->s_down));
    H->read_pending = false;
  }
  if ((rc = check_limits(H, n, w, ht))) return rc;
  const bool alias = memory == 1 && ((uintptr_t)imgs % 16 == 0) && stride % 16 == 0 && (n == 1 || frame_stride % 16 == 0);
  // host frames with 16-byte-multiple strides are uploaded with one linear copy and read in place
  const bool linear = memory == 0 && stride % 16 == 0 && (n == 1 || frame_stride % 16 == 0);
  if ((rc = configure(H, w, ht, n, !alias && !linear))) return rc;
  ExtractParams& P = H->P;
  if (alias) {
    P.l0 = imgs; P.l0_pitch = stride; P.l0_frame = frame_stride;
  } else if (linear) {
    if ((rc = H->l0.reserve((size_t)n * frame_stride + 256))) return rc;
    const size_t bytes = (size_t)(n - 1) * frame_stride + stride * (size_t)(ht - 1) + (size_t)w;
    ORB_CUDA_TRY(cudaMemcpyAsync(H->l0.p, imgs, bytes, cudaMemcpyHostToDevice, H->stream));
    P.l0 = H->l0.p; P.l0_pitch = stride; P.l0_frame = frame_stride;
  } else {
    const cudaMemcpyKind kind = memory ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (H->l0_frame == H->l0_pitch * (size_t)ht && frame_stride == stride * (size_t)ht) {
      ORB_CUDA_TRY(cudaMemcpy2DAsync(H->l0.p, H->l0_pitch, imgs, stride, (size_t)w, (size_t)ht * n, kind, H->stream));
    } else {
      for (int f = 0; f < n; ++f)
        ORB_CUDA_TRY(cudaMemcpy2DAsync(H->l0.p + (size_t)f * H->l0_frame, H->l0_pitch, imgs + (size_t)f * frame_stride, stride,
                                       (size_t)w, (size_t)ht, kind, H->stream));
    }
    P.l0 = H->l0.p; P.l0_pitch = H->l0_pitch; P.l0_frame = H->l0_frame;
  }
  int launches = 0;
  ORB_CUDA_TRY(cudaEventRecord(H->ev0, H->stream));
  typedef int (*stage_fn)(const ExtractParams&, cudaStream_t, int*);
  const stage_fn stages[ORBX_N_STAGES] = {launch_pyramid, launch_fast, launch_quadtree, launch_blur_after_quadtree, launch_describe};
  const bool timing = H->stage_timing;
  if (timing) ORB_CUDA_TRY(cudaEventRecord(H->stage_ev[0], H->stream));
  if (timing) {
    for (int s = 0; s < ORBX_N_STAGES; ++s) {
      if ((rc = stages[s](P, H->stream, &launches))) return rc;
      if (s == 0 && (rc = stage_pyramid(H, P, H->stream, &launches))) return rc;
      ORB_CUDA_TRY(cudaEventRecord(H->stage_ev[s + 1], H->stream));
    }
  } else if ((rc = run_stages(H, P, H->stream, 0, &launches))) {
    return rc;
  }
  H->stage_valid = timing;
  ORB_CUDA_TRY(cudaEventRecord(H->ev1, H->stream));
  H->n_frames = n;
  H->last_launches = launches;
  return ORB_OK;
}

int orbx_extract_batch_host(orbx_handle* H, const uint8_t* imgs, int n, int w, int ht, size_t stride, size_t frame_stride,
                            orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* counts, int chunk_frames) {
  NvtxRange nvtx_range("orbx_extract_batch_host");
  if (!H || n < 0 || !counts || (n > 0 && (!kps || !desc))) return set_error(ORB_E_INVALID, "orbx_extract_batch_host: bad argument");
  H->counts_valid = false;
  if (n == 0 || !imgs || w <= 0 || ht <= 0) {
    H->n_frames = 0;
    for (int f = 0; f < n; ++f) counts[f] = 0;
    return ORB_OK;
  }
  if (stride < (size_t)w || (n > 1 && frame_stride < stride * (size_t)(ht - 1) + w))
    return set_error(ORB_E_INVALID, "orbx_extract_batch_host: stride smaller than a row / frame");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  if (H->pyr_pending) {  // the previous frame's staging copy reads the buffers this call rewrites
    ORB_CUDA_TRY(cudaStreamWaitEvent(H->stream, H->ev_pyr_done, 0));
    H->pyr_pending = false;
  }
  if (H->read_pending) {  // so does an unfinished orbx_batch_read_begin
    ORB_CUDA_TRY(cudaStreamSynchronize(H->s_down));
    H->read_pending = false;
  }
  if ((rc = check_limits(H, n, w, ht))) return rc;
  // frames whose rows and frame stride are 16-byte multiples are uploaded as they are (one linear
  // copy per chunk) and read in place; other layouts go through pitched 2-D copies
  const bool linear = stride % 16 == 0 && (n == 1 || frame_stride % 16 == 0);
  if ((rc = configure(H, w, ht, n, !linear))) return rc;
  ExtractParams& P = H->P;
  if (cap_per_frame < P.out_cap)
    return set_error(ORB_E_CAPACITY, "orbx_extract_batch_host: cap_per_frame %d < %d (per-frame output capacity)", cap_per_frame, P.out_cap);
  if (linear) {
    if ((rc = H->l0.reserve((size_t)n * frame_stride + 256))) return rc;
    P.l0 = H->l0.p; P.l0_pitch = stride; P.l0_frame = frame_stride;
  } else {
    P.l0 = H->l0.p; P.l0_pitch = H->l0_pitch; P.l0_frame = H->l0_frame;
  }
  if (H->pinned_counts_cap < n) {
    if (H->pinned_counts) cudaFreeHost(H->pinned_counts);
    H->pinned_counts = nullptr; H->pinned_counts_cap = 0;
    ORB_CUDA_TRY(cudaMallocHost(&H->pinned_counts, sizeof(int) * (size_t)n));
    H->pinned_counts_cap = n;
  }
  // chunk schedule: the caller's size, else a sixteenth of the batch (at least 32 frames)
  std::vector<int> chunk_begin;
  static const char* sched_env = getenv("ORBX_CHUNK_SCHEDULE");  // developer knob: "16,48,64" (last size repeats)
  static const int e2e_skip = getenv("ORBX_E2E_SKIP") ? atoi(getenv("ORBX_E2E_SKIP")) : 0;  // developer knob: 1 no upload, 2 no download, 4 no kernels
  if (chunk_frames > 0) {
    for (int f0 = 0; f0 < n; f0 += chunk_frames) chunk_begin.push_back(f0);
  } else if (sched_env && *sched_env) {
    const char* q = sched_env;
    int size = 32;
    for (int f0 = 0; f0 < n; f0 += size) {
      if (*q) { size = std::max(1, atoi(q)); while (*q && *q != ',') ++q; if (*q == ',') ++q; }
      chunk_begin.push_back(f0);
    }
  } else {
    // a sixteenth of the batch, at least 32 frames (measured best for 640x480 with four compute streams: 64 of 1024,
    // tools/exp_e2e_streams.py), but no more than about 20 MB of frames per chunk, so that a batch of a few large
    // frames (16 x 3840x2160) is still pipelined: upload of chunk c+1 beside the kernels of chunk c
    const size_t frame_bytes = std::max<size_t>(1, frame_stride);
    const int size = std::max(1, std::min(std::max(32, n / 16), (int)std::max<size_t>(1, (20u << 20) / frame_bytes)));
    const int k = div_up(n, size);
    if (k >= 8 && size >= 16) {
      // The pipeline is bound by the upload stream and ends one chunk's extraction after the last upload, so the chunks
      // shrink towards the end -- 1.5x the mean size for the first three eighths, then 1, 0.75, 0.5 and 0.25x (96 x 6,
      // 64 x 4, 48 x 2, 32 x 2, 16 x 2 frames for 1024): large chunks where throughput counts, little work in flight when
      // the last byte arrives (1024 frames: 6.72 -> 6.54 ms per call against equal 64-frame chunks).
      std::vector<double> wgt(k);
      double sum = 0;
      for (int i = 0; i < k; ++i) {
        const double p = (double)i / k;
        wgt[i] = p < 0.375 ? 1.5 : p < 0.625 ? 1.0 : p < 0.75 ? 0.75 : p < 0.875 ? 0.5 : 0.25;
        sum += wgt[i];
      }
      const int unit = size >= 64 ? 8 : 1;  // whole multiples of 8 frames where chunks are large
      double acc = 0;
      for (int i = 0; i < k; ++i) {
        chunk_begin.push_back((int)(acc / unit + 0.5) * unit);
        acc += (double)n * wgt[i] / sum;
      }
    } else {
      for (int f0 = 0; f0 < n; f0 += size) chunk_begin.push_back(f0);
    }
  }
  chunk_begin.push_back(n);
  const int n_chunks = (int)chunk_begin.size() - 1;
  if (!H->s_up) ORB_CUDA_TRY(cudaStreamCreateWithFlags(&H->s_up, cudaStreamNonBlocking));
  if (!H->s_down) ORB_CUDA_TRY(cudaStreamCreateWithFlags(&H->s_down, cudaStreamNonBlocking));
  if (!H->s_alt) ORB_CUDA_TRY(cudaStreamCreateWithFlags(&H->s_alt, cudaStreamNonBlocking));
  static const int n_comp_env = getenv("ORBX_COMPUTE_STREAMS") ? atoi(getenv("ORBX_COMPUTE_STREAMS")) : 4;
  const int n_comp = std::min(std::max(n_comp_env, 1), 8);
  while ((int)H->s_more.size() < n_comp - 2) {
    cudaStream_t t = nullptr;
    ORB_CUDA_TRY(cudaStreamCreateWithFlags(&t, cudaStreamNonBlocking));
    H->s_more.push_back(t);
  }
  while ((int)H->ev_up.size() < n_chunks) {
    cudaEvent_t a = nullptr, b = nullptr;
    static const bool trace = getenv("ORBX_E2E_TRACE") != nullptr;  // developer knob: timed events, timeline on stderr
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&a, trace ? cudaEventDefault : cudaEventDisableTiming));
    H->ev_up.push_back(a);
    ORB_CUDA_TRY(cudaEventCreateWithFlags(&b, trace ? cudaEventDefault : cudaEventDisableTiming));
    H->ev_done.push_back(b);
  }
  int launches = 0;
  const int oc = P.out_cap;
  // the copy streams must not run ahead of work already queued on the handle's stream
  ORB_CUDA_TRY(cudaEventRecord(H->ev0, H->stream));
  ORB_CUDA_TRY(cudaStreamWaitEvent(H->s_up, H->ev0, 0));
  for (int c = 0; c < n_chunks; ++c) {
    const int f0 = chunk_begin[c], nf = chunk_begin[c + 1] - f0;
    static int skip_calls = 0;  // the first batch is always uploaded, so the kernels see real frames
    if ((e2e_skip & 1) && skip_calls >= n_chunks) {
    } else if (linear) {
      const size_t bytes = (size_t)(nf - 1) * frame_stride + stride * (size_t)(ht - 1) + (size_t)w;
      ORB_CUDA_TRY(cudaMemcpyAsync(H->l0.p + (size_t)f0 * frame_stride, imgs + (size_t)f0 * frame_stride, bytes, cudaMemcpyHostToDevice, H->s_up));
    } else {
      for (int f = f0; f < f0 + nf; ++f)
        ORB_CUDA_TRY(cudaMemcpy2DAsync(H->l0.p + (size_t)f * H->l0_frame, H->l0_pitch, imgs + (size_t)f * frame_stride, stride, (size_t)w,
                                       (size_t)ht, cudaMemcpyHostToDevice, H->s_up));
    }
    ++skip_calls;
    ORB_CUDA_TRY(cudaEventRecord(H->ev_up[c], H->s_up));
    // consecutive chunks rotate over the compute streams: one chunk's tails (few CTAs left,
    // latency-bound quadtree) overlap the streaming kernels of the chunks behind it
    const int lane = c % n_comp;
    cudaStream_t cs = lane == 0 ? H->stream : lane == 1 ? H->s_alt : H->s_more[lane - 2];
    ORB_CUDA_TRY(cudaStreamWaitEvent(cs, H->ev_up[c], 0));
    const ExtractParams Q = frame_range(P, f0, nf);
    if (!(e2e_skip & 4) && (rc = run_stages(H, Q, cs, lane, &launches))) return rc;
    ORB_CUDA_TRY(cudaEventRecord(H->ev_done[c], cs));
    if (e2e_skip & 2) continue;
    ORB_CUDA_TRY(cudaStreamWaitEvent(H->s_down, H->ev_done[c], 0));
    ORB_CUDA_TRY(cudaMemcpyAsync(H->pinned_counts + f0, H->out_count.p + f0, sizeof(int) * (size_t)nf, cudaMemcpyDeviceToHost, H->s_down));
    if (cap_per_frame == oc) {  // same row count on both sides: linear copies
      ORB_CUDA_TRY(cudaMemcpyAsync(kps + (size_t)f0 * oc, H->out_kp.p + (size_t)f0 * oc, sizeof(orbx_keypoint) * (size_t)oc * nf,
                                   cudaMemcpyDeviceToHost, H->s_down));
      ORB_CUDA_TRY(cudaMemcpyAsync(desc + (size_t)f0 * oc * 32, H->out_desc.p + (size_t)f0 * oc * 32, 32 * (size_t)oc * nf,
                                   cudaMemcpyDeviceToHost, H->s_down));
    } else {
      ORB_CUDA_TRY(cudaMemcpy2DAsync(kps + (size_t)f0 * cap_per_frame, sizeof(orbx_keypoint) * (size_t)cap_per_frame,
                                     H->out_kp.p + (size_t)f0 * oc, sizeof(orbx_keypoint) * (size_t)oc, sizeof(orbx_keypoint) * (size_t)oc,
                                     (size_t)nf, cudaMemcpyDeviceToHost, H->s_down));
      ORB_CUDA_TRY(cudaMemcpy2DAsync(desc + (size_t)f0 * cap_per_frame * 32, 32 * (size_t)cap_per_frame, H->out_desc.p + (size_t)f0 * oc * 32,
                                     32 * (size_t)oc, 32 * (size_t)oc, (size_t)nf, cudaMemcpyDeviceToHost, H->s_down));
    }
  }
  // the other compute streams join the handle's stream once, at the end (a join per chunk would
  // hold the next chunk of the handle's stream back until its predecessor has drained)
  for (int c = std::max(0, n_chunks - n_comp); c < n_chunks; ++c)
    if (c % n_comp != 0) ORB_CUDA_TRY(cudaStreamWaitEvent(H->stream, H->ev_done[c], 0));
  ORB_CUDA_TRY(cudaEventRecord(H->ev1, H->stream));
  H->stage_valid = false;
  H->n_frames = n;
  H->last_launches = launches;
  ORB_CUDA_TRY(cudaStreamSynchronize(H->s_down));
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  memcpy(counts, H->pinned_counts, sizeof(int) * (size_t)n);
  if (getenv("ORBX_E2E_TRACE")) {
    static int call = 0;
    if (++call == atoi(getenv("ORBX_E2E_TRACE"))) {
      float end = 0;
      cudaEventElapsedTime(&end, H->ev0, H->ev1);
      fprintf(stderr, "e2e trace (ms after the call started; %d chunks, stream end %.3f)\n", n_chunks, end);
      for (int c = 0; c < n_chunks; ++c) {
        float up = 0, done = 0;
        cudaEventElapsedTime(&up, H->ev0, H->ev_up[c]);
        cudaEventElapsedTime(&done, H->ev0, H->ev_done[c]);
        fprintf(stderr, "  chunk %2d frames %4d..%4d  uploaded %.3f  extracted %.3f  (+%.3f)\n", c, chunk_begin[c], chunk_begin[c + 1], up, done, done - up);
      }
    }
  }
  return ORB_OK;
}

int orbx_frame_capacity(orbx_handle* H, int w, int ht, int* cap) {
  if (!H || !cap || w <= 0 || ht <= 0) return set_error(ORB_E_INVALID, "orbx_frame_capacity: bad argument");
  int total = 0;
  for (int l = 0; l < H->cfg.nlevels; ++l) {  // per level: quota + 4 * roots + 3 (the last split may overshoot by 3)
    const int lw = cv_round((float)w * H->inv_scale[l]), lh = cv_round((float)ht * H->inv_scale[l]);
    const int bw = lw - 2 * (kEdge - 3), bh = lh - 2 * (kEdge - 3);
    if (bw < 30 || bh < 30) return set_error(ORB_E_INVALID, "orbx_frame_capacity: level %d is %dx%d, too small", l, lw, lh);
    const int nini = std::max(1, (int)roundf((float)bw / (float)bh));
    total += std::max(H->quota[l], 0) + 4 * nini + 3;
  }
  *cap = total;
  return ORB_OK;
}

int orbx_sync(orbx_handle* H) {
  if (!H) return set_error(ORB_E_INVALID, "orbx_sync: null");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  return ORB_OK;
}

static int fetch_counts(orbx_handle* H) {
  if (H->counts_valid) return ORB_OK;
  H->host_counts.assign((size_t)H->n_frames, 0);
  if (H->n_frames) {
    ORB_CUDA_TRY(cudaMemcpyAsync(H->host_counts.data(), H->out_count.p, sizeof(int) * (size_t)H->n_frames,
                                 cudaMemcpyDeviceToHost, H->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  }
  H->counts_valid = true;
  return ORB_OK;
}

int orbx_batch_counts(orbx_handle* H, int* counts) {
  if (!H || !counts) return set_error(ORB_E_INVALID, "orbx_batch_counts: bad argument");
  int rc = use_device(H->cfg.device);
  if (rc || (rc = fetch_counts(H))) return rc;
  memcpy(counts, H->host_counts.data(), sizeof(int) * (size_t)H->n_frames);
  return ORB_OK;
}

int orbx_batch_read(orbx_handle* H, int frame, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out) {
  if (!H || !n_out || cap < 0) return set_error(ORB_E_INVALID, "orbx_batch_read: bad argument");
  if (H->n_frames == 0) { *n_out = 0; return ORB_OK; }
  if (frame < 0 || frame >= H->n_frames) return set_error(ORB_E_INVALID, "orbx_batch_read: frame %d out of range", frame);
  int rc = use_device(H->cfg.device);
  if (rc || (rc = fetch_counts(H))) return rc;
  const int n = H->host_counts[frame];
  *n_out = n;
  if (n > cap) return set_error(ORB_E_CAPACITY, "orbx_batch_read: %d keypoints, capacity %d", n, cap);
  if (n == 0) return ORB_OK;
  if (!kps || !desc) return set_error(ORB_E_INVALID, "orbx_batch_read: null output");
  const size_t at = (size_t)frame * H->P.out_cap;
  ORB_CUDA_TRY(cudaMemcpyAsync(kps, H->out_kp.p + at, sizeof(orbx_keypoint) * (size_t)n, cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(desc, H->out_desc.p + at * 32, (size_t)n * 32, cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  return ORB_OK;
}

int orbx_batch_read_all(orbx_handle* H, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* counts) {
  if (!H || !kps || !desc || !counts) return set_error(ORB_E_INVALID, "orbx_batch_read_all: bad argument");
  if (H->n_frames == 0) return ORB_OK;
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  const int oc = H->P.out_cap;
  if (cap_per_frame < oc)
    return set_error(ORB_E_CAPACITY, "orbx_batch_read_all: cap_per_frame %d < %d (per-frame output capacity)", cap_per_frame, oc);
  ORB_CUDA_TRY(cudaMemcpyAsync(counts, H->out_count.p, sizeof(int) * (size_t)H->n_frames, cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaMemcpy2DAsync(kps, sizeof(orbx_keypoint) * (size_t)cap_per_frame, H->out_kp.p, sizeof(orbx_keypoint) * (size_t)oc,
                                 sizeof(orbx_keypoint) * (size_t)oc, (size_t)H->n_frames, cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaMemcpy2DAsync(desc, 32 * (size_t)cap_per_frame, H->out_desc.p, 32 * (size_t)oc, 32 * (size_t)oc,
                                 (size_t)H->n_frames, cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  return ORB_OK;
}

int orbx_batch_read_begin(orbx_handle* H, orbx_keypoint* kps, uint8_t* desc, int cap_per_frame, int* counts) {
  if (!H || !kps || !desc || !counts) return set_error(ORB_E_INVALID, "orbx_batch_read_begin: bad argument");
  if (H->n_frames == 0) return ORB_OK;
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  const int oc = H->P.out_cap;
  if (cap_per_frame < oc)
    return set_error(ORB_E_CAPACITY, "orbx_batch_read_begin: cap_per_frame %d < %d (per-frame output capacity)", cap_per_frame, oc);
  if (!H->s_down) ORB_CUDA_TRY(cudaStreamCreateWithFlags(&H->s_down, cudaStreamNonBlocking));
  if (!H->ev_read) ORB_CUDA_TRY(cudaEventCreateWithFlags(&H->ev_read, cudaEventDisableTiming));
  // behind the extraction, beside whatever the handle's stream does next
  ORB_CUDA_TRY(cudaEventRecord(H->ev_read, H->stream));
  ORB_CUDA_TRY(cudaStreamWaitEvent(H->s_down, H->ev_read, 0));
  ORB_CUDA_TRY(cudaMemcpyAsync(counts, H->out_count.p, sizeof(int) * (size_t)H->n_frames, cudaMemcpyDeviceToHost, H->s_down));
  ORB_CUDA_TRY(cudaMemcpy2DAsync(kps, sizeof(orbx_keypoint) * (size_t)cap_per_frame, H->out_kp.p, sizeof(orbx_keypoint) * (size_t)oc,
                                 sizeof(orbx_keypoint) * (size_t)oc, (size_t)H->n_frames, cudaMemcpyDeviceToHost, H->s_down));
  ORB_CUDA_TRY(cudaMemcpy2DAsync(desc, 32 * (size_t)cap_per_frame, H->out_desc.p, 32 * (size_t)oc, 32 * (size_t)oc,
                                 (size_t)H->n_frames, cudaMemcpyDeviceToHost, H->s_down));
  H->read_pending = true;
  return ORB_OK;
}

int orbx_batch_read_end(orbx_handle* H) {
  if (!H) return set_error(ORB_E_INVALID, "orbx_batch_read_end: null");
  if (!H->read_pending) return ORB_OK;
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaStreamSynchronize(H->s_down));
  H->read_pending = false;
  return ORB_OK;
}

int orbx_extract(orbx_handle* H, const uint8_t* img, int w, int ht, size_t stride, orbx_keypoint* kps, uint8_t* desc,
                 int cap, int* n_out) {
  if (!H || !n_out) return set_error(ORB_E_INVALID, "orbx_extract: null argument");
  *n_out = 0;
  int rc = orbx_extract_batch(H, img, 1, w, ht, stride, stride * (size_t)std::max(ht, 0), 0);
  if (rc) return rc;
  return orbx_batch_read(H, 0, kps, desc, cap, n_out);
}

}  // extern "C"
namespace orbfront {
int pyramid_view(orbx_handle* H, int frame, PyramidView* out) {
  if (!H || !out) return set_error(ORB_E_INVALID, "pyramid_view: bad argument");
  if (H->n_frames == 0) return set_error(ORB_E_STATE, "stereo matching: nothing extracted yet");
  if (frame < 0 || frame >= H->n_frames) return set_error(ORB_E_INVALID, "stereo matching: frame %d out of range", frame);
  if (H->cfg.nlevels > PyramidView::kLevels) return set_error(ORB_E_INVALID, "stereo matching: more than %d levels", PyramidView::kLevels);
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  const ExtractParams& P = H->P;
  for (int l = 0; l < H->cfg.nlevels; ++l) {
    const LevelGeom& G = P.lv[l];
    out->px[l] = l == 0 ? P.l0 + (size_t)frame * P.l0_frame : P.pyr + (size_t)frame * P.pyr_frame + G.off;
    out->pitch[l] = l == 0 ? (int)P.l0_pitch : G.pitch;
    out->w[l] = G.w; out->h[l] = G.h;
    out->scale[l] = H->scale[l]; out->inv_scale[l] = H->inv_scale[l];
  }
  out->n_levels = H->cfg.nlevels;
  out->device = H->cfg.device;
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  return ORB_OK;
}
int batch_view(orbx_handle* H, BatchView* out) {
  if (!H || !out) return set_error(ORB_E_INVALID, "batch_view: bad argument");
  if (H->n_frames == 0) return set_error(ORB_E_STATE, "nothing extracted yet");
  const ExtractParams& P = H->P;
  out->kp = P.out_kp; out->desc = P.out_desc; out->count = P.out_count;
  out->out_cap = P.out_cap; out->n_frames = H->n_frames; out->device = H->cfg.device;
  out->level0_frame_stride = P.l0_frame; out->level_frame_stride = P.pyr_frame;
  out->stream = H->stream;
  return ORB_OK;
}
}  // namespace orbfront
extern "C" {

int orbx_level_size(const orbx_handle* H, int level, int* w, int* ht) {
  if (!H || !w || !ht || level < 0 || level >= H->cfg.nlevels) return set_error(ORB_E_INVALID, "orbx_level_size: bad argument");
  if (H->w == 0) return set_error(ORB_E_STATE, "orbx_level_size: no frame size configured yet");
  *w = H->P.lv[level].w;
  *ht = H->P.lv[level].h;
  return ORB_OK;
}

int orbx_get_level(orbx_handle* H, int frame, int level, uint8_t* dst, size_t dst_stride, int with_border) {
  if (!H || !dst || level < 0 || level >= H->cfg.nlevels) return set_error(ORB_E_INVALID, "orbx_get_level: bad argument");
  if (H->n_frames == 0) return set_error(ORB_E_STATE, "orbx_get_level: nothing extracted yet");
  if (frame < 0 || frame >= H->n_frames) return set_error(ORB_E_INVALID, "orbx_get_level: frame out of range");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  const ExtractParams& P = H->P;
  const LevelGeom& G = P.lv[level];
  const uint8_t* src = level == 0 ? P.l0 + (size_t)frame * P.l0_frame : P.pyr + (size_t)frame * P.pyr_frame + G.off;
  const size_t pitch = level == 0 ? P.l0_pitch : (size_t)G.pitch;
  if (!with_border) {
    if (dst_stride < (size_t)G.w) return set_error(ORB_E_INVALID, "orbx_get_level: dst_stride too small");
    ORB_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_stride, src, pitch, (size_t)G.w, (size_t)G.h, cudaMemcpyDeviceToHost, H->stream));
  } else {
    const int W = G.w + 2 * kEdge, Hh = G.h + 2 * kEdge;
    if (dst_stride < (size_t)W) return set_error(ORB_E_INVALID, "orbx_get_level: dst_stride too small");
    if ((rc = H->border_tmp.reserve((size_t)W * Hh))) return rc;
    if ((rc = launch_border_copy(src, G.w, G.h, pitch, H->border_tmp.p, H->stream))) return rc;
    ORB_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_stride, H->border_tmp.p, (size_t)W, (size_t)W, (size_t)Hh, cudaMemcpyDeviceToHost, H->stream));
  }
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  return ORB_OK;
}

int orbx_set_pyramid_staging(orbx_handle* H, int enabled) {
  if (!H) return set_error(ORB_E_INVALID, "orbx_set_pyramid_staging: null");
  H->pyr_staging = enabled != 0;
  return ORB_OK;
}

int orbx_get_pyramid_host(orbx_handle* H, int level, const uint8_t** pixel00, int* w, int* ht, size_t* stride) {
  if (!H || !pixel00 || !w || !ht || !stride || level < 0 || level >= H->cfg.nlevels)
    return set_error(ORB_E_INVALID, "orbx_get_pyramid_host: bad argument");
  if (!H->pyr_staging || H->n_frames != 1 || !H->pyr_stage_host || !H->ev_pyr_done)
    return set_error(ORB_E_STATE, "orbx_get_pyramid_host: no single frame extracted with pyramid staging enabled");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaEventSynchronize(H->ev_pyr_done));
  const LevelGeom& G = H->P.lv[level];
  const size_t W = (size_t)G.w + 2 * kEdge;
  *pixel00 = H->pyr_stage_host + H->pyr_stage_off[level] + (size_t)kEdge * W + kEdge;
  *w = G.w; *ht = G.h; *stride = W;
  return ORB_OK;
}

int orbx_read_frame0(orbx_handle* H, orbx_keypoint* kps, uint8_t* desc, int cap, int* n_out) {
  if (!H || !n_out || cap < 0) return set_error(ORB_E_INVALID, "orbx_read_frame0: bad argument");
  *n_out = 0;
  if (H->n_frames == 0) return ORB_OK;
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  const size_t oc = (size_t)H->P.out_cap, need = 256 + oc * (sizeof(orbx_keypoint) + 32);
  if (need > H->out_stage_cap) {
    if (H->out_stage_host) cudaFreeHost(H->out_stage_host);
    H->out_stage_host = nullptr; H->out_stage_cap = 0;
    ORB_CUDA_TRY(cudaMallocHost(&H->out_stage_host, need));
    H->out_stage_cap = need;
  }
  uint8_t* st = H->out_stage_host;
  ORB_CUDA_TRY(cudaMemcpyAsync(st, H->out_count.p, sizeof(int), cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(st + 256, H->out_kp.p, oc * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaMemcpyAsync(st + 256 + oc * sizeof(orbx_keypoint), H->out_desc.p, oc * 32, cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  const int n = *reinterpret_cast<const int*>(st);
  *n_out = n;
  if (n > cap) return set_error(ORB_E_CAPACITY, "orbx_read_frame0: %d keypoints, capacity %d", n, cap);
  if (n == 0) return ORB_OK;
  if (!kps || !desc) return set_error(ORB_E_INVALID, "orbx_read_frame0: null output");
  memcpy(kps, st + 256, (size_t)n * sizeof(orbx_keypoint));
  memcpy(desc, st + 256 + oc * sizeof(orbx_keypoint), (size_t)n * 32);
  return ORB_OK;
}

int orbx_get_blur_level(orbx_handle* H, int frame, int level, uint8_t* dst, size_t dst_stride) {
  if (!H || !dst || level < 0 || level >= H->cfg.nlevels) return set_error(ORB_E_INVALID, "orbx_get_blur_level: bad argument");
  if (H->n_frames == 0) return set_error(ORB_E_STATE, "orbx_get_blur_level: nothing extracted yet");
  if (frame < 0 || frame >= H->n_frames) return set_error(ORB_E_INVALID, "orbx_get_blur_level: frame out of range");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  const ExtractParams& P = H->P;
  const LevelGeom& G = P.lv[level];
  if (dst_stride < (size_t)G.w) return set_error(ORB_E_INVALID, "orbx_get_blur_level: dst_stride too small");
  ORB_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_stride, P.blur + (size_t)frame * P.blur_frame + G.off, (size_t)G.pitch, (size_t)G.w, (size_t)G.h,
                                 cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  return ORB_OK;
}

int orbx_get_candidates(orbx_handle* H, int frame, int level, int32_t* xys, int cap, int* n_out) {
  if (!H || !n_out || level < 0 || level >= H->cfg.nlevels || cap < 0) return set_error(ORB_E_INVALID, "orbx_get_candidates: bad argument");
  if (H->n_frames == 0) return set_error(ORB_E_STATE, "orbx_get_candidates: nothing extracted yet");
  if (frame < 0 || frame >= H->n_frames) return set_error(ORB_E_INVALID, "orbx_get_candidates: frame out of range");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  const ExtractParams& P = H->P;
  int n = 0;
  ORB_CUDA_TRY(cudaMemcpyAsync(&n, P.cand_count + frame * P.nlevels + level, sizeof(int), cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  *n_out = n;
  if (n > cap) return set_error(ORB_E_CAPACITY, "orbx_get_candidates: %d candidates, capacity %d", n, cap);
  if (n == 0) return ORB_OK;
  if (!xys) return set_error(ORB_E_INVALID, "orbx_get_candidates: null output");
  std::vector<unsigned long long> tmp((size_t)n);
  ORB_CUDA_TRY(cudaMemcpyAsync(tmp.data(), P.cand + (size_t)frame * P.cand_frame + P.lv[level].cand_off, sizeof(unsigned long long) * (size_t)n,
                               cudaMemcpyDeviceToHost, H->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(H->stream));
  for (int i = 0; i < n; ++i) {
    const uint32_t pay = (uint32_t)tmp[i];
    xys[3 * i] = (int)(pay & 0xfff);
    xys[3 * i + 1] = (int)((pay >> 12) & 0xfff);
    xys[3 * i + 2] = (int)(pay >> 24);
  }
  return ORB_OK;
}

int orbx_last_batch_ms(orbx_handle* H, float* ms, int* launches) {
  if (!H || H->n_frames == 0) return set_error(ORB_E_STATE, "orbx_last_batch_ms: nothing extracted yet");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaEventSynchronize(H->ev1));
  if (ms) ORB_CUDA_TRY(cudaEventElapsedTime(ms, H->ev0, H->ev1));
  if (launches) *launches = H->last_launches;
  return ORB_OK;
}

int orbx_get_stream(orbx_handle* H, void** cuda_stream) {
  if (!H || !cuda_stream) return set_error(ORB_E_INVALID, "orbx_get_stream: bad argument");
  *cuda_stream = (void*)H->stream;
  return ORB_OK;
}

int orbx_set_stage_timing(orbx_handle* H, int enabled) {
  if (!H) return set_error(ORB_E_INVALID, "orbx_set_stage_timing: null");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  if (enabled)
    for (cudaEvent_t& e : H->stage_ev)
      if (!e) ORB_CUDA_TRY(cudaEventCreate(&e));
  H->stage_timing = enabled != 0;
  H->stage_valid = false;
  return ORB_OK;
}

int orbx_last_stage_ms(orbx_handle* H, float* ms) {
  if (!H || !ms) return set_error(ORB_E_INVALID, "orbx_last_stage_ms: bad argument");
  if (!H->stage_valid) return set_error(ORB_E_STATE, "orbx_last_stage_ms: no batch has run with stage timing enabled");
  int rc = use_device(H->cfg.device);
  if (rc) return rc;
  ORB_CUDA_TRY(cudaEventSynchronize(H->stage_ev[ORBX_N_STAGES]));
  for (int s = 0; s < ORBX_N_STAGES; ++s) ORB_CUDA_TRY(cudaEventElapsedTime(&ms[s], H->stage_ev[s], H->stage_ev[s + 1]));
  return ORB_OK;
}

}  // extern "C"
