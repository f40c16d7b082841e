// oracle/ref_harness.cc -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C-ABI wrapper around the reference's own ORB_SLAM2::ORBextractor, compiled from
// /root/reference/src/ORBextractor.cc UNMODIFIED (by oracle/Makefile, into oracle/_ref/)
// against oracle/cvshim.  Used to validate the restatement in orb_port.hpp and, in the
// -O3 build, as the timed CPU baseline ("kind": "reference").
//
// With -DORBREF_ARENA the global operator new is a monotonic arena while a call is in
// flight, so ExtractorNode addresses increase in creation order and the pointer tie-break
// of std::sort at src/ORBextractor.cc:684 becomes deterministic.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <new>

#include "ORBextractor.h"

#ifdef ORBREF_ARENA
namespace {
const size_t kArenaBytes = (size_t)3 << 30;  // virtual; touched lazily
char* g_base = nullptr;
size_t g_off = 0;
bool g_on = false;
inline bool in_arena(void* p) { return g_base && (char*)p >= g_base && (char*)p < g_base + kArenaBytes; }
void* arena_alloc(size_t n) {
  if (!g_on) {
    void* p = malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
  }
  size_t a = (g_off + 15) & ~(size_t)15;
  if (a + n > kArenaBytes) { fprintf(stderr, "ref_harness: arena exhausted\n"); abort(); }
  g_off = a + n;
  return g_base + a;
}
struct ArenaScope {
  ArenaScope() {
    if (!g_base) g_base = (char*)malloc(kArenaBytes);  // overcommitted, pages touched on use
    g_off = 0;
    g_on = true;
  }
  ~ArenaScope() { g_on = false; g_off = 0; }
};
}  // namespace
void* operator new(size_t n) { return arena_alloc(n); }
void* operator new[](size_t n) { return arena_alloc(n); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) free(p); }
void operator delete[](void* p) noexcept { if (p && !in_arena(p)) free(p); }
void operator delete(void* p, size_t) noexcept { if (p && !in_arena(p)) free(p); }
void operator delete[](void* p, size_t) noexcept { if (p && !in_arena(p)) free(p); }
#else
namespace { struct ArenaScope {}; }
#endif

namespace {
// Derived only to reach the protected stage methods for stage-level parity checks.
class Probe : public ORB_SLAM2::ORBextractor {
 public:
  using ORB_SLAM2::ORBextractor::ORBextractor;
  int quota(int l) const { return mnFeaturesPerLevel[l]; }
  int umax_at(int v) const { return umax[v]; }
  void stage_keypoints(std::vector<std::vector<cv::KeyPoint>>& all) { ComputeKeyPointsOctTree(all); }
  void stage_pyramid(const cv::Mat& im) { ComputePyramid(im); }
};
}  // namespace

struct RefKp { float x, y, size, angle, response; int octave, class_id; };

extern "C" {

void* orbref_create(int nf, float sf, int nl, int ini, int mn) { return new Probe(nf, sf, nl, ini, mn); }
void orbref_destroy(void* h) { delete (Probe*)h; }

// operator(): returns the keypoint count (or -needed if cap is too small).
int orbref_extract(void* h, const uint8_t* img, int w, int h_, size_t stride, RefKp* kps, uint8_t* desc, int cap) {
  Probe* e = (Probe*)h;
  int n;
  {
    ArenaScope scope;
    cv::Mat im(h_, w, CV_8U, (void*)img, stride);
    std::vector<cv::KeyPoint> k;
    cv::Mat d;
    (*e)(im, cv::Mat(), k, d);
    n = (int)k.size();
    if (n <= cap) {
      for (int i = 0; i < n; ++i) {
        kps[i] = RefKp{k[i].pt.x, k[i].pt.y, k[i].size, k[i].angle, k[i].response, k[i].octave, k[i].class_id};
        memcpy(desc + 32 * (size_t)i, d.ptr(i), 32);
      }
    } else {
      n = -n;
    }
  }
  return n;
}

// Per-level kept keypoints (level coordinates incl. border offset, before scaling and
// descriptors): stage output of ComputeKeyPointsOctTree.  Requires a prior extract/pyramid.
int orbref_stage_keypoints(void* h, const uint8_t* img, int w, int h_, size_t stride, RefKp* kps, int* level_counts, int cap) {
  Probe* e = (Probe*)h;
  int n = 0;
  {
    ArenaScope scope;
    cv::Mat im(h_, w, CV_8U, (void*)img, stride);
    e->stage_pyramid(im);
    std::vector<std::vector<cv::KeyPoint>> all;
    e->stage_keypoints(all);
    for (size_t l = 0; l < all.size(); ++l) {
      level_counts[l] = (int)all[l].size();
      for (const auto& k : all[l]) {
        if (n < cap) kps[n] = RefKp{k.pt.x, k.pt.y, k.size, k.angle, k.response, k.octave, k.class_id};
        ++n;
      }
    }
  }
  return n <= cap ? n : -n;
}

int orbref_level_size(void* h, int l, int* w, int* h_) {
  Probe* e = (Probe*)h;
  if (l < 0 || l >= e->GetLevels() || e->mvImagePyramid[l].empty()) return -1;
  *w = e->mvImagePyramid[l].cols;
  *h_ = e->mvImagePyramid[l].rows;
  return 0;
}

// Copies level l; with_border != 0 copies the (h+38)x(w+38) framed buffer.
int orbref_get_level(void* h, int l, uint8_t* dst, size_t dstep, int with_border) {
  Probe* e = (Probe*)h;
  if (l < 0 || l >= e->GetLevels() || e->mvImagePyramid[l].empty()) return -1;
  const cv::Mat& m = e->mvImagePyramid[l];
  const int b = with_border ? 19 : 0;
  for (int y = -b; y < m.rows + b; ++y)
    memcpy(dst + (size_t)(y + b) * dstep, m.data + (ptrdiff_t)y * (ptrdiff_t)m.step - b, (size_t)(m.cols + 2 * b));
  return 0;
}

int orbref_quota(void* h, int l) { return ((Probe*)h)->quota(l); }
int orbref_umax(void* h, int v) { return ((Probe*)h)->umax_at(v); }
float orbref_scale(void* h, int l) { return ((Probe*)h)->GetScaleFactors()[l]; }
int orbref_has_arena(void) {
#ifdef ORBREF_ARENA
  return 1;
#else
  return 0;
#endif
}
}
