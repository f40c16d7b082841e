// tests/cpp/voc_harness.cc -- TEST INFRASTRUCTURE.
//
// Drives the product's include/ORBVocabulary.h (GPU transform) and, from the very same object, the
// vendored reference implementation it derives from (DBoW2::TemplatedVocabulary<FORB>::transform,
// compiled UNMODIFIED from /root/reference/Thirdparty/DBoW2 over oracle/cvshim), so pytest can compare
// BowVector and FeatureVector entry by entry.  Built only where /root/reference exists
// (_build/libdropin_voc.so travels to the GPU box with the snapshot).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "ORBVocabulary.h"

using ORB_SLAM2::ORBVocabulary;
typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> RefVocabulary;

static thread_local std::string g_err;

extern "C" {

const char* voc_error() { return g_err.c_str(); }

void* voc_load(const char* path) {
  try {
    ORBVocabulary* v = new ORBVocabulary();
    if (!v->loadFromTextFile(path) || v->empty()) { delete v; g_err = "loadFromTextFile failed"; return nullptr; }
    return v;
  } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void voc_free(void* v) { delete (ORBVocabulary*)v; }
int voc_words(void* v) { return (int)((ORBVocabulary*)v)->size(); }

// which: 0 = product (GPU descent), 1 = reference (DBoW2's own transform on the host)
int voc_transform(void* vp, const uint8_t* desc, int n, int levelsup, int which, uint32_t* bow_id, double* bow_val, int* n_bow,
                  uint32_t* fv_node, int32_t* fv_off, uint32_t* fv_idx, int* n_fv) {
  try {
    ORBVocabulary* v = (ORBVocabulary*)vp;
    std::vector<cv::Mat> feats;
    for (int i = 0; i < n; ++i) feats.push_back(cv::Mat(1, 32, CV_8U, (void*)(desc + 32 * (size_t)i), 32));
    DBoW2::BowVector bv;
    DBoW2::FeatureVector fv;
    if (which == 0) v->transform(feats, bv, fv, levelsup);
    else static_cast<RefVocabulary*>(v)->RefVocabulary::transform(feats, bv, fv, levelsup);
    int k = 0;
    for (DBoW2::BowVector::const_iterator it = bv.begin(); it != bv.end(); ++it, ++k) { bow_id[k] = it->first; bow_val[k] = it->second; }
    *n_bow = k;
    int m = 0, pos = 0;
    fv_off[0] = 0;
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++m) {
      fv_node[m] = it->first;
      for (size_t j = 0; j < it->second.size(); ++j) fv_idx[pos++] = it->second[j];
      fv_off[m + 1] = pos;
    }
    *n_fv = m;
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

}  // extern "C"
