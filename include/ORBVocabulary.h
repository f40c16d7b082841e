/* ORBVocabulary.h -- drop-in for the reference's include/ORBVocabulary.h:27-31.
 *
 * The reference's ORBVocabulary is a typedef of DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>.
 * Here it is a subclass of that same template (the vendored Thirdparty/DBoW2 stays in the build: the
 * text loader, scoring objects, BowVector / FeatureVector are used as they are), which overrides the
 * one hot member: transform(features, BowVector&, FeatureVector&, levelsup)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1196), called once per frame / keyframe by
 * Frame::ComputeBoW (src/Frame.cc:410-417) and KeyFrame::ComputeBoW (src/KeyFrame.cc:70-79).
 * The per-feature tree descent (:1217-1259, k x L Hamming distances per descriptor) runs on the GPU
 * over a device-resident copy of the tree (orbm_vocab_*); the BowVector / FeatureVector are then
 * filled on the host in feature order with the reference's own addWeight / addFeature / normalize,
 * so every double in the BowVector is produced by the same sequence of operations.
 * No CPU fallback: without a CUDA device transform() throws.
 */
#ifndef ORBVOCABULARY_H
#define ORBVOCABULARY_H

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "Thirdparty/DBoW2/DBoW2/FORB.h"
#include "Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h"

#include "orbfront.h"

namespace ORB_SLAM2 {

class ORBVocabulary : public DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> {
  typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> Base;

 public:
  ORBVocabulary() : dev_(NULL) {}
  ORBVocabulary(const ORBVocabulary&) = delete;
  ORBVocabulary& operator=(const ORBVocabulary&) = delete;
  virtual ~ORBVocabulary() { orbm_vocab_destroy(dev_); }

  /* System.cc:69-77 loads the text vocabulary through this name; the tree goes to the device on the
   * first transform. */
  bool loadFromTextFile(const std::string& filename) {
    Drop();
    return Base::loadFromTextFile(filename);
  }

  using Base::transform;
  virtual void transform(const std::vector<DBoW2::FORB::TDescriptor>& features, DBoW2::BowVector& v, DBoW2::FeatureVector& fv,
                         int levelsup) const {
    v.clear();
    fv.clear();
    if (this->empty()) return;
    const int n = (int)features.size();
    std::vector<uint8_t> desc((size_t)n * 32);
    for (int i = 0; i < n; ++i) memcpy(&desc[(size_t)i * 32], features[i].data, 32);
    std::vector<int32_t> word((size_t)n), node((size_t)n);
    std::vector<double> weight((size_t)n);
    if (n > 0) {
      const int rc = orbm_vocab_transform(Device(), desc.data(), n, levelsup, word.data(), node.data(), weight.data());
      if (rc != ORB_OK) throw std::runtime_error(std::string("ORBVocabulary::transform: ") + orb_last_error());
    }
    DBoW2::LNorm norm;
    const bool must = this->m_scoring_object->mustNormalize(norm);
    if (this->m_weighting == DBoW2::TF || this->m_weighting == DBoW2::TF_IDF) {
      for (int i = 0; i < n; ++i) {
        if (weight[i] > 0) {  // not stopped
          v.addWeight((DBoW2::WordId)word[i], weight[i]);
          fv.addFeature((DBoW2::NodeId)node[i], (unsigned int)i);
        }
      }
      if (!v.empty() && !must) {
        const double nd = v.size();
        for (DBoW2::BowVector::iterator vit = v.begin(); vit != v.end(); vit++) vit->second /= nd;
      }
    } else {  // IDF || BINARY
      for (int i = 0; i < n; ++i) {
        if (weight[i] > 0) {
          v.addIfNotExist((DBoW2::WordId)word[i], weight[i]);
          fv.addFeature((DBoW2::NodeId)node[i], (unsigned int)i);
        }
      }
    }
    if (must) v.normalize(norm);
  }

 private:
  mutable orbm_vocab* dev_;
  mutable std::mutex mutex_;  // ComputeBoW runs on the tracking and the local-mapping thread

  void Drop() {
    std::lock_guard<std::mutex> lock(mutex_);
    orbm_vocab_destroy(dev_);
    dev_ = NULL;
  }
  /* The tree as loadFromTextFile built it (m_nodes: parent links, children in file order; m_words:
   * the leaves in file order), flattened for orbm_vocab_create. */
  orbm_vocab* Device() const {
    std::lock_guard<std::mutex> lock(mutex_);
    if (dev_) return dev_;
    const int n_nodes = (int)this->m_nodes.size();
    std::vector<int32_t> parent((size_t)n_nodes, 0);
    std::vector<uint8_t> is_word((size_t)n_nodes, 0), desc((size_t)n_nodes * 32, 0);
    std::vector<double> weight((size_t)n_nodes, 0.0);
    for (int i = 1; i < n_nodes; ++i) {
      parent[i] = (int32_t)this->m_nodes[i].parent;
      weight[i] = this->m_nodes[i].weight;
      if (!this->m_nodes[i].descriptor.empty()) memcpy(&desc[(size_t)i * 32], this->m_nodes[i].descriptor.data, 32);
    }
    for (size_t w = 0; w < this->m_words.size(); ++w) is_word[this->m_words[w]->id] = 1;
    int device = 0;
    if (const char* e = getenv("ORBFRONT_DEVICE")) device = atoi(e);
    const int rc = orbm_vocab_create(device, this->m_k, this->m_L, n_nodes, parent.data(), is_word.data(), desc.data(), weight.data(),
                                     &dev_);
    if (rc != ORB_OK) throw std::runtime_error(std::string("ORBVocabulary: ") + orb_last_error());
    return dev_;
  }
};

}  // namespace ORB_SLAM2

#endif  // ORBVOCABULARY_H
