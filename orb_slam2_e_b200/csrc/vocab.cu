// Vocabulary-tree descent (DBoW2 TemplatedVocabulary<FORB>::transform):
// Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1217-1259 -- every descriptor walks from the root to
// a leaf, taking at each node the FIRST child with the smallest Hamming distance (`d < best_d`),
// and reports the word at the leaf, its weight, and the node it passed `levelsup` levels above the
// leaves (the key of DBoW2::FeatureVector that SearchByBoW groups by).
//
// The tree (ORBvoc: k = 10, L = 6, 1.1 M nodes, 35 MB of descriptors) stays resident in HBM.  One
// warp per descriptor: lane c scores child c (k <= 32), a warp min over `distance << 8 | c` picks
// the child; L dependent rounds per descriptor, thousands of descriptors in flight.
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"

struct orbm_vocab {
  int device = 0, k = 0, L = 0, n_nodes = 0, n_words = 0;
  uint4* desc = nullptr;      // [n_nodes][2]
  int* child_off = nullptr;   // [n_nodes + 1]
  int* child_idx = nullptr;   // children in the reference's order (ascending node id = file order)
  int* word_id = nullptr;     // [n_nodes] (-1: not a word)
  std::vector<double> weight; // host copy, indexed by node
  std::vector<int> host_word; // host copy of word_id
};

namespace orbfront {
namespace {

__global__ void __launch_bounds__(128) vocab_descend_kernel(const uint4* __restrict__ node_desc, const int* __restrict__ child_off,
                                                            const int* __restrict__ child_idx, const uint4* __restrict__ feat, int n,
                                                            int nid_level, int* __restrict__ leaf, int* __restrict__ nid) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= n) return;
  const uint4 qa = __ldg(&feat[2 * i]), qb = __ldg(&feat[2 * i + 1]);
  int node = 0, level = 0, at_level = 0;  // TemplatedVocabulary.h:1226-1227: nid_level <= 0 -> root
  for (;;) {
    const int b = __ldg(&child_off[node]), e = __ldg(&child_off[node + 1]);
    if (b == e) break;  // leaf (:1254)
    ++level;
    uint32_t key = 0xffffffffu;
    if (lane < e - b) {
      const int c = __ldg(&child_idx[b + lane]);
      const uint4 a = __ldg(&node_desc[2 * c]), d = __ldg(&node_desc[2 * c + 1]);
      const int dist = __popc(qa.x ^ a.x) + __popc(qa.y ^ a.y) + __popc(qa.z ^ a.z) + __popc(qa.w ^ a.w) + __popc(qb.x ^ d.x) +
                       __popc(qb.y ^ d.y) + __popc(qb.z ^ d.z) + __popc(qb.w ^ d.w);
      key = ((uint32_t)dist << 8) | (uint32_t)lane;
    }
    for (int s = 16; s > 0; s >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, s));
    node = __ldg(&child_idx[b + (int)(key & 0xffu)]);
    if (level == nid_level) at_level = node;  // :1251-1252
  }
  if (lane == 0) { leaf[i] = node; nid[i] = at_level; }
}

}  // namespace
}  // namespace orbfront

using namespace orbfront;

extern "C" {

int orbm_vocab_create(int device, int k, int L, int n_nodes, const int32_t* parent, const uint8_t* is_word, const uint8_t* desc,
                      const double* weight, orbm_vocab** out) {
  if (!out || k < 2 || k > 32 || L < 1 || n_nodes < 1 || !parent || !is_word || !desc || !weight)
    return set_error(ORB_E_INVALID, "orbm_vocab_create: bad argument (k must be 2..32)");
  for (int i = 1; i < n_nodes; ++i)
    if (parent[i] < 0 || parent[i] >= i) return set_error(ORB_E_INVALID, "orbm_vocab_create: node %d has parent %d (parents precede children)", i, parent[i]);
  int rc = use_device(device);
  if (rc) return rc;
  // children in ascending node id, as loadFromTextFile appends them (TemplatedVocabulary.h:1377-1391)
  std::vector<int> off((size_t)n_nodes + 1, 0), idx((size_t)std::max(n_nodes - 1, 0)), word((size_t)n_nodes, -1);
  for (int i = 1; i < n_nodes; ++i) ++off[(size_t)parent[i] + 1];
  for (int i = 0; i < n_nodes; ++i) {
    if (off[(size_t)i + 1] > k) return set_error(ORB_E_INVALID, "orbm_vocab_create: node %d has %d children, k = %d", i, off[(size_t)i + 1], k);
    off[(size_t)i + 1] += off[i];
  }
  std::vector<int> fill(off.begin(), off.end() - 1);
  for (int i = 1; i < n_nodes; ++i) idx[(size_t)fill[parent[i]]++] = i;
  int n_words = 0;
  for (int i = 1; i < n_nodes; ++i)
    if (is_word[i]) word[i] = n_words++;  // :1407-1414: word ids in file order
  orbm_vocab* V = new orbm_vocab();
  V->device = device; V->k = k; V->L = L; V->n_nodes = n_nodes; V->n_words = n_words;
  V->weight.assign(weight, weight + n_nodes);
  V->host_word = word;
  cudaError_t e = cudaMalloc(&V->desc, (size_t)n_nodes * 32);
  if (e == cudaSuccess) e = cudaMalloc(&V->child_off, ((size_t)n_nodes + 1) * 4);
  if (e == cudaSuccess) e = cudaMalloc(&V->child_idx, std::max<size_t>(idx.size(), 1) * 4);
  if (e == cudaSuccess) e = cudaMalloc(&V->word_id, (size_t)n_nodes * 4);
  if (e == cudaSuccess) e = cudaMemcpy(V->desc, desc, (size_t)n_nodes * 32, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(V->child_off, off.data(), off.size() * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess && !idx.empty()) e = cudaMemcpy(V->child_idx, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(V->word_id, word.data(), word.size() * 4, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    orbm_vocab_destroy(V);
    return set_error(ORB_E_CUDA, "orbm_vocab_create: %s", cudaGetErrorString(e));
  }
  *out = V;
  return ORB_OK;
}

int orbm_vocab_load_text(int device, const char* path, orbm_vocab** out) {
  if (!path || !out) return set_error(ORB_E_INVALID, "orbm_vocab_load_text: bad argument");
  FILE* f = fopen(path, "r");
  if (!f) return set_error(ORB_E_INVALID, "orbm_vocab_load_text: cannot open %s", path);
  int k = 0, L = 0, n1 = 0, n2 = 0;
  if (fscanf(f, "%d %d %d %d", &k, &L, &n1, &n2) != 4 || k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) {
    fclose(f);
    return set_error(ORB_E_INVALID, "orbm_vocab_load_text: %s is not a vocabulary text file", path);  // :1359-1363
  }
  std::vector<int32_t> parent(1, 0);
  std::vector<uint8_t> is_word(1, 0), desc(32, 0);
  std::vector<double> weight(1, 0.0);
  for (;;) {  // one node per line: parent, leaf flag, 32 descriptor bytes, weight (:1375-1419)
    int pid, leaf;
    if (fscanf(f, "%d %d", &pid, &leaf) != 2) break;
    uint8_t d[32];
    bool ok = true;
    for (int i = 0; i < 32; ++i) {
      int v;
      if (fscanf(f, "%d", &v) != 1) { ok = false; break; }
      d[i] = (uint8_t)v;
    }
    double w;
    if (!ok || fscanf(f, "%lf", &w) != 1) break;
    parent.push_back(pid);
    is_word.push_back(leaf > 0);
    desc.insert(desc.end(), d, d + 32);
    weight.push_back(w);
  }
  fclose(f);
  return orbm_vocab_create(device, k, L, (int)parent.size(), parent.data(), is_word.data(), desc.data(), weight.data(), out);
}

int orbm_vocab_info(const orbm_vocab* V, int* k, int* L, int* n_nodes, int* n_words) {
  if (!V) return set_error(ORB_E_INVALID, "orbm_vocab_info: null");
  if (k) *k = V->k;
  if (L) *L = V->L;
  if (n_nodes) *n_nodes = V->n_nodes;
  if (n_words) *n_words = V->n_words;
  return ORB_OK;
}

int orbm_vocab_destroy(orbm_vocab* V) {
  if (!V) return ORB_OK;
  cudaSetDevice(V->device);
  if (V->desc) cudaFree(V->desc);
  if (V->child_off) cudaFree(V->child_off);
  if (V->child_idx) cudaFree(V->child_idx);
  if (V->word_id) cudaFree(V->word_id);
  delete V;
  return ORB_OK;
}

int orbm_vocab_transform(orbm_vocab* V, const uint8_t* desc, int n, int levelsup, int32_t* word_id, int32_t* node_id,
                         double* weight) {
  if (!V || n < 0 || (n > 0 && (!desc || !word_id || !node_id || !weight)))
    return set_error(ORB_E_INVALID, "orbm_vocab_transform: bad argument");
  if (n == 0) return ORB_OK;
  int rc = use_device(V->device);
  if (rc) return rc;
  Layout L;
  const size_t o_d = L.place((size_t)n * 32), o_res = L.place((size_t)n * 8);
  CallScratch* S = nullptr;
  if ((rc = call_scratch(V->device, L.bytes, &S))) return rc;
  memcpy(S->host + o_d, desc, (size_t)n * 32);
  ORB_CUDA_TRY(cudaMemcpyAsync(S->dev + o_d, S->host + o_d, (size_t)n * 32, cudaMemcpyHostToDevice, S->stream));
  int* leaf = (int*)(S->dev + o_res);
  vocab_descend_kernel<<<div_up(n, 4), 128, 0, S->stream>>>(V->desc, V->child_off, V->child_idx, (const uint4*)(S->dev + o_d), n,
                                                           V->L - levelsup, leaf, leaf + n);
  ORB_LAUNCH_CHECK();
  ORB_CUDA_TRY(cudaMemcpyAsync(S->host + o_res, S->dev + o_res, (size_t)n * 8, cudaMemcpyDeviceToHost, S->stream));
  ORB_CUDA_TRY(cudaStreamSynchronize(S->stream));
  const int* hl = reinterpret_cast<const int*>(S->host + o_res);
  for (int i = 0; i < n; ++i) {
    word_id[i] = V->host_word[hl[i]];  // :1256-1258 (a childless root has word 0 / weight 0 in the reference)
    if (word_id[i] < 0) word_id[i] = 0;
    weight[i] = V->weight[hl[i]];
    node_id[i] = hl[n + i];
  }
  return ORB_OK;
}

}  // extern "C"
