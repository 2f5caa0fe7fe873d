// bow.cu — host side of the DBoW2 vocabulary entry points of include/orbfe.h: the tree lives on
// the device once (ORBvoc: 1.1 M nodes x 32 B = 36 MB of the 180 GB), a transform uploads the descriptors (or takes
// the extractor's device-resident ones), runs k_bow.cu and brings the two small containers back.  No result is
// computed on the host.  One vocabulary is shared by the Tracking and LocalMapping threads in the reference
// (Frame::ComputeBoW, KeyFrame::ComputeBoW), so a transform holds the handle's mutex.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"
#include "match_kernels.h"
#include "../../include/orbfe.h"

namespace orbfe {

struct Vocabulary {
  int device = 0;
  int k = 0, L = 0, scoring = 0, weighting = 0, n_nodes = 0, n_words = 0;
  cudaStream_t stream = nullptr;
  DevVocab dv{};
  void* d_tree = nullptr;   // one allocation behind dv
  uint8_t* d_work = nullptr;  // per-call workspace
  size_t work_cap = 0;
  int32_t* h_counts = nullptr;  // pinned
  std::mutex mu;
};

static int ensure_work(Vocabulary* v, size_t bytes) {
  if (bytes <= v->work_cap) return ORBFE_OK;
  if (v->d_work) cudaFree(v->d_work);
  v->d_work = nullptr; v->work_cap = 0;
  ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&v->d_work), bytes + bytes / 2));
  v->work_cap = bytes + bytes / 2;
  return ORBFE_OK;
}


// whole file into memory; C stdio only (the process that loads this library may carry another libstdc++)
static bool read_file(const char* path, std::vector<char>& buf) {
  FILE* f = std::fopen(path, "rb");
  if (!f) return false;
  std::fseek(f, 0, SEEK_END);
  const long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  buf.resize((size_t)(sz > 0 ? sz : 0) + 1);
  const size_t got = sz > 0 ? std::fread(buf.data(), 1, (size_t)sz, f) : 0;
  std::fclose(f);
  buf[got] = 0;
  return true;
}

// "k L scoring weighting" then one line per node "parent isLeaf d0 .. d31 weight"; blank lines are skipped
static bool parse_vocabulary_text(char* text, int hdr[4], std::vector<int32_t>& parent, std::vector<uint8_t>& leaf,
                                  std::vector<uint8_t>& desc, std::vector<double>& weight) {
  char* p = text;
  char* line_end = std::strchr(p, '\n');
  if (line_end) *line_end = 0;
  hdr[0] = hdr[1] = hdr[2] = hdr[3] = -1;
  {
    char* e = p;
    for (int i = 0; i < 4; ++i) {
      const long v = std::strtol(p, &e, 10);
      if (e == p) break;
      hdr[i] = (int)v;
      p = e;
    }
  }
  parent.assign(1, 0);
  leaf.assign(1, 0);
  desc.assign(32, 0);
  weight.assign(1, 0.0);
  if (!line_end) return true;
  p = line_end + 1;
  while (*p) {
    char* nl = std::strchr(p, '\n');
    if (nl) *nl = 0;
    char* e = p;
    const long pid = std::strtol(p, &e, 10);
    if (e != p) {  // not a blank line
      p = e;
      const long is_leaf = std::strtol(p, &e, 10); p = e;
      parent.push_back((int32_t)pid);
      leaf.push_back(is_leaf > 0);
      for (int i = 0; i < 32; ++i) { desc.push_back((uint8_t)std::strtol(p, &e, 10)); p = e; }  // FORB::fromString
      weight.push_back(std::strtod(p, &e));
    }
    if (!nl) break;
    p = nl + 1;
  }
  return true;
}

}  // namespace orbfe

using namespace orbfe;

extern "C" {

int orbfe_vocabulary_create(int device, int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent,
                            const uint8_t* is_leaf, const uint8_t* desc, const double* weight, orbfe_vocabulary** out) {
  ORBFE_ARG(out);
  *out = nullptr;
  ORBFE_ARG(n_nodes >= 1 && parent && is_leaf && desc && weight);
  ORBFE_ARG(scoring >= 0 && scoring <= 5 && weighting >= 0 && weighting <= 3 && L >= 1);
  int ndev = 0;
  ORBFE_CUDA(cudaGetDeviceCount(&ndev));
  ORBFE_ARG(device >= 0 && device < ndev);
  // children in push_back order = ascending node id (TemplatedVocabulary.h:1385-1392)
  std::vector<int32_t> start((size_t)n_nodes + 1, 0), list((size_t)std::max(n_nodes - 1, 1), 0);
  std::vector<uint32_t> word((size_t)n_nodes, 0);
  int n_words = 0;
  for (int i = 1; i < n_nodes; ++i) {
    ORBFE_ARG(parent[i] >= 0 && parent[i] < i);
    start[(size_t)parent[i] + 1]++;
    if (is_leaf[i]) word[i] = (uint32_t)n_words++;
  }
  for (int i = 0; i < n_nodes; ++i) start[i + 1] += start[i];
  {
    std::vector<int32_t> fill(start.begin(), start.end() - 1);
    for (int i = 1; i < n_nodes; ++i) list[fill[parent[i]]++] = i;
  }
  Vocabulary* v = new (std::nothrow) Vocabulary();
  ORBFE_ARG(v != nullptr);
  v->device = device; v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
  v->n_nodes = n_nodes; v->n_words = n_words;
  DeviceGuard guard(device);
  const size_t N = (size_t)n_nodes;
  const size_t o_start = 0, o_list = align_up(o_start + (N + 1) * 4, 256), o_desc = align_up(o_list + N * 4, 256),
               o_word = align_up(o_desc + N * 32, 256), o_weight = align_up(o_word + N * 4, 256),
               total = o_weight + N * 8;
  if (cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaMalloc(&v->d_tree, total) != cudaSuccess ||
      cudaHostAlloc(reinterpret_cast<void**>(&v->h_counts), 64, cudaHostAllocDefault) != cudaSuccess) {
    set_error("vocabulary: device allocation of %zu bytes failed", total);
    orbfe_vocabulary_destroy(reinterpret_cast<orbfe_vocabulary*>(v));
    return ORBFE_ERR_CUDA;
  }
  uint8_t* base = static_cast<uint8_t*>(v->d_tree);
  cudaMemcpyAsync(base + o_start, start.data(), (N + 1) * 4, cudaMemcpyHostToDevice, v->stream);
  cudaMemcpyAsync(base + o_list, list.data(), (N > 1 ? N - 1 : 1) * 4, cudaMemcpyHostToDevice, v->stream);
  cudaMemcpyAsync(base + o_desc, desc, N * 32, cudaMemcpyHostToDevice, v->stream);
  cudaMemcpyAsync(base + o_word, word.data(), N * 4, cudaMemcpyHostToDevice, v->stream);
  cudaMemcpyAsync(base + o_weight, weight, N * 8, cudaMemcpyHostToDevice, v->stream);
  if (cudaStreamSynchronize(v->stream) != cudaSuccess) {
    set_error("vocabulary: upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    orbfe_vocabulary_destroy(reinterpret_cast<orbfe_vocabulary*>(v));
    return ORBFE_ERR_CUDA;
  }
  v->dv.n_nodes = n_nodes;
  v->dv.child_start = reinterpret_cast<const int32_t*>(base + o_start);
  v->dv.child_list = reinterpret_cast<const int32_t*>(base + o_list);
  v->dv.desc = base + o_desc;
  v->dv.word_id = reinterpret_cast<const uint32_t*>(base + o_word);
  v->dv.weight = reinterpret_cast<const double*>(base + o_weight);
  *out = reinterpret_cast<orbfe_vocabulary*>(v);
  return ORBFE_OK;
}

// TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1338-1424): "k L scoring weighting" then one line per
// node "parent isLeaf d0 .. d31 weight"; node ids are line numbers, word ids count the leaf lines.  A trailing empty
// line is ignored (the reference builds one more node from uninitialised locals there).
int orbfe_vocabulary_load_text(int device, const char* path, orbfe_vocabulary** out) {
  ORBFE_ARG(out && path);
  *out = nullptr;
  std::vector<char> buf;
  if (!read_file(path, buf)) { set_error("vocabulary: cannot open %s", path); return ORBFE_ERR_ARG; }
  int h[4];
  std::vector<int32_t> parent;
  std::vector<uint8_t> leaf, desc;
  std::vector<double> weight;
  parse_vocabulary_text(buf.data(), h, parent, leaf, desc, weight);
  if (h[0] < 0 || h[0] > 20 || h[1] < 1 || h[1] > 10 || h[2] < 0 || h[2] > 5 || h[3] < 0 || h[3] > 3) {  // :1359-1363
    set_error("vocabulary: %s is not a correct text file", path);
    return ORBFE_ERR_ARG;
  }
  return orbfe_vocabulary_create(device, h[0], h[1], h[2], h[3], (int)parent.size(), parent.data(), leaf.data(),
                                 desc.data(), weight.data(), out);
}

void orbfe_vocabulary_destroy(orbfe_vocabulary* h) {
  if (!h) return;
  Vocabulary* v = reinterpret_cast<Vocabulary*>(h);
  DeviceGuard guard(v->device);
  if (v->stream) cudaStreamSynchronize(v->stream);
  if (v->d_tree) cudaFree(v->d_tree);
  if (v->d_work) cudaFree(v->d_work);
  if (v->h_counts) cudaFreeHost(v->h_counts);
  if (v->stream) cudaStreamDestroy(v->stream);
  delete v;
}

int orbfe_vocabulary_info(const orbfe_vocabulary* h, int32_t* info6) {
  ORBFE_ARG(h && info6);
  const Vocabulary* v = reinterpret_cast<const Vocabulary*>(h);
  info6[0] = v->k; info6[1] = v->L; info6[2] = v->scoring; info6[3] = v->weighting;
  info6[4] = v->n_nodes; info6[5] = v->n_words;
  return ORBFE_OK;
}

static int transform_impl(Vocabulary* v, const uint8_t* desc, bool desc_on_device, int n, int levelsup,
                          uint32_t* bow_ids, double* bow_vals, int bow_cap, int32_t* n_bow, uint32_t* fv_nodes,
                          int32_t* fv_offsets, int32_t* fv_features, int fv_cap, int32_t* n_fv, uint32_t* word_of,
                          uint32_t* node_of, double* weight_of) {
  ORBFE_ARG(n >= 0 && n_bow && n_fv && bow_cap >= 0 && fv_cap >= 0);
  *n_bow = 0;
  *n_fv = 0;
  if (fv_offsets) fv_offsets[0] = 0;
  if (n == 0 || v->n_nodes <= 1) return ORBFE_OK;  // v.clear(); fv.clear(); if(empty()) return;
  ORBFE_ARG(desc && bow_ids && bow_vals && fv_nodes && fv_offsets && fv_features);
  if (n > bow_build_capacity()) {
    set_error("vocabulary transform: %d descriptors exceed the %d the shared-memory sort holds", n, bow_build_capacity());
    return ORBFE_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(v->mu);
  DeviceGuard guard(v->device);
  const size_t N = (size_t)n;
  const size_t o_desc = 0, o_word = align_up(o_desc + N * 32, 256), o_node = align_up(o_word + N * 4, 256),
               o_w = align_up(o_node + N * 4, 256), o_bid = align_up(o_w + N * 8, 256),
               o_bval = align_up(o_bid + N * 4, 256), o_fn = align_up(o_bval + N * 8, 256),
               o_fo = align_up(o_fn + N * 4, 256), o_ff = align_up(o_fo + (N + 1) * 4, 256),
               o_cnt = align_up(o_ff + N * 4, 256), total = o_cnt + 64;
  int rc = ensure_work(v, total);
  if (rc != ORBFE_OK) return rc;
  uint8_t* w = v->d_work;
  const uint8_t* d_desc = desc;
  if (!desc_on_device) {
    ORBFE_CUDA(cudaMemcpyAsync(w + o_desc, desc, N * 32, cudaMemcpyHostToDevice, v->stream));
    d_desc = w + o_desc;
  }
  uint32_t* d_word = reinterpret_cast<uint32_t*>(w + o_word);
  uint32_t* d_node = reinterpret_cast<uint32_t*>(w + o_node);
  double* d_w = reinterpret_cast<double*>(w + o_w);
  uint32_t* d_bid = reinterpret_cast<uint32_t*>(w + o_bid);
  double* d_bval = reinterpret_cast<double*>(w + o_bval);
  uint32_t* d_fn = reinterpret_cast<uint32_t*>(w + o_fn);
  int32_t* d_fo = reinterpret_cast<int32_t*>(w + o_fo);
  int32_t* d_ff = reinterpret_cast<int32_t*>(w + o_ff);
  int32_t* d_cnt = reinterpret_cast<int32_t*>(w + o_cnt);
  const int must = v->scoring != 5;  // ScoringObject.h:74-89: every scoring but DOT_PRODUCT normalises
  const int l1 = v->scoring != 1;    // ... with the L1 norm except L2_NORM
  launch_bow_descend(v->dv, d_desc, n, v->L - levelsup, d_word, d_node, d_w, v->stream);
  if (launch_bow_build(n, d_word, d_node, d_w, v->weighting, must, l1, d_bid, d_bval, d_fn, d_fo, d_ff, d_cnt,
                       v->stream) != 0) {
    set_error("vocabulary transform: cannot configure the build kernel for %d descriptors", n);
    return ORBFE_ERR_CUDA;
  }
  ORBFE_CUDA(cudaMemcpyAsync(v->h_counts, d_cnt, 8, cudaMemcpyDeviceToHost, v->stream));
  ORBFE_CUDA(cudaStreamSynchronize(v->stream));
  ORBFE_CUDA(cudaGetLastError());
  const int nb = v->h_counts[0], nf = v->h_counts[1];
  if (nb > bow_cap || nf > fv_cap) {
    set_error("vocabulary transform: capacities %d / %d too small for %d words / %d nodes", bow_cap, fv_cap, nb, nf);
    return ORBFE_ERR_ARG;
  }
  if (nb > 0) {
    ORBFE_CUDA(cudaMemcpyAsync(bow_ids, d_bid, (size_t)nb * 4, cudaMemcpyDeviceToHost, v->stream));
    ORBFE_CUDA(cudaMemcpyAsync(bow_vals, d_bval, (size_t)nb * 8, cudaMemcpyDeviceToHost, v->stream));
  }
  ORBFE_CUDA(cudaMemcpyAsync(fv_offsets, d_fo, ((size_t)nf + 1) * 4, cudaMemcpyDeviceToHost, v->stream));
  if (nf > 0) ORBFE_CUDA(cudaMemcpyAsync(fv_nodes, d_fn, (size_t)nf * 4, cudaMemcpyDeviceToHost, v->stream));
  if (word_of) ORBFE_CUDA(cudaMemcpyAsync(word_of, d_word, N * 4, cudaMemcpyDeviceToHost, v->stream));
  if (node_of) ORBFE_CUDA(cudaMemcpyAsync(node_of, d_node, N * 4, cudaMemcpyDeviceToHost, v->stream));
  if (weight_of) ORBFE_CUDA(cudaMemcpyAsync(weight_of, d_w, N * 8, cudaMemcpyDeviceToHost, v->stream));
  ORBFE_CUDA(cudaStreamSynchronize(v->stream));
  // the number of non-stopped features = fv_offsets[nf]
  const int n_valid = fv_offsets[nf];
  if (n_valid > 0) {
    ORBFE_CUDA(cudaMemcpyAsync(fv_features, d_ff, (size_t)n_valid * 4, cudaMemcpyDeviceToHost, v->stream));
    ORBFE_CUDA(cudaStreamSynchronize(v->stream));
  }
  *n_bow = nb;
  *n_fv = nf;
  return ORBFE_OK;
}

int orbfe_vocabulary_transform(orbfe_vocabulary* h, const uint8_t* desc, int n, int levelsup, uint32_t* bow_ids,
                               double* bow_vals, int bow_cap, int32_t* n_bow, uint32_t* fv_nodes, int32_t* fv_offsets,
                               int32_t* fv_features, int fv_cap, int32_t* n_fv, uint32_t* word_of, uint32_t* node_of,
                               double* weight_of) {
  ORBFE_ARG(h);
  return transform_impl(reinterpret_cast<Vocabulary*>(h), desc, false, n, levelsup, bow_ids, bow_vals, bow_cap, n_bow,
                        fv_nodes, fv_offsets, fv_features, fv_cap, n_fv, word_of, node_of, weight_of);
}

int orbfe_vocabulary_transform_device(orbfe_vocabulary* h, const uint8_t* d_desc, int n, int levelsup,
                                      uint32_t* bow_ids, double* bow_vals, int bow_cap, int32_t* n_bow,
                                      uint32_t* fv_nodes, int32_t* fv_offsets, int32_t* fv_features, int fv_cap,
                                      int32_t* n_fv) {
  ORBFE_ARG(h);
  return transform_impl(reinterpret_cast<Vocabulary*>(h), d_desc, true, n, levelsup, bow_ids, bow_vals, bow_cap, n_bow,
                        fv_nodes, fv_offsets, fv_features, fv_cap, n_fv, nullptr, nullptr, nullptr);
}

}  // extern "C"
