// orb_bow.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h).
//
// Restatement of the DBoW2 vocabulary-tree transform that Frame::ComputeBoW / KeyFrame::ComputeBoW call
// (reference: Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1262 transform, :1338-1424 loadFromTextFile;
// BowVector.cpp:34-84 addWeight / addIfNotExist / normalize; FeatureVector.cpp:31-45 addFeature;
// FORB.cpp:81-101 distance; ScoringObject.h:74-89 which scoring types normalise).  std::map keeps the reference's
// container semantics: weights of a word are accumulated by repeated addition in feature order, the norm is summed in
// ascending word id.
//
// Two places where the reference's behaviour is undefined and this restatement picks the defined reading:
//  * loadFromTextFile's `while(!f.eof()) getline(...)` makes one more pass over the empty string after the last line
//    and builds a node from uninitialised locals (pid, nIsLeaf); here a trailing empty line is ignored;
//  * transform() leaves *nid unset when the descent reaches a leaf above level L - levelsup; here nid = 0 (root).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "orb_internal.h"
#include "orb_oracle.h"

struct orc_vocabulary {
  int k = 0, L = 0, scoring = 0, weighting = 0;
  struct Node {
    double weight = 0;
    std::vector<int> children;
    int parent = 0;
    uint8_t desc[32] = {0};
    unsigned word_id = 0;
  };
  std::vector<Node> nodes;
  int n_words = 0;
};


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

extern "C" {

orc_vocabulary* orc_vocabulary_create(int k, int L, int scoring, int weighting, int n_nodes, const int32_t* parent,
                                      const uint8_t* is_leaf, const uint8_t* desc, const double* weight) {
  if (n_nodes < 1) return nullptr;
  orc_vocabulary* v = new orc_vocabulary();
  v->k = k; v->L = L; v->scoring = scoring; v->weighting = weighting;
  v->nodes.resize(n_nodes);
  for (int nid = 1; nid < n_nodes; ++nid) {  // :1385-1418
    const int pid = parent[nid];
    if (pid < 0 || pid >= nid) { delete v; return nullptr; }
    v->nodes[nid].parent = pid;
    v->nodes[pid].children.push_back(nid);
    std::memcpy(v->nodes[nid].desc, desc + (size_t)nid * 32, 32);
    v->nodes[nid].weight = weight[nid];
    if (is_leaf[nid]) v->nodes[nid].word_id = (unsigned)v->n_words++;
  }
  return v;
}

// TemplatedVocabulary::loadFromTextFile (:1338-1424)
orc_vocabulary* orc_vocabulary_load_text(const char* path) {
  std::vector<char> buf;
  if (!read_file(path, buf)) return nullptr;
  int h[4];
  std::vector<int32_t> parent;
  std::vector<uint8_t> leaf, desc;
  std::vector<double> weight;
  parse_vocabulary_text(buf.data(), h, parent, leaf, desc, weight);
  if (h[0] < 0 || h[0] > 20 || h[1] < 1 || h[1] > 10 || h[2] < 0 || h[2] > 5 || h[3] < 0 || h[3] > 3) return nullptr;
  return orc_vocabulary_create(h[0], h[1], h[2], h[3], (int)parent.size(), parent.data(), leaf.data(), desc.data(),
                               weight.data());
}

void orc_vocabulary_destroy(orc_vocabulary* v) { delete v; }

void orc_vocabulary_info(const orc_vocabulary* v, int* k, int* L, int* scoring, int* weighting, int* n_nodes,
                         int* n_words) {
  *k = v->k; *L = v->L; *scoring = v->scoring; *weighting = v->weighting;
  *n_nodes = (int)v->nodes.size(); *n_words = v->n_words;
}

// transform(features, BowVector&, FeatureVector&, levelsup) (:1127-1196).  Outputs: the BowVector as ascending
// (id, value) pairs, the FeatureVector as CSR (ascending node ids), and per feature the word / node / weight of the
// single-feature transform (:1216-1262).  Returns 0, or -1 when a capacity is too small.
int orc_vocabulary_transform(const orc_vocabulary* v, const uint8_t* desc, int n, int levelsup, uint32_t* bow_ids,
                             double* bow_vals, int bow_cap, int32_t* n_bow, uint32_t* fv_nodes, int32_t* fv_offsets,
                             int32_t* fv_features, int fv_cap, int32_t* n_fv, uint32_t* word_of, uint32_t* node_of,
                             double* weight_of) {
  *n_bow = 0;
  *n_fv = 0;
  if (fv_cap >= 0 && fv_offsets) fv_offsets[0] = 0;
  if (v->nodes.size() <= 1) return 0;  // empty()
  const bool must = v->scoring != 5;   // ScoringObject.h:74-89: all but DOT_PRODUCT normalise
  const bool l1 = v->scoring != 1;     // L2 only for L2_NORM
  std::map<unsigned, double> bow;
  std::map<unsigned, std::vector<unsigned>> fv;
  const int nid_level = v->L - levelsup;
  for (int i = 0; i < n; ++i) {
    const uint8_t* d = desc + (size_t)i * 32;
    unsigned nid = 0, final_id = 0;  // root
    int current_level = 0;
    do {
      ++current_level;
      const std::vector<int>& ch = v->nodes[final_id].children;
      final_id = ch[0];
      double best_d = orc::descriptor_distance(d, v->nodes[final_id].desc);
      for (size_t c = 1; c < ch.size(); ++c) {
        const double dd = orc::descriptor_distance(d, v->nodes[ch[c]].desc);
        if (dd < best_d) { best_d = dd; final_id = ch[c]; }
      }
      if (current_level == nid_level) nid = final_id;
    } while (!v->nodes[final_id].children.empty());
    const unsigned id = v->nodes[final_id].word_id;
    const double w = v->nodes[final_id].weight;
    if (word_of) word_of[i] = id;
    if (node_of) node_of[i] = nid;
    if (weight_of) weight_of[i] = w;
    if (w > 0) {  // not stopped
      if (v->weighting == 0 || v->weighting == 1) {  // TF_IDF, TF: addWeight
        auto it = bow.lower_bound(id);
        if (it != bow.end() && it->first == id) it->second += w; else bow.insert(it, std::make_pair(id, w));
      } else {  // IDF, BINARY: addIfNotExist
        if (!bow.count(id)) bow[id] = w;
      }
      fv[nid].push_back((unsigned)i);
    }
  }
  if ((v->weighting == 0 || v->weighting == 1) && !bow.empty() && !must) {
    const double nd = (double)bow.size();
    for (auto& kv : bow) kv.second /= nd;
  }
  if (must) {  // BowVector::normalize
    double norm = 0.0;
    if (l1) { for (auto& kv : bow) norm += std::fabs(kv.second); }
    else { for (auto& kv : bow) norm += kv.second * kv.second; norm = std::sqrt(norm); }
    if (norm > 0.0) for (auto& kv : bow) kv.second /= norm;
  }
  if ((int)bow.size() > bow_cap || (int)fv.size() > fv_cap) return -1;
  int j = 0;
  for (auto& kv : bow) { bow_ids[j] = kv.first; bow_vals[j] = kv.second; ++j; }
  *n_bow = j;
  j = 0;
  int off = 0;
  for (auto& kv : fv) {
    fv_nodes[j] = kv.first;
    for (unsigned f : kv.second) fv_features[off++] = (int32_t)f;
    fv_offsets[++j] = off;
  }
  *n_fv = j;
  return 0;
}

}  // extern "C"
