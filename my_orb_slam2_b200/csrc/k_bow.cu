// k_bow.cu — DBoW2 vocabulary-tree transform (reference:
// Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1127-1262, called from Frame::ComputeBoW src/Frame.cc:420-427 and
// KeyFrame::ComputeBoW src/KeyFrame.cc:59-68): every descriptor walks the tree to its word, the words become the
// L1/L2-normalised BowVector and the nodes `levelsup` levels above the leaves the FeatureVector that SearchByBoW and
// SearchForTriangulation consume.
//
//  * bow_descend_kernel: one warp per descriptor; at each level the lanes take one child each (8 POPC), the warp
//    takes the minimum of (distance << 16 | child position) — the first child wins ties like the reference's strict
//    '<' (:1240-1249).
//  * bow_build_kernel: one CTA per descriptor set; the two std::map containers become two bitonic sorts of
//    (word << 32 | feature) resp. (node << 32 | feature) in shared memory.  Weights of a word are accumulated by
//    REPEATED addition and the norm is summed in ascending word id by one thread, because double addition is not
//    associative and BowVector::addWeight / normalize (BowVector.cpp:34-84) fix exactly that order.
#include "search_common.cuh"

namespace orbfe {

__global__ void __launch_bounds__(256) bow_descend_kernel(const DevVocab v, const uint8_t* __restrict__ desc, int n,
                                                          int nid_level, uint32_t* __restrict__ word_of,
                                                          uint32_t* __restrict__ node_of,
                                                          double* __restrict__ weight_of) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= n) return;
  const uint4* q = reinterpret_cast<const uint4*>(desc + (size_t)i * 32);
  const uint4 q0 = __ldg(q), q1 = __ldg(q + 1);
  int final_id = 0, nid = 0, level = 0;
  int cs = v.child_start[0], ce = v.child_start[1];
  do {
    ++level;
    unsigned best = 0xFFFFFFFFu;
    for (int c = cs + lane; c < ce; c += 32) {
      const int id = v.child_list[c];
      const uint4* p = reinterpret_cast<const uint4*>(v.desc + (size_t)id * 32);
      const uint4 a0 = __ldg(p), a1 = __ldg(p + 1);
      const int d = __popc(a0.x ^ q0.x) + __popc(a0.y ^ q0.y) + __popc(a0.z ^ q0.z) + __popc(a0.w ^ q0.w) +
                    __popc(a1.x ^ q1.x) + __popc(a1.y ^ q1.y) + __popc(a1.z ^ q1.z) + __popc(a1.w ^ q1.w);
      best = min(best, ((unsigned)d << 16) | (unsigned)(c - cs));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    final_id = v.child_list[cs + (int)(best & 0xFFFFu)];
    if (level == nid_level) nid = final_id;
    cs = v.child_start[final_id];
    ce = v.child_start[final_id + 1];
  } while (ce > cs);  // !isLeaf()
  if (lane == 0) {
    word_of[i] = v.word_id[final_id];
    node_of[i] = (uint32_t)nid;
    weight_of[i] = v.weight[final_id];
  }
}

__device__ void bitonic_sort_u64(unsigned long long* keys, int P) {
  const int tid = threadIdx.x, T = blockDim.x;
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < P; i += T) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = keys[i], b = keys[ixj];
          const bool up = (i & k) == 0;
          if ((a > b) == up) { keys[i] = b; keys[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
}

constexpr unsigned long long KEY_STOPPED = ~0ull;

// out_counts: [0] = BowVector size, [1] = FeatureVector nodes
__global__ void __launch_bounds__(1024) bow_build_kernel(int n, int P, const uint32_t* __restrict__ word_of,
                                                         const uint32_t* __restrict__ node_of,
                                                         const double* __restrict__ weight_of, int weighting, int must,
                                                         int l1, uint32_t* bow_ids, double* bow_vals, uint32_t* fv_nodes,
                                                         int32_t* fv_offsets, int32_t* fv_features,
                                                         int32_t* out_counts) {
  extern __shared__ unsigned long long s_keys[];
  int32_t* s_flag = reinterpret_cast<int32_t*>(s_keys + P);
  __shared__ int wtmp[32];
  __shared__ int s_nvalid;
  const int tid = threadIdx.x, T = blockDim.x;
  if (tid == 0) s_nvalid = 0;
  __syncthreads();
  const bool tf = weighting == 0 || weighting == 1;  // TF_IDF, TF accumulate; IDF, BINARY keep the first weight
  // ---- BowVector: std::map<WordId, WordValue> ----
  for (int i = tid; i < P; i += T) {
    const bool valid = i < n && weight_of[i] > 0.0;  // "not stopped" (:1157)
    s_keys[i] = valid ? (((unsigned long long)word_of[i] << 32) | (unsigned)i) : KEY_STOPPED;
    if (valid) atomicAdd(&s_nvalid, 1);
  }
  __syncthreads();
  bitonic_sort_u64(s_keys, P);
  for (int i = tid; i < P; i += T) {
    const unsigned long long k = s_keys[i];
    s_flag[i] = (k != KEY_STOPPED && (i == 0 || (s_keys[i - 1] >> 32) != (k >> 32))) ? 1 : 0;
  }
  __syncthreads();
  int n_words = 0;
  for (int base = 0; base < P; base += 8192) {  // cta_excl_scan handles 8 items per thread
    const int len = min(8192, P - base);
    const unsigned long long* keys = s_keys + base;
    int32_t* flag = s_flag + base;
    // heads were marked before the scan overwrites the flags: remember them in a register per item
    bool head[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { const int i = tid + r * T; head[r] = i < len && flag[i] != 0; }
    __syncthreads();
    const int total = cta_excl_scan(flag, len, wtmp);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int i = tid + r * T;
      if (!head[r]) continue;
      const unsigned word = (unsigned)(keys[i] >> 32);
      const double w = weight_of[(unsigned)(keys[i] & 0xFFFFFFFFu)];
      double val = w;
      if (tf)
        for (int j = base + i + 1; j < P && (unsigned)(s_keys[j] >> 32) == word && s_keys[j] != KEY_STOPPED; ++j)
          val = __dadd_rn(val, w);  // vit->second += v, once per further occurrence
      const int slot = n_words + flag[i];
      bow_ids[slot] = word;
      bow_vals[slot] = val;
    }
    n_words += total;
    __syncthreads();
  }
  if (tid == 0) {
    if (tf && n_words > 0 && !must) {  // :1166-1172
      const double nd = (double)n_words;
      for (int j = 0; j < n_words; ++j) bow_vals[j] = __ddiv_rn(bow_vals[j], nd);
    }
    if (must) {  // BowVector::normalize, in ascending word id
      double norm = 0.0;
      if (l1) {
        for (int j = 0; j < n_words; ++j) norm = __dadd_rn(norm, fabs(bow_vals[j]));
      } else {
        for (int j = 0; j < n_words; ++j) norm = __dadd_rn(norm, __dmul_rn(bow_vals[j], bow_vals[j]));
        norm = __dsqrt_rn(norm);
      }
      wtmp[0] = norm > 0.0 ? 1 : 0;
      reinterpret_cast<double*>(s_flag)[0] = norm;  // flags are free now; P >= 32 keeps this in bounds
    } else {
      wtmp[0] = 0;
    }
    out_counts[0] = n_words;
  }
  __syncthreads();
  if (wtmp[0]) {
    const double norm = reinterpret_cast<double*>(s_flag)[0];
    for (int j = tid; j < n_words; j += T) bow_vals[j] = __ddiv_rn(bow_vals[j], norm);
  }
  __syncthreads();
  // ---- FeatureVector: std::map<NodeId, std::vector<unsigned>> ----
  for (int i = tid; i < P; i += T)
    s_keys[i] = (i < n && weight_of[i] > 0.0) ? (((unsigned long long)node_of[i] << 32) | (unsigned)i) : KEY_STOPPED;
  __syncthreads();
  bitonic_sort_u64(s_keys, P);
  for (int i = tid; i < P; i += T) {
    const unsigned long long k = s_keys[i];
    s_flag[i] = (k != KEY_STOPPED && (i == 0 || (s_keys[i - 1] >> 32) != (k >> 32))) ? 1 : 0;
    if (k != KEY_STOPPED) fv_features[i] = (int32_t)(k & 0xFFFFFFFFu);  // valid keys sort first: position = CSR offset
  }
  __syncthreads();
  int n_nodes = 0;
  for (int base = 0; base < P; base += 8192) {
    const int len = min(8192, P - base);
    const unsigned long long* keys = s_keys + base;
    int32_t* flag = s_flag + base;
    bool head[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) { const int i = tid + r * T; head[r] = i < len && flag[i] != 0; }
    __syncthreads();
    const int total = cta_excl_scan(flag, len, wtmp);
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const int i = tid + r * T;
      if (!head[r]) continue;
      const int slot = n_nodes + flag[i];
      fv_nodes[slot] = (uint32_t)(keys[i] >> 32);
      fv_offsets[slot] = base + i;
    }
    n_nodes += total;
    __syncthreads();
  }
  if (tid == 0) {
    fv_offsets[n_nodes] = s_nvalid;
    out_counts[1] = n_nodes;
  }
}

void launch_bow_descend(const DevVocab& v, const uint8_t* desc, int n, int nid_level, uint32_t* word_of,
                        uint32_t* node_of, double* weight_of, cudaStream_t s) {
  if (n > 0) bow_descend_kernel<<<(n + 7) / 8, 256, 0, s>>>(v, desc, n, nid_level, word_of, node_of, weight_of);
}

int bow_build_capacity() { return 16384; }

int launch_bow_build(int n, const uint32_t* word_of, const uint32_t* node_of, const double* weight_of, int weighting,
                     int must, int l1, uint32_t* bow_ids, double* bow_vals, uint32_t* fv_nodes, int32_t* fv_offsets,
                     int32_t* fv_features, int32_t* out_counts, cudaStream_t s) {
  int P = 32;
  while (P < n) P <<= 1;
  if (P > bow_build_capacity()) return -1;
  const size_t smem = (size_t)P * 12;
  if (smem > 40 * 1024 &&
      cudaFuncSetAttribute(bow_build_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384 * 12) != cudaSuccess)
    return -1;
  bow_build_kernel<<<1, 1024, smem, s>>>(n, P, word_of, node_of, weight_of, weighting, must, l1, bow_ids, bow_vals,
                                         fv_nodes, fv_offsets, fv_features, out_counts);
  return 0;
}

}  // namespace orbfe
