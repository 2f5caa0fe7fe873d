// k_quadtree.cu — DistributeOctTree + ExtractorNode::DivideNode (reference:
// src/ORBextractor.cc:539-765, :481-537), one CTA per (pyramid level, image).
//
// The reference keeps a std::list of nodes, each owning a vector of keypoints,
// and mutates it node by node.  Here the list is an array held in shared memory
// IN LIST ORDER (position == list position), and the keypoints never move: each
// one only remembers the position of the node that owns it (node_of[k]).  One
// round of the reference's loop becomes a few data-parallel steps:
//   A. every key of an expandable node (count > 1) votes for its quadrant ->
//      speculative child counts (shared-memory atomics, order independent);
//   B. the reference's visiting order is turned into a rank per expandable node:
//      phase 1 (:603-664) visits in list order; phase 2 (:678-740) visits by
//      (size, creation order) descending (sort at :686, reverse loop :687).
//      All expandable nodes of a round were created in the previous round, where
//      a later push_front means a smaller list position, so "creation order
//      descending" == "list position ascending" (the deterministic tie rule of
//      SURVEY.md B.1 — the reference itself ties by heap address);
//   C. an exclusive scan of the non-empty child counts in visiting order gives
//      every child its push_front sequence number, hence its new list position
//      (children: reverse push order at the front; survivors keep their order
//      behind them), and phase 2's early break (:732) is the first rank at which
//      the running size reaches N;
//   D. keys of processed nodes re-point to their child, others to the moved slot.
// Finally the best response per node is taken (first wins on ties, :746-762),
// where "first" is the reference's candidate order: cell row, cell column, then
// row-major inside the cell (:801-846) — computed from (x,y), since k_fast.cu
// appends cells in arbitrary order.
#include "common.cuh"
#include "extractor.h"
#include "kernels.h"

namespace orbfe {

constexpr int QT_T = 512;

struct QtSmem {
  int4* box[2];
  int* cnt[2];
  int* cc;      // [4C] child counts, then child slots
  int* nrank;   // [C] visiting rank of node (or -1)
  int* nne;     // [C] non-empty children by rank
  int* push;    // [C] exclusive scan of nne by rank
  int* surv;    // [C] survivor flag scan -> new slot of survivors
  int* elist;   // [C] expandable nodes in list order
  int* vis;     // [C] node by visiting rank
  unsigned long long* best;  // [C]
};

size_t quadtree_smem_bytes(int node_cap) { return (size_t)node_cap * 88 + 64; }

__device__ __forceinline__ int block_excl_scan(int* a, int n, int* wtmp) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int per = (n + QT_T - 1) / QT_T;
  const int beg = min(tid * per, n), end = min(beg + per, n);
  int s = 0;
  for (int i = beg; i < end; ++i) s += a[i];
  int v = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  if (lane == 31) wtmp[wid] = v;
  __syncthreads();
  if (wid == 0) {
    int w = lane < QT_T / 32 ? wtmp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    wtmp[lane] = w;
  }
  __syncthreads();
  int run = (wid > 0 ? wtmp[wid - 1] : 0) + v - s;
  const int total = wtmp[QT_T / 32 - 1];
  for (int i = beg; i < end; ++i) { const int t = a[i]; a[i] = run; run += t; }
  __syncthreads();
  return total;
}

__device__ __forceinline__ int quadrant(const int4 bx, int x, int y) {
  const int mx = bx.x + ((bx.y - bx.x + 1) >> 1);  // UL.x + ceil((UR.x-UL.x)/2) (:483)
  const int my = bx.z + ((bx.w - bx.z + 1) >> 1);
  return (x < mx ? 0 : 1) + (y < my ? 0 : 2);      // n1,n2,n3,n4 (:513-525)
}

// MINB = 1: as many registers as the kernel wants (55: two CTAs per SM), the fastest single CTA — single frames, where
// level 0's CTA is the critical path.  MINB = 3: 40 registers, three CTAs per SM — batches with more CTAs than two
// per SM, where the kernel is latency-bound and a third resident CTA pays (0.0583 -> 0.0556 ms per 64 images,
// +1.1 % frames/s; it costs a single pair 6 us).
template <int MINB>
__global__ void __launch_bounds__(QT_T, MINB) quadtree_kernel(const __grid_constant__ PyrGeom g,
                                                        uint32_t* cand, const uint8_t* __restrict__ cell_hi,
                                                        uint16_t* __restrict__ node_of, int32_t* cand_count,
                                                        uint32_t* __restrict__ kept, int32_t* __restrict__ kept_count,
                                                        int32_t* __restrict__ flags) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int wtmp[32];
  __shared__ int s_misc[4];  // 0: first rank reaching N (phase 2), 1: children with count > 1, 2: filtered candidates

  const int tid = threadIdx.x;
  // level-major CTA order (x = image is the fast index): the long level-0 CTAs of ALL images start in the first wave
  // and the short upper levels fill the slots they free; image-major order made two full-length waves
  const int lvl = blockIdx.y, b = blockIdx.x;
  const LevelGeom& L = g.lv[lvl];
  const int C = g.max_node_cap;
  const int N = L.quota;
  int n = cand_count[b * kMaxLevels + lvl];
  n = min(n, L.cand_cap);
  if (n <= 0 || L.n_cols <= 0) {
    if (tid == 0) kept_count[b * kMaxLevels + lvl] = 0;
    return;
  }
  uint32_t* ck = cand + (size_t)b * g.cand_total + L.cand_off;
  // ---- the per-cell threshold choice (:830-837): k_fast.cu appended every cell-local maximum with score >=
  // min(iniTh, minTh); a cell that has a maximum >= iniTh keeps only those, any other cell keeps those >= minTh.
  // Compacted in place, chunk by chunk (a chunk is read completely before anything is written into it). ----
  {
    const uint8_t* hi = cell_hi + (size_t)b * g.total_cells + L.cell_base;
    if (tid == 0) s_misc[2] = 0;
    __syncthreads();
    for (int base = 0; base < n; base += QT_T) {
      const int k = base + tid;
      uint32_t c = 0;
      bool keep = false;
      if (k < n) {
        c = ck[k];
        const int cell = ((cand_y(c) - 3) / L.h_cell) * L.n_cols + (cand_x(c) - 3) / L.w_cell;
        keep = cand_score(c) >= (hi[cell] ? g.ini_th : g.min_th);
      }
      __syncthreads();
      const unsigned mk = __ballot_sync(0xffffffffu, keep);
      int slot = 0;
      if ((tid & 31) == 0 && mk) slot = atomicAdd(&s_misc[2], __popc(mk));
      slot = __shfl_sync(0xffffffffu, slot, 0) + __popc(mk & ((1u << (tid & 31)) - 1u));
      if (keep) ck[slot] = c;
    }
    __syncthreads();
    n = s_misc[2];
    if (tid == 0) cand_count[b * kMaxLevels + lvl] = n;
    if (n <= 0) {
      if (tid == 0) kept_count[b * kMaxLevels + lvl] = 0;
      return;
    }
  }
  QtSmem sm;
  {
    unsigned char* p = smem_raw;
    sm.box[0] = reinterpret_cast<int4*>(p); p += (size_t)C * 16;
    sm.box[1] = reinterpret_cast<int4*>(p); p += (size_t)C * 16;
    sm.best = reinterpret_cast<unsigned long long*>(p); p += (size_t)C * 8;
    sm.cnt[0] = reinterpret_cast<int*>(p); p += (size_t)C * 4;
    sm.cnt[1] = reinterpret_cast<int*>(p); p += (size_t)C * 4;
    sm.cc = reinterpret_cast<int*>(p); p += (size_t)C * 16;
    sm.nrank = reinterpret_cast<int*>(p); p += (size_t)C * 4;
    sm.nne = reinterpret_cast<int*>(p); p += (size_t)C * 4;
    sm.push = reinterpret_cast<int*>(p); p += (size_t)C * 4;
    sm.surv = reinterpret_cast<int*>(p); p += (size_t)C * 4;
    sm.elist = reinterpret_cast<int*>(p); p += (size_t)C * 4;
    sm.vis = reinterpret_cast<int*>(p);
  }
  uint16_t* nk = node_of + (size_t)b * g.cand_total + L.cand_off;

  // ---- roots (:543-583) ----
  const int n_ini = L.n_ini;
  const float h_x = L.h_x;
  const int box_h = L.max_by - kMinBorder;
  for (int i = tid; i < n_ini; i += QT_T) {
    sm.box[1][i] = make_int4((int)__fmul_rn(h_x, (float)i), (int)__fmul_rn(h_x, (float)(i + 1)), 0, box_h);
    sm.cnt[1][i] = 0;
  }
  __syncthreads();
  for (int k = tid; k < n; k += QT_T) {
    int ri = (int)__fdiv_rn((float)cand_x(ck[k]), h_x);  // :569
    ri = min(ri, n_ini - 1);
    nk[k] = (uint16_t)ri;
    atomicAdd(&sm.cnt[1][ri], 1);
  }
  __syncthreads();
  for (int i = tid; i < n_ini; i += QT_T) sm.surv[i] = sm.cnt[1][i] > 0 ? 1 : 0;
  __syncthreads();
  ORBFE_DCHECK(n_ini <= C);
  int S = block_excl_scan(sm.surv, n_ini, wtmp);
  for (int i = tid; i < n_ini; i += QT_T)
    if (sm.cnt[1][i] > 0) { sm.box[0][sm.surv[i]] = sm.box[1][i]; sm.cnt[0][sm.surv[i]] = sm.cnt[1][i]; }
  __syncthreads();
  if (S != n_ini) {
    for (int k = tid; k < n; k += QT_T) nk[k] = (uint16_t)sm.surv[nk[k]];
    __syncthreads();
  }

  int cur = 0, phase = 1;
  bool overflow = false;
  while (true) {
    int4* box = sm.box[cur];
    int* cnt = sm.cnt[cur];
    ORBFE_DCHECK(S >= 1 && S <= C);
    // ---- A: speculative child counts ----
    for (int i = tid; i < 4 * S; i += QT_T) sm.cc[i] = 0;
    for (int i = tid; i < S; i += QT_T) { sm.elist[i] = cnt[i] > 1 ? 1 : 0; sm.nrank[i] = -1; }
    if (tid == 0) { s_misc[0] = 0x7fffffff; s_misc[1] = 0; }
    __syncthreads();
    for (int k = tid; k < n; k += QT_T) {
      const int nd = nk[k];
      ORBFE_DCHECK(nd < S);
      if (cnt[nd] > 1) {
        const uint32_t c = ck[k];
        atomicAdd(&sm.cc[nd * 4 + quadrant(box[nd], cand_x(c), cand_y(c))], 1);
      }
    }
    // ---- B: visiting order ----
    // elist currently holds flags; scan them into slots, then turn into the list of expandable nodes
    for (int i = tid; i < S; i += QT_T) sm.surv[i] = sm.elist[i];  // keep flags
    __syncthreads();
    const int E = block_excl_scan(sm.elist, S, wtmp);
    if (E == 0) break;  // nothing to split: size unchanged (:669)
    for (int i = tid; i < S; i += QT_T) sm.push[i] = sm.elist[i];  // slot of node i among expandables
    __syncthreads();
    for (int i = tid; i < S; i += QT_T)
      if (sm.surv[i]) sm.elist[sm.push[i]] = i;
    __syncthreads();
    for (int a = tid; a < E; a += QT_T) {
      const int i = sm.elist[a];
      int r = a;
      if (phase == 2) {
        const int ci = cnt[i];
        r = 0;
        for (int a2 = 0; a2 < E; ++a2) {
          const int c2 = cnt[sm.elist[a2]];
          r += (c2 > ci || (c2 == ci && a2 < a)) ? 1 : 0;
        }
      }
      sm.nrank[i] = r;
      sm.vis[r] = i;
    }
    __syncthreads();
    // ---- C: push sequence numbers, early break, new positions ----
    for (int r = tid; r < E; r += QT_T) {
      const int* c4 = sm.cc + sm.vis[r] * 4;
      const int v = (c4[0] > 0) + (c4[1] > 0) + (c4[2] > 0) + (c4[3] > 0);
      sm.nne[r] = v;
      sm.push[r] = v;
    }
    __syncthreads();
    const int pushed_all = block_excl_scan(sm.push, E, wtmp);
    int P = E;
    if (phase == 2) {
      for (int r = tid; r < E; r += QT_T) {
        const int grown = sm.push[r] + sm.nne[r] - (r + 1);  // size change after visiting rank r
        if (S + grown >= N) atomicMin(&s_misc[0], r);
      }
      __syncthreads();
      if (s_misc[0] != 0x7fffffff) P = s_misc[0] + 1;  // break after this node (:732)
    }
    const int TP = (P == E) ? pushed_all : sm.push[P];
    for (int i = tid; i < S; i += QT_T) {
      const int r = sm.nrank[i];
      sm.surv[i] = (r >= 0 && r < P) ? 0 : 1;
    }
    __syncthreads();
    const int NS = block_excl_scan(sm.surv, S, wtmp);
    const int S2 = TP + NS;
    if (S2 > C) { overflow = true; break; }
    int4* nbox = sm.box[cur ^ 1];
    int* ncnt = sm.cnt[cur ^ 1];
    for (int i = tid; i < S; i += QT_T) {
      const int r = sm.nrank[i];
      if (r >= 0 && r < P) {
        const int4 bx = box[i];
        const int mx = bx.x + ((bx.y - bx.x + 1) >> 1);
        const int my = bx.z + ((bx.w - bx.z + 1) >> 1);
        int pi = sm.push[r];
        int many = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int c = sm.cc[i * 4 + q];
          if (c > 0) {
            const int pos = TP - 1 - pi;  // push_front: later pushes sit nearer the front (:623-662)
            ++pi;
            ORBFE_DCHECK(pos >= 0 && pos < S2 && S2 <= C);
            nbox[pos] = make_int4((q & 1) ? mx : bx.x, (q & 1) ? bx.y : mx, (q & 2) ? my : bx.z, (q & 2) ? bx.w : my);
            ncnt[pos] = c;
            sm.cc[i * 4 + q] = pos;
            many += c > 1;
          }
        }
        if (many) atomicAdd(&s_misc[1], many);
      } else {
        const int pos = TP + sm.surv[i];
        ORBFE_DCHECK(pos >= 0 && pos < S2 && S2 <= C);
        nbox[pos] = box[i];
        ncnt[pos] = cnt[i];
        sm.surv[i] = pos;
      }
    }
    __syncthreads();
    // ---- D: re-point the keys ----
    for (int k = tid; k < n; k += QT_T) {
      const int nd = nk[k];
      const int r = sm.nrank[nd];
      if (r >= 0 && r < P) {
        const uint32_t c = ck[k];
        nk[k] = (uint16_t)sm.cc[nd * 4 + quadrant(box[nd], cand_x(c), cand_y(c))];
      } else {
        nk[k] = (uint16_t)sm.surv[nd];
      }
    }
    const int n_expand = s_misc[1];
    __syncthreads();
    const int prev = S;
    S = S2;
    cur ^= 1;
    if (S >= N || S == prev) break;                       // :669-672, :736-737
    if (phase == 1 && S + 3 * n_expand > N) phase = 2;  // :675
  }
  __syncthreads();
  if (overflow) {
    if (tid == 0) { atomicOr(&flags[b], FLAG_NODE_OVERFLOW); kept_count[b * kMaxLevels + lvl] = 0; }
    return;
  }
  // ---- best response per node, first in the reference's candidate order wins (:746-762) ----
  for (int i = tid; i < S; i += QT_T) sm.best[i] = 0ull;
  __syncthreads();
  const int w_cell = L.w_cell, h_cell = L.h_cell;
  auto order_key = [&](uint32_t c) -> unsigned long long {
    const int x = cand_x(c), y = cand_y(c);
    const unsigned long long cy = (unsigned)((y - 3) / h_cell), cx = (unsigned)((x - 3) / w_cell);
    const unsigned long long ord = (cy << 32) | (cx << 24) | ((unsigned long long)y << 12) | (unsigned long long)x;
    return ((unsigned long long)cand_score(c) << 40) | (~ord & 0xFFFFFFFFFFull);
  };
  for (int k = tid; k < n; k += QT_T) {
    ORBFE_DCHECK(nk[k] < S && S <= C);
    atomicMax(&sm.best[nk[k]], order_key(ck[k]));
  }
  __syncthreads();
  uint32_t* out = kept + (size_t)b * g.kept_total + L.kept_off;
  for (int k = tid; k < n; k += QT_T) {
    const uint32_t c = ck[k];
    const int nd = nk[k];
    if (sm.best[nd] == order_key(c) && nd < L.kept_cap) out[nd] = c;
  }
  if (tid == 0) kept_count[b * kMaxLevels + lvl] = min(S, L.kept_cap);
}

int configure_kernels(const PyrGeom& g) {
  const size_t need = quadtree_smem_bytes(g.max_node_cap);
  if (need > 220 * 1024) {
    set_error("nfeatures too large: quadtree needs %zu bytes of shared memory", need);
    return ORBFE_ERR_ARG;
  }
  if (need > 48 * 1024) {
    ORBFE_CUDA(cudaFuncSetAttribute(quadtree_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need));
    ORBFE_CUDA(cudaFuncSetAttribute(quadtree_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need));
  }
  if (upload_pattern() != 0) { set_error("rBRIEF pattern upload failed"); return ORBFE_ERR_CUDA; }
  if (configure_stereo_kernels(g) != 0) { set_error("cudaFuncSetAttribute(stereo_bucket_kernel) failed"); return ORBFE_ERR_CUDA; }
  if (configure_fast_kernel() != 0) { set_error("cudaFuncSetAttribute(fast_detect_kernel) failed"); return ORBFE_ERR_CUDA; }
  if (configure_describe_kernel() != 0) { set_error("cudaFuncSetAttribute(describe_kernel) failed"); return ORBFE_ERR_CUDA; }
  return ORBFE_OK;
}

void launch_quadtree(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s) {
  dim3 grid(batch, g.nlevels);
  const size_t smem = quadtree_smem_bytes(g.max_node_cap);
  if ((long long)batch * g.nlevels > 2 * 148)
    quadtree_kernel<3><<<grid, QT_T, smem, s>>>(g, a.cand, a.cell_hi, a.node_of, a.cand_count, a.kept, a.kept_count, a.flags);
  else
    quadtree_kernel<1><<<grid, QT_T, smem, s>>>(g, a.cand, a.cell_hi, a.node_of, a.cand_count, a.kept, a.kept_count, a.flags);
}

}  // namespace orbfe
