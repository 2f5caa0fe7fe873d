// k_fast.cu — the cell loop of ComputeKeyPointsOctTree (reference:
// src/ORBextractor.cc:776-850): cv::FAST(cell ROI, iniThFAST, NMS) and, when a
// cell returns nothing, cv::FAST(cell ROI, minThFAST, NMS) (:830-837).
//
// One CTA per 30-px cell (the reference's own unit: NMS never crosses a cell,
// SURVEY.md A.3 / B.2).  The ROI (cell + 3 px FAST margin on each side) is staged
// in shared memory.  The work is done in compacted stages so that warps stay
// convergent:
//   1. quick reject on the two opposite ring pairs (0,8) and (4,12): any 9-arc
//      contains one pixel of each pair -> warp-ballot compaction into a list;
//   2. corner strength m = max over the 16 arcs of min|diff| for listed pixels
//      (sliding-window min/max by doubling, 4 steps), score = m-1 into a score
//      tile; pixels with m > min(iniTh,minTh) go to a second list;
//   3. strict 8-neighbour maximum test on that list (score tile is 0 outside
//      the detection area, so suppression is cell-local as in the reference);
//   4. threshold choice per cell: the set of strict local maxima does not
//      depend on the threshold (a neighbour below the threshold is below the
//      centre anyway), so FAST(t) = {maxima with score >= t}; emit the iniTh
//      set if it is non-empty, else the minTh set.
// Candidates are appended (one atomicAdd per CTA) to the level's list as
// x | y<<12 | score<<24 with x,y relative to minBorder, as the reference stores
// them (:841-846).  List order is not the reference's; the consumer
// (k_quadtree.cu) breaks ties by the reference's order key computed from (x,y).
#include <cstdio>

#include "kernels.h"

namespace orbfe {

constexpr int FAST_T = 128;
constexpr int ROI_MAX = 66;  // w_cell <= 60 (+6)
constexpr int RS = 68;       // smem row stride
constexpr int LIST_MAX = 60 * 60;

__device__ __forceinline__ int warp_push(bool pred, int* counter) {
  const unsigned m = __ballot_sync(0xffffffffu, pred);
  if (m == 0) return -1;
  const int lane = threadIdx.x & 31;
  const int leader = __ffs(m) - 1;
  int base = 0;
  if (lane == leader) base = atomicAdd(counter, __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  return pred ? base + __popc(m & ((1u << lane) - 1u)) : -1;
}

// max over the 16 circular 9-arcs of max(min(d), min(-d)); 0 if negative
__device__ __forceinline__ int corner_strength(const uint8_t* p) {
  const int v = p[0];
  int d[16];
  d[0] = v - p[3 * RS];       d[1] = v - p[3 * RS + 1];   d[2] = v - p[2 * RS + 2];   d[3] = v - p[RS + 3];
  d[4] = v - p[3];            d[5] = v - p[-RS + 3];      d[6] = v - p[-2 * RS + 2];  d[7] = v - p[-3 * RS + 1];
  d[8] = v - p[-3 * RS];      d[9] = v - p[-3 * RS - 1];  d[10] = v - p[-2 * RS - 2]; d[11] = v - p[-RS - 3];
  d[12] = v - p[-3];          d[13] = v - p[RS - 3];      d[14] = v - p[2 * RS - 2];  d[15] = v - p[3 * RS - 1];
  // min over every 9-arc of d (darker ring) and of e = -d (brighter ring).  NB: written with two min chains on
  // purpose: ptxas 12.9 drops the negation when max(a, -b) is fused into VIMNMX3 on sm_100a (measured).
  int e[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) e[i] = -d[i];
  int a1[16], b1[16], a2[16], b2[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { a1[i] = min(d[i], d[(i + 1) & 15]); b1[i] = min(e[i], e[(i + 1) & 15]); }
#pragma unroll
  for (int i = 0; i < 16; ++i) { a2[i] = min(a1[i], a1[(i + 2) & 15]); b2[i] = min(b1[i], b1[(i + 2) & 15]); }
#pragma unroll
  for (int i = 0; i < 16; ++i) { a1[i] = min(a2[i], a2[(i + 4) & 15]); b1[i] = min(b2[i], b2[(i + 4) & 15]); }
  int best = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    best = max(best, min(a1[i], d[(i + 8) & 15]));
    best = max(best, min(b1[i], e[(i + 8) & 15]));
  }
  return best;
}

__global__ void __launch_bounds__(FAST_T) fast_cells_kernel(const __grid_constant__ PyrGeom g,
                                                            const uint8_t* __restrict__ pyr,
                                                            uint32_t* __restrict__ cand,
                                                            int32_t* __restrict__ cand_count,
                                                            int32_t* __restrict__ flags) {
  __shared__ __align__(16) uint8_t roi[ROI_MAX * RS];
  __shared__ __align__(16) uint8_t score[ROI_MAX * RS];
  __shared__ uint16_t list1[LIST_MAX];
  __shared__ uint16_t list2[LIST_MAX];
  __shared__ int n1, n2, n3, n_hi, out_base;

  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  int cell = blockIdx.x;
  int lvl = 0;
  while (lvl + 1 < g.nlevels && cell >= g.lv[lvl + 1].cell_base) ++lvl;
  const LevelGeom& L = g.lv[lvl];
  cell -= L.cell_base;
  if (L.n_cols <= 0 || cell >= L.n_cols * L.n_rows) return;
  const int ci = cell / L.n_cols, cj = cell - ci * L.n_cols;
  const int x0 = kMinBorder + cj * L.w_cell, y0 = kMinBorder + ci * L.h_cell;
  if (y0 >= L.max_by - 3 || x0 >= L.max_bx - 6) return;  // :806, :815
  const int cw = min(x0 + L.w_cell + 6, L.max_bx) - x0;
  const int ch = min(y0 + L.h_cell + 6, L.max_by) - y0;
  if (cw < 7 || ch < 7) return;  // cv::FAST finds nothing in such a sliver

  if (tid == 0) { n1 = 0; n2 = 0; n3 = 0; n_hi = 0; }
  // i / cw and i / dw below: multiply by ceil(2^32 / d) (exact for i < 2^32 / d; here i < 4489)
  const unsigned inv_cw = 0xFFFFFFFFu / (unsigned)cw + 1u;
  const uint8_t* img = pyr + L.lvl_base + (size_t)b * L.lvl_bytes + (size_t)y0 * L.pitch + x0;
  for (int i = tid; i < ch * cw; i += FAST_T) {
    const int r = (int)__umulhi((unsigned)i, inv_cw), c = i - r * cw;
    roi[r * RS + c] = __ldg(img + (size_t)r * L.pitch + c);
  }
  for (int i = tid; i < ch * (RS / 4); i += FAST_T) reinterpret_cast<uint32_t*>(score)[i] = 0u;
  __syncthreads();

  const int t_lo = min(g.ini_th, g.min_th);
  const int dw = cw - 6, dh = ch - 6, npx = dw * dh;
  const unsigned inv_dw = 0xFFFFFFFFu / (unsigned)dw + 1u;
  // stage 1: quick reject
  for (int base = 0; base < npx; base += FAST_T) {
    const int i = base + tid;
    bool keep = false;
    int pos = 0;
    if (i < npx) {
      const int r = (int)__umulhi((unsigned)i, inv_dw), c = i - r * dw;
      pos = (r + 3) * RS + (c + 3);
      const uint8_t* p = roi + pos;
      const int v = p[0], hi = v + t_lo, lo = v - t_lo;
      const int a = p[3 * RS], e = p[-3 * RS], f = p[3], h = p[-3];
      const bool br = (a > hi || e > hi) && (f > hi || h > hi);
      const bool dk = (a < lo || e < lo) && (f < lo || h < lo);
      keep = br || dk;
    }
    const int slot = warp_push(keep, &n1);
    if (slot >= 0) list1[slot] = (uint16_t)pos;
  }
  __syncthreads();
  // stage 2: corner strength
  const int c1 = n1;
  for (int base = 0; base < c1; base += FAST_T) {
    const int i = base + tid;
    bool keep = false;
    int pos = 0;
    if (i < c1) {
      pos = list1[i];
      const int m = corner_strength(roi + pos);
      if (m > t_lo) { keep = true; score[pos] = (uint8_t)(m - 1); }
    }
    const int slot = warp_push(keep, &n2);
    if (slot >= 0) list2[slot] = (uint16_t)pos;
  }
  __syncthreads();
  // stage 3: strict maximum over the 8 neighbours; survivors overwrite list1
  const int c2 = n2;
  for (int base = 0; base < c2; base += FAST_T) {
    const int i = base + tid;
    bool keep = false;
    int pos = 0;
    if (i < c2) {
      pos = list2[i];
      const uint8_t* s = score + pos;
      const int c = s[0];
      keep = c > s[-1] && c > s[1] && c > s[-RS - 1] && c > s[-RS] && c > s[-RS + 1] && c > s[RS - 1] && c > s[RS] &&
             c > s[RS + 1];
      if (keep && c >= g.ini_th) n_hi = 1;  // benign race: all writers store 1
    }
    const int slot = warp_push(keep, &n3);
    if (slot >= 0) list1[slot] = (uint16_t)pos;
  }
  __syncthreads();
  // stage 4: threshold choice + emit
  const int c3 = n3;
  if (c3 == 0) return;
  const int t_emit = n_hi ? g.ini_th : g.min_th;
  __syncthreads();
  if (tid == 0) n2 = 0;
  __syncthreads();
  for (int base = 0; base < c3; base += FAST_T) {
    const int i = base + tid;
    bool keep = false;
    int pos = 0;
    if (i < c3) {
      pos = list1[i];
      keep = (int)score[pos] >= t_emit;
    }
    const int slot = warp_push(keep, &n2);
    if (slot >= 0) list2[slot] = (uint16_t)pos;
  }
  __syncthreads();
  const int ne = n2;
#ifdef ORBFE_DEBUG_FAST
  if (blockIdx.x == 0 && b == 0 && tid == 0) {
    printf("DBG n1 %d nc %d n3 %d n_hi %d ne %d\n", n1, c2, c3, n_hi, ne);
    for (int i = 0; i < ne; ++i) printf("DBG emit pos %d -> x %d y %d score %d\n", list2[i], list2[i] % RS, list2[i] / RS, score[list2[i]]);
  }
#endif
  if (ne == 0) return;
  if (tid == 0) out_base = atomicAdd(&cand_count[b * kMaxLevels + lvl], ne);
  __syncthreads();
  const int ob = out_base;
  uint32_t* out = cand + (size_t)b * g.cand_total + L.cand_off;
  for (int i = tid; i < ne; i += FAST_T) {
    const int pos = list2[i];
    const int r = (int)__umulhi((unsigned)pos, 0xFFFFFFFFu / RS + 1u), c = pos - r * RS;
    if (ob + i < L.cand_cap) out[ob + i] = pack_cand(c + cj * L.w_cell, r + ci * L.h_cell, score[pos]);
    else atomicOr(&flags[b], FLAG_CAND_OVERFLOW);
  }
}

void launch_fast_cells(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s) {
  dim3 grid(g.total_cells, batch);
  fast_cells_kernel<<<grid, FAST_T, 0, s>>>(g, a.pyr, a.cand, a.cand_count, a.flags);
}

}  // namespace orbfe
