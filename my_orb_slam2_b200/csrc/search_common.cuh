// search_common.cuh — device helpers shared by k_search.cu and k_search2.cu: Hamming distance of two descriptors,
// the CTA scan, the Frame grid lookups (src/Frame.cc:243-258, :351-417), the rotation histogram
// (src/ORBmatcher.cc:1669-1710) and the packed candidate word.
#pragma once
#include <climits>

#include "match_kernels.h"

namespace orbfe {

constexpr int GRID_COLS = ORBFE_GRID_COLS, GRID_ROWS = ORBFE_GRID_ROWS, GRID_CELLS = GRID_COLS * GRID_ROWS;
constexpr int TH_HIGH = ORBFE_TH_HIGH, TH_LOW = ORBFE_TH_LOW, HISTO = ORBFE_HISTO_LENGTH;

__device__ __forceinline__ int hamming32(const uint8_t* a, const uint8_t* b) {
  const uint4* pa = reinterpret_cast<const uint4*>(a);
  const uint4* pb = reinterpret_cast<const uint4*>(b);
  const uint4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pb), b1 = __ldg(pb + 1);
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// exclusive scan of a[0..n) in place by one CTA of 1024 threads; returns the total
__device__ inline int cta_excl_scan(int32_t* a, int n, int* wtmp) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, T = blockDim.x;
  const int per = (n + T - 1) / T;
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
    int w = lane < (T >> 5) ? wtmp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    wtmp[lane] = w;
  }
  __syncthreads();
  int run = (wid > 0 ? wtmp[wid - 1] : 0) + v - s;
  const int total = wtmp[(T >> 5) - 1];
  for (int i = beg; i < end; ++i) { const int t = a[i]; a[i] = run; run += t; }
  __syncthreads();
  return total;
}

// ---- Frame::AssignFeaturesToGrid + PosInGrid (src/Frame.cc:243-258, :407-417) ----
__device__ __forceinline__ int pos_in_grid(const DevFrame& f, const orbfe_keypoint& kp) {
  const int px = (int)roundf(__fmul_rn(__fsub_rn(kp.x, f.min_x), f.inv_w));
  const int py = (int)roundf(__fmul_rn(__fsub_rn(kp.y, f.min_y), f.inv_h));
  if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) return -1;
  return px * GRID_ROWS + py;  // mGrid[x][y]
}

// Frame::GetFeaturesInArea (src/Frame.cc:351-405) as a visitor; fn(idx) is called in the reference's order
template <typename Fn>
__device__ __forceinline__ void features_in_area(const DevFrame& f, float x, float y, float r, int min_level,
                                                 int max_level, Fn fn) {
  const int min_cx = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, f.min_x), r), f.inv_w)));
  if (min_cx >= GRID_COLS) return;
  const int max_cx = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, f.min_x), r), f.inv_w)));
  if (max_cx < 0) return;
  const int min_cy = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, f.min_y), r), f.inv_h)));
  if (min_cy >= GRID_ROWS) return;
  const int max_cy = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, f.min_y), r), f.inv_h)));
  if (max_cy < 0) return;
  const bool check_levels = (min_level > 0) || (max_level >= 0);
  for (int ix = min_cx; ix <= max_cx; ++ix) {
    const int b = f.cell_start[ix * GRID_ROWS + min_cy], e = f.cell_start[ix * GRID_ROWS + max_cy + 1];
    for (int p = b; p < e; ++p) {  // cells iy = min_cy..max_cy of column ix are contiguous in the CSR
      const int idx = f.cell_items[p];
      const orbfe_keypoint kp = f.keys[idx];
      if (check_levels) {
        if (kp.octave < min_level) continue;
        if (max_level >= 0 && kp.octave > max_level) continue;
      }
      const float dx = __fsub_rn(kp.x, x), dy = __fsub_rn(kp.y, y);
      if (fabsf(dx) < r && fabsf(dy) < r) fn(idx, kp.octave);
    }
  }
}

__device__ __forceinline__ uint32_t pack_sc(int idx, int dist, int oct) {
  return (uint32_t)idx | ((uint32_t)dist << 16) | ((uint32_t)oct << 25);
}

// ---- rotation histogram helpers (:1499-1509, :1669-1710) ----
__device__ __forceinline__ int rot_bin(float a1, float a2) {
  const float factor = 1.0f / HISTO;
  float rot = __fsub_rn(a1, a2);
  if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
  int bin = (int)roundf(__fmul_rn(rot, factor));
  if (bin == HISTO) bin = 0;
  return bin;
}

// ORBmatcher::ComputeThreeMaxima on the bin sizes
__device__ inline void three_maxima(const int* histo, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  ind1 = ind2 = ind3 = -1;
  for (int i = 0; i < HISTO; ++i) {
    const int s = histo[i];
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
  else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}

}  // namespace orbfe
