// k_stereo.cu — Frame::ComputeStereoMatches (reference: src/Frame.cc:496-686) on the
// device-resident outputs of the left and right extractor handles.
//
// Kernel 0, one CTA per image: the row table of :511-531 as buckets of 8 image rows (CSR in global memory): right
// keypoint iR goes into every bucket its band [floor(yR-r), ceil(yR+r)] touches; its exact band, x and octave are
// stored next to it.  Kernel 1, one warp per left keypoint, scans only the bucket of row int(vL):
//   * row-band candidates (:511-531,:549): right keypoint iR is a candidate of row
//     int(vL) iff floor(yR - r) <= int(vL) <= ceil(yR + r), r = 2*scale[octave(iR)];
//     the per-row index vectors of the reference are in ascending iR, and the best
//     Hamming distance is taken with a strict '<' (:581), so the winner is the
//     lexicographic minimum of (distance, iR) — a warp min-reduction of dist<<16|iR;
//   * 11x11 SAD over shifts -5..+5 on the left keypoint's pyramid level (:591-631),
//     computed in int32 (all operands are integers), one warp reduction
//     per shift; parabola fit, disparity gate and depth (:633-667) by lane 0 with
//     explicitly rounded fp32 operations.
// Kernel 2, one CTA per image: median of the kept SADs and the 1.5*1.4*median cut
// (:672-685).  Order of vDistIdx does not matter: the cut removes every match whose
// SAD is >= the threshold.
#include <climits>

#include "kernels.h"

namespace orbfe {

constexpr int ST_WARPS = 1;  // one warp per CTA: a slot frees as soon as its keypoint is done (4 warps 0.116 ms, 2 warps 0.112, 1 warp 0.109)
constexpr int ST_BROWS = 8;       // image rows per bucket of the row table
constexpr int TH_HIGH = ORBFE_TH_HIGH, TH_LOW = ORBFE_TH_LOW;

__device__ __forceinline__ int hamming256(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// row table (:511-531) as 8-row buckets; one CTA per image
__global__ void __launch_bounds__(512) stereo_bucket_kernel(const __grid_constant__ PyrGeom g,
                                                            const orbfe_keypoint* __restrict__ kpR,
                                                            const int32_t* __restrict__ cntR, int4* __restrict__ band,
                                                            int32_t* __restrict__ bucket_start,
                                                            int4* __restrict__ bucket_items, int items_per_kp) {
  // counting sort by (bucket, octave): a left keypoint of level L only scans the items of octaves L-1 .. L+1 of its
  // bucket (:565-566), a contiguous slice of the item array
  extern __shared__ int s_cnt[];  // nb * nlevels + 1
  const int b = blockIdx.x, tid = threadIdx.x;
  const int nR = cntR[b];
  const int n_rows = g.lv[0].h;
  const int nb = (n_rows + ST_BROWS - 1) / ST_BROWS;
  const int nl = g.nlevels, nk = nb * nl;
  const orbfe_keypoint* kr = kpR + (size_t)b * g.out_cap;
  int4* bands = band + (size_t)b * g.out_cap;
  int32_t* bstart = bucket_start + (size_t)b * (nk + 1);
  int4* bitems = bucket_items + (size_t)b * g.out_cap * items_per_kp;
  for (int i = tid; i <= nk; i += 512) s_cnt[i] = 0;
  __syncthreads();
  for (int i = tid; i < nR; i += 512) {
    const orbfe_keypoint k = kr[i];
    const float r = __fmul_rn(2.0f, g.lv[k.octave].scale);  // :525
    const int maxr = (int)ceilf(__fadd_rn(k.y, r)), minr = (int)floorf(__fsub_rn(k.y, r));
    bands[i] = make_int4(minr, maxr, __float_as_int(k.x), k.octave);
    const int lo = max(minr, 0) / ST_BROWS, hi = min(maxr, n_rows - 1) / ST_BROWS;
    for (int q = lo; q <= hi; ++q) atomicAdd(&s_cnt[q * nl + k.octave], 1);
  }
  __syncthreads();
  // exclusive scan of s_cnt[0..nk) by warp 0
  if (tid < 32) {
    const int per = (nk + 31) / 32;
    const int beg = min(tid * per, nk), end = min(beg + per, nk);
    int s = 0;
    for (int i = beg; i < end; ++i) s += s_cnt[i];
    int v = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (tid >= o) v += t;
    }
    int run = v - s;
    for (int i = beg; i < end; ++i) { const int t = s_cnt[i]; s_cnt[i] = run; run += t; }
    if (tid == 31) s_cnt[nk] = v;
  }
  __syncthreads();
  for (int i = tid; i <= nk; i += 512) bstart[i] = s_cnt[i];
  __syncthreads();
  const int cap_items = g.out_cap * items_per_kp;
  for (int i = tid; i < nR; i += 512) {
    const int4 bd = bands[i];
    const int lo = max(bd.x, 0) / ST_BROWS, hi = min(bd.y, n_rows - 1) / ST_BROWS;
    for (int q = lo; q <= hi; ++q) {
      const int slot = atomicAdd(&s_cnt[q * nl + bd.w], 1);
      if (slot < cap_items) bitems[slot] = make_int4(bd.x, bd.y, bd.z, bd.w | (i << 8));  // band inline: no second lookup
    }
  }
}

// 32 warps per SM (64 registers): measured best (40 / 48 / 51 registers spill the SAD accumulators, 80 lose latency hiding)
__global__ void __launch_bounds__(ST_WARPS * 32, 32) stereo_match_kernel(
    const __grid_constant__ PyrGeom g, const uint8_t* __restrict__ pyrL, const uint8_t* __restrict__ pyrR,
    const orbfe_keypoint* __restrict__ kpL, const uint8_t* __restrict__ descL, const int32_t* __restrict__ cntL,
    const orbfe_keypoint* __restrict__ kpR, const uint8_t* __restrict__ descR, const int4* __restrict__ band,
    const int32_t* __restrict__ bucket_start, const int4* __restrict__ bucket_items, int items_per_kp,
    float mbf, float mb, float* __restrict__ u_right, float* __restrict__ depth, int32_t* __restrict__ sad_out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int nL = cntL[b];
  const int iL = blockIdx.x * ST_WARPS + warp;
  if (iL >= nL) return;
  const bool active = true;
  const orbfe_keypoint* kl = kpL + (size_t)b * g.out_cap;
  const orbfe_keypoint* kr = kpR + (size_t)b * g.out_cap;
  const int n_rows = g.lv[0].h;                  // mvImagePyramid[0].rows (:503)
  const float maxD = __fdiv_rn(mbf, mb);         // :536
  const float uL = kl[iL].x, vL = kl[iL].y;
  const int levelL = kl[iL].octave;
  const int rowi = (int)vL;  // :549
  const uint4* dpl = reinterpret_cast<const uint4*>(descL + ((size_t)b * g.out_cap + iL) * 32);
  const uint4 d0 = __ldg(dpl), d1 = __ldg(dpl + 1);
  const float minU = __fsub_rn(uL, maxD), maxU = uL;  // :554-555 (minD = 0)
  const bool searchable = rowi >= 0 && rowi < n_rows && !(maxU < 0.f);
  unsigned best = ((unsigned)TH_HIGH << 16);  // bestDist = TH_HIGH, strict '<' below (:560,:581)
  int best_x = 0;  // x bits of the lane's best right keypoint (rides along: no dependent kr[bestIdxR] load afterwards)
  if (searchable) {
    const int nb = (n_rows + ST_BROWS - 1) / ST_BROWS, nl = g.nlevels;
    const int32_t* bstart = bucket_start + (size_t)b * (nb * nl + 1);
    const int4* bitems = bucket_items + (size_t)b * g.out_cap * items_per_kp;
    // the bucket's items are sorted by octave: only octaves levelL-1 .. levelL+1 are candidates (:565-566)
    const int k0 = (rowi / ST_BROWS) * nl;
    const int beg = bstart[k0 + max(levelL - 1, 0)], end = bstart[k0 + min(levelL + 1, nl - 1) + 1];
    ORBFE_DCHECK(levelL >= 0 && levelL < nl && beg >= 0 && beg <= end && end <= g.out_cap * items_per_kp);
    const uint8_t* dR = descR + (size_t)b * g.out_cap * 32;
    // 4 items per lane in flight: the item loads of a 128-item chunk are issued together, then the descriptor loads
    for (int base = beg; base < end; base += 128) {
      int4 bd[4];
      bool cand[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int p = base + lane + 32 * u;
        bd[u] = p < end ? __ldg(bitems + p) : make_int4(1, 0, 0, 0);  // minr > maxr: never a candidate
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float uR = __int_as_float(bd[u].z);
        cand[u] = rowi >= bd[u].x && rowi <= bd[u].y && uR >= minU && uR <= maxU;
      }
#pragma unroll
      for (int h = 0; h < 4; h += 2) {  // descriptor loads two at a time (register budget)
        uint4 r0[2], r1[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (cand[h + u]) {
            const uint4* dp = reinterpret_cast<const uint4*>(dR + (size_t)(bd[h + u].w >> 8) * 32);
            r0[u] = __ldg(dp);
            r1[u] = __ldg(dp + 1);
          }
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          if (cand[h + u]) {
            const unsigned dist = (unsigned)hamming256(d0, d1, r0[u], r1[u]);
            const unsigned key = (dist << 16) | (unsigned)(bd[h + u].w >> 8);
            if (key < best) { best = key; best_x = bd[h + u].z; }
          }
        }
      }
    }
  }
  if (!active) return;
  const unsigned my_best = best;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  // the winning lane (keys are unique: they carry the right keypoint's index) hands over its keypoint's x
  const unsigned owners = __ballot_sync(0xffffffffu, my_best == best);
  best_x = __shfl_sync(0xffffffffu, best_x, __ffs(owners) - 1);

  const size_t oidx = (size_t)b * g.out_cap + iL;
  float res_u = -1.f, res_d = -1.f;
  int res_sad = -1;
  const int bestDist = (int)(best >> 16);
  if (bestDist < (TH_HIGH + TH_LOW) / 2) {  // :591 (bestDist stays TH_HIGH when not searchable)
    const int bestIdxR = (int)(best & 0xFFFFu);
    ORBFE_DCHECK(bestIdxR < g.out_cap);
    const float uR0 = __int_as_float(best_x);  // == kr[bestIdxR].x (the bucket item carries the bits)
    const LevelGeom& LV = g.lv[levelL];
    const float sf = LV.inv_scale;
    const float scaleduL = roundf(__fmul_rn(uL, sf));
    const float scaledvL = roundf(__fmul_rn(vL, sf));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const int w = 5, Lh = 5;
    const int cuL = (int)scaleduL, cvL = (int)scaledvL, cuR = (int)scaleduR0;
    const float iniu = scaleduR0 + Lh - w, endu = scaleduR0 + Lh + w + 1;
    const bool ok = !(cvL - w < 0 || cvL + w >= LV.h || cuL - w < 0 || cuL + w >= LV.w) &&
                    !(iniu < 0.f || endu >= (float)LV.w) && !(cuR - Lh - w < 0);  // :612-615 + memory guards
    if (ok) {
      const uint8_t* imL = pyrL + LV.lvl_base + (size_t)b * LV.lvl_bytes;
      const uint8_t* imR = pyrR + LV.lvl_base + (size_t)b * LV.lvl_bytes;
      // the 11 left and 11 right window rows are 22 different cache lines, mostly L2 / DRAM misses: request them all
      // at once (the loop below walks them three rows per iteration and would otherwise take the misses in turn)
      if (lane < 22) {
        const uint8_t* row = lane < 11 ? imL + (size_t)(cvL - w + lane) * LV.pitch + cuL - w
                                       : imR + (size_t)(cvL - w + lane - 11) * LV.pitch + cuR - w - Lh;
        asm volatile("prefetch.global.L1 [%0];" ::"l"(row));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(row + 20));
      }
      ORBFE_DCHECK(cvL - w >= 0 && cvL + w < LV.h && cuL - w >= 0 && cuL + w < LV.w && cuR - w - Lh >= 0 &&
                   cuR + w + Lh < LV.pitch);
      const int centerL = __ldg(imL + (size_t)cvL * LV.pitch + cuL);
      int sad[11];
#pragma unroll
      for (int s = 0; s < 11; ++s) sad[s] = 0;
#pragma unroll  // 4 rounds: all their loads are issued before the first round's arithmetic
      for (int p = lane; p < 121; p += 32) {
        const int dy = p / 11 - w, dx = p - (p / 11) * 11 - w;
        const int lv = (int)__ldg(imL + (size_t)(cvL + dy) * LV.pitch + cuL + dx) - centerL;
        const uint8_t* rr = imR + (size_t)(cvL + dy) * LV.pitch + cuR + dx;
#pragma unroll
        for (int s = 0; s < 11; ++s) {
          const int centerR = __ldg(imR + (size_t)cvL * LV.pitch + cuR + s - Lh);
          sad[s] = (int)__sad(lv, (int)__ldg(rr + s - Lh) - centerR, (unsigned)sad[s]);  // |a - b| + c: one VABSDIFF
        }
      }
      // warp sums: a total is at most 121 * 510 = 61 710 < 2^16, so two of them share a register through the butterfly
      // (6 registers x 5 steps instead of 11 x 5 shuffles)
      {
        unsigned pk[6];
#pragma unroll
        for (int i = 0; i < 5; ++i) pk[i] = (unsigned)sad[2 * i] | ((unsigned)sad[2 * i + 1] << 16);
        pk[5] = (unsigned)sad[10];
#pragma unroll
        for (int i = 0; i < 6; ++i) {
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) pk[i] += __shfl_xor_sync(0xffffffffu, pk[i], o);
        }
#pragma unroll
        for (int i = 0; i < 5; ++i) { sad[2 * i] = (int)(pk[i] & 0xFFFFu); sad[2 * i + 1] = (int)(pk[i] >> 16); }
        sad[10] = (int)pk[5];
      }
      // strict '<': first minimum (:624); its two neighbours ride along (no dynamically indexed local array)
      int bestSad = sad[0], bestInc = -Lh, sPrev = 0, sNext = 0;
#pragma unroll
      for (int s = 1; s < 11; ++s) {
        const bool lt = sad[s] < bestSad;
        bestSad = lt ? sad[s] : bestSad;
        bestInc = lt ? s - Lh : bestInc;
        sPrev = lt ? sad[s - 1] : sPrev;
        sNext = lt ? (s < 10 ? sad[s + 1] : 0) : sNext;
      }
      if (!(bestInc == -Lh || bestInc == Lh)) {  // :633
        const float d1f = (float)sPrev, d2f = (float)bestSad, d3f = (float)sNext;
        // deltaR = (d1-d3)/(2*(d1+d3-2*d2))  (:644), every step rounded to fp32
        const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1f, d3f), __fmul_rn(2.0f, d2f)));
        const float deltaR = __fdiv_rn(__fsub_rn(d1f, d3f), den);
        if (!(deltaR < -1.f || deltaR > 1.f)) {  // NaN passes, as in the reference (:646)
          float bestuR = __fmul_rn(LV.scale, __fadd_rn(__fadd_rn(scaleduR0, (float)bestInc), deltaR));  // :653
          float disparity = __fsub_rn(uL, bestuR);
          if (disparity >= 0.f && disparity < maxD) {
            if (disparity <= 0.f) {
              disparity = (float)0.01;                       // :661
              bestuR = (float)((double)uL - 0.01);           // :662
            }
            res_d = __fdiv_rn(mbf, disparity);
            res_u = bestuR;
            res_sad = bestSad;
          }
        }
      }
    }
  }
  if (lane == 0) { u_right[oidx] = res_u; depth[oidx] = res_d; sad_out[oidx] = res_sad; }
}

// median cut (:672-685)
__global__ void __launch_bounds__(1024) stereo_median_kernel(const __grid_constant__ PyrGeom g,
                                                             const int32_t* __restrict__ cntL,
                                                             float* __restrict__ u_right, float* __restrict__ depth,
                                                             const int32_t* __restrict__ sad,
                                                             int32_t* __restrict__ n_matched) {
  extern __shared__ int s_sad[];  // out_cap
  __shared__ int s_n, s_median, s_removed;
  const int b = blockIdx.x, tid = threadIdx.x;
  const int nL = cntL[b];
  const size_t off = (size_t)b * g.out_cap;
  if (tid == 0) { s_n = 0; s_median = 0; s_removed = 0; }
  // the reference initialises mvuRight / mvDepth to -1 for all N (:498-499): entries past this image's keypoint
  // count are defined too, so a caller that downloads `capacity` entries never sees stale values
  for (int i = nL + tid; i < g.out_cap; i += blockDim.x) { u_right[off + i] = -1.f; depth[off + i] = -1.f; }
  __syncthreads();
  for (int i = tid; i < nL; i += blockDim.x) {
    const int v = sad[off + i];
    if (v >= 0) s_sad[atomicAdd(&s_n, 1)] = v;
  }
  __syncthreads();
  const int n = s_n;
  if (n == 0) {  // the reference indexes an empty vector here; we return (SURVEY.md 8(c) deviation 2)
    if (tid == 0) n_matched[b] = 0;
    return;
  }
  const int kth = n / 2;
  for (int i = tid; i < n; i += blockDim.x) {
    const int v = s_sad[i];
    int less = 0, leq = 0;
    for (int j = 0; j < n; ++j) { const int t = s_sad[j]; less += t < v; leq += t <= v; }
    if (less <= kth && kth < leq) s_median = v;  // all writers store the same value
  }
  __syncthreads();
  const float thDist = 1.5f * 1.4f * (float)s_median;  // :674
  for (int i = tid; i < nL; i += blockDim.x) {
    const int v = sad[off + i];
    if (v >= 0 && !((float)v < thDist)) {
      u_right[off + i] = -1.f;
      depth[off + i] = -1.f;
      atomicAdd(&s_removed, 1);
    }
  }
  __syncthreads();
  if (tid == 0) n_matched[b] = n - s_removed;
}

void launch_stereo(const PyrGeom& g, const ExtractorArena& l, const ExtractorArena& r, int batch, float mbf, float mb,
                   cudaStream_t s, cudaEvent_t* ev) {
  if (ev) cudaEventRecord(ev[0], s);
  const size_t bucket_smem = ((size_t)((g.lv[0].h + ST_BROWS - 1) / ST_BROWS) * g.nlevels + 1) * sizeof(int);
  stereo_bucket_kernel<<<batch, 512, bucket_smem, s>>>(g, r.out_kp, r.out_count, l.st_band, l.st_bucket_start, l.st_bucket_items,
                                            l.st_items_per_kp);
  dim3 grid((g.out_cap + ST_WARPS - 1) / ST_WARPS, batch);
  stereo_match_kernel<<<grid, ST_WARPS * 32, 0, s>>>(g, l.pyr, r.pyr, l.out_kp, l.out_desc, l.out_count, r.out_kp,
                                                     r.out_desc, l.st_band, l.st_bucket_start, l.st_bucket_items,
                                                     l.st_items_per_kp, mbf, mb, l.st_uright, l.st_depth, l.st_sad);
  if (ev) cudaEventRecord(ev[1], s);
  stereo_median_kernel<<<batch, 1024, (size_t)g.out_cap * sizeof(int), s>>>(g, l.out_count, l.st_uright, l.st_depth,
                                                                            l.st_sad, l.st_count);
  if (ev) cudaEventRecord(ev[2], s);
}

// the bucket kernel's counters (buckets x levels) can exceed the 48 KB default for very tall images
int configure_stereo_kernels(const PyrGeom& g) {
  const size_t need = ((size_t)((g.lv[0].h + ST_BROWS - 1) / ST_BROWS) * g.nlevels + 1) * sizeof(int);
  if (need <= 48 * 1024) return 0;
  return cudaFuncSetAttribute(stereo_bucket_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need) == cudaSuccess ? 0 : -1;
}

int stereo_items_per_kp(const PyrGeom& g) {
  float smax = 1.f;
  for (int i = 0; i < g.nlevels; ++i) smax = g.lv[i].scale > smax ? g.lv[i].scale : smax;
  return (int)((4.f * smax + 3.f) / ST_BROWS) + 2;  // buckets one band can touch
}

}  // namespace orbfe
