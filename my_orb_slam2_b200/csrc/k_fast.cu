// k_fast.cu — the cell loop of ComputeKeyPointsOctTree (reference:
// src/ORBextractor.cc:776-850): cv::FAST(cell ROI, iniThFAST, NMS) and, when a
// cell returns nothing, cv::FAST(cell ROI, minThFAST, NMS) (:830-837).
//
// The corner score of a pixel does not depend on the cell it is in; only the
// non-maximum suppression (cell-local, because cv::FAST runs on the cell ROI) and
// the threshold choice do.  So detection is ONE dense,
// streaming kernel over whole levels, and the per-cell threshold choice is a
// filter in front of the quadtree (k_quadtree.cu):
//
//  fast_detect_kernel — a warp owns a strip of one level: it computes 128 columns
//  x 32 rows and owns the inner 120 x 30 (1-px halo for the 3x3 suppression, padded
//  to whole words).  Warps are independent (no CTA barrier).
//   1. dense quick reject.  A lane owns one aligned 32-bit word (4 pixels) and
//      walks down the rows with the last 7 centre words in registers.  Per row it
//      loads three aligned words (coalesced 128 B per warp), forms the x-3 / x+3
//      neighbours with funnel shifts and tests four pixels at once: any 9-arc of
//      the 16-ring contains one pixel of each opposite pair (0,8) and (4,12), so a
//      corner needs |p-v| > t on one of each pair — 4 x VABSDIFF4 + SWAR "byte >
//      t".  Survivors (~14 % on textured images) stay a 128-bit register mask per
//      lane; the rows are also copied into a shared-memory tile for step 2.
//   2. exact corner strength, one lane per survivor (masks flattened into a
//      per-warp list so the lanes are balanced): m = max over the 16 arcs of
//      min|diff| = max(max_arc(min ring) - v, v - min_arc(max ring)); with
//      x = p | (255-p)<<16 both halves are arc MINIMA, so 40 VIMNMX3.U16x2 do all
//      32 arcs.  score = m-1 goes into a shared-memory score tile.
//   3. strict 8-neighbour maximum on the score tile for the owned pixels with a
//      score; a neighbour in another cell's detection window counts as 0.  The set
//      of strict local maxima does not depend on the threshold (a neighbour below
//      the threshold is below the centre anyway), so FAST(t) = {maxima with score
//      >= t}.
//  The kernel runs twice: at iniTh over every strip (each maximum is appended to the level's list and flags its
//  cell), then at minTh, where a strip returns at once unless one of the cells under its owned pixels has no flag,
//  and appends maxima of unflagged cells only.  On textured images the second launch does nothing, and the first
//  one computes exact corner strengths for the iniTh survivors only (a third of the minTh survivors); on smooth
//  regions the minTh pass has few survivors anyway.  The consumer keeps score >= iniTh in flagged cells and
//  score >= minTh in the others (:830-837).
// Candidates are x | y<<12 | score<<24 with x,y relative to minBorder, as the
// reference stores them (:841-846).  List order is not the reference's; the
// consumer (k_quadtree.cu) breaks ties by the reference's order key computed
// from (x,y).
#include <cuda.h>

#include <cstdio>

#include "kernels.h"

namespace orbfe {

constexpr int FD_WARPS = 1;             // independent warps per CTA: one (a slot frees as soon as its strip ends; 2 warps 0.168 ms, 4 warps 0.173 ms, 1 warp 0.159 ms)
constexpr int FD_T = FD_WARPS * 32;
constexpr int FD_ROWS = kFastOwnRows + 2;  // computed rows per strip (multiple of 8)
constexpr int FD_ACC = FD_ROWS / 8;     // 32-bit survivor masks per lane (8 rows x 4 pixels each)
static_assert(FD_ROWS % 8 == 0 && FD_ROWS <= 32, "strip rows");
constexpr int FD_TROWS = FD_ROWS + 6;   // image tile rows (FAST ring: +-3)
constexpr int FD_TS = kFastTileW;       // image tile stride in bytes (the TMA box: 136 bytes used, from column -4 of the strip)
constexpr int FD_TW = FD_TS / 4;
static_assert(FD_TROWS == kFastTileH, "TMA box rows");
constexpr int FD_CAP = FD_ROWS * 64;            // list entries per pass (a strip with more survivors takes several passes)

struct __align__(128) FdSmem {
  uint8_t img[FD_TROWS * FD_TS];   // TMA destination (128-byte aligned)
  uint8_t score[FD_ROWS * 128];
  uint16_t list[FD_CAP];
  unsigned long long bar;          // mbarrier the tile's TMA load completes on
};

// ---- TMA: one elected lane loads the strip's whole image tile (kFastTileW x kFastTileH bytes of image b of a level;
// rows / columns past the tensor are zero-filled and only feed masked pixels) and the warp waits on the mbarrier ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void tma_load_tile(void* dst, const void* tmap, int x, int y, int z, unsigned long long* bar) {
  const unsigned b = smem_u32(bar);
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(b) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(b), "r"(FD_TROWS * FD_TS) : "memory");
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n" ::
          "r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(b)
      : "memory");
}
__device__ __forceinline__ void mbar_wait_phase0(unsigned long long* bar) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "FD_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
      "@p bra FD_DONE;\n"
      "bra FD_WAIT;\n"
      "FD_DONE:\n"
      "}\n" ::"r"(smem_u32(bar))
      : "memory");
}

// exact corner strength of the pixel at p in the image tile: max over the 16 circular 9-arcs of the
// minimum of (ring - v) resp. (v - ring) over the arc, 0 if negative.  Unsigned 16x2 min/max only (no negations:
// ptxas 12.9 drops the negation when max(a, -b) is fused into VIMNMX3 on sm_100a).
__device__ __forceinline__ int corner_strength(const uint8_t* p) {
  constexpr int S = FD_TS;
  unsigned x[16];
  x[0] = p[3 * S];       x[1] = p[3 * S + 1];   x[2] = p[2 * S + 2];    x[3] = p[S + 3];
  x[4] = p[3];           x[5] = p[-S + 3];      x[6] = p[-2 * S + 2];   x[7] = p[-3 * S + 1];
  x[8] = p[-3 * S];      x[9] = p[-3 * S - 1];  x[10] = p[-2 * S - 2];  x[11] = p[-S - 3];
  x[12] = p[-3];         x[13] = p[S - 3];      x[14] = p[2 * S - 2];   x[15] = p[3 * S - 1];
  const int v = p[0];
#pragma unroll
  for (int i = 0; i < 16; ++i) x[i] = x[i] * 0xFFFF0001u + 0x00FF0000u;  // low half p, high half 255-p; one IMAD (FMA pipe)
  unsigned a3[16], a9[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a3[i] = __vimin3_u16x2(x[i], x[(i + 1) & 15], x[(i + 2) & 15]);
#pragma unroll
  for (int i = 0; i < 16; ++i) a9[i] = __vimin3_u16x2(a3[i], a3[(i + 3) & 15], a3[(i + 6) & 15]);
  unsigned m0 = __vimax3_u16x2(a9[0], a9[1], a9[2]);
  const unsigned m1 = __vimax3_u16x2(a9[3], a9[4], a9[5]);
  const unsigned m2 = __vimax3_u16x2(a9[6], a9[7], a9[8]);
  unsigned m3 = __vimax3_u16x2(a9[9], a9[10], a9[11]);
  const unsigned m4 = __vimax3_u16x2(a9[12], a9[13], a9[14]);
  m0 = __vimax3_u16x2(m0, m1, m2);
  m3 = __vimax3_u16x2(m3, m4, a9[15]);
  const unsigned mm = __vimax3_u16x2(m0, m3, m3);
  const int bright = (int)(mm & 0xFFFFu) - v;  // max over arcs of (min ring) - v
  const int dark = (int)(mm >> 16) - 255 + v;  // v - min over arcs of (max ring)
  return max(max(bright, dark), 0);
}

// bit 7 of a byte is set iff that byte of x or of y is > t (K = (127 - t) * 0x01010101, t <= 127); other bits: garbage
__device__ __forceinline__ unsigned add_fma(unsigned a, unsigned b, unsigned one);
__device__ __forceinline__ unsigned swar_gt2(unsigned x, unsigned y, unsigned K, unsigned one) {
  return add_fma(x & 0x7F7F7F7Fu, K, one) | x | add_fma(y & 0x7F7F7F7Fu, K, one) | y;
}
// survivor masks: row r of a strip keeps its four pixel flags (bit 7 of each byte of m) in acc[r >> 3], pixel k at bit
// 8 * k + (r & 7): one shift and one AND-OR per row (a 4-bit nibble per row took a multiply, a shift and two logic ops)
__device__ __forceinline__ void put_row(unsigned (&acc)[FD_ACC], int r, unsigned m) {
  const int q = r & 7;  // compile-time after unrolling
  acc[r >> 3] |= (q == 7) ? (m & 0x80808080u) : ((m >> (7 - q)) & (0x01010101u << q));
}
__device__ __forceinline__ constexpr unsigned row_bit(int r, int k) { return 1u << (8 * k + (r & 7)); }

// exclusive prefix over the lanes of the per-lane set-bit counts; returns the warp total
__device__ __forceinline__ int mask_scan(const unsigned (&acc)[FD_ACC], int lane, int* excl) {
  int cnt = 0;
#pragma unroll
  for (int q = 0; q < FD_ACC; ++q) cnt += __popc(acc[q]);
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  *excl = incl - cnt;
  return __shfl_sync(0xffffffffu, incl, 31);
}

// a + C as IMAD (a * one + C, `one` = 1 from a kernel parameter so that ptxas keeps the multiply): the kernel is bound by
// the ALU pipe (stall_math / not_selected everywhere in the ncu source view) while the FMA pipe that runs IMAD idles
template <unsigned C>
__device__ __forceinline__ unsigned add_imm_fma(unsigned a, unsigned one) {
  unsigned d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "n"(C));
  return d;
}
__device__ __forceinline__ unsigned add_fma(unsigned a, unsigned b, unsigned one) {
  unsigned d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
  return d;
}

// the whole-list flatten (no window): a running shared-memory byte address instead of an index, and both per-entry
// additions (entry value, address bump) on the FMA pipe: STS + 2 IMAD per slot.  Slot S = row S/4, column bit S%4.
template <int S>
__device__ __forceinline__ void flatten_slot(const unsigned (&acc)[FD_ACC], unsigned col, unsigned& p, unsigned one) {
  if constexpr (S < FD_ROWS * 4) {
    constexpr int r = S >> 2, k = S & 3;
    if (acc[r >> 3] & row_bit(r, k)) {
      const unsigned v = add_imm_fma<(unsigned)((r << 7) + k)>(col, one);
      asm volatile("st.shared.u16 [%0], %1;" ::"r"(p), "h"((unsigned short)v) : "memory");
      p = add_imm_fma<2u>(p, one);
    }
    flatten_slot<S + 1>(acc, col, p, one);
  }
}
__device__ __forceinline__ void flatten_all(const unsigned (&acc)[FD_ACC], int lane, int excl, uint16_t* list, unsigned one) {
  unsigned p = (unsigned)__cvta_generic_to_shared(list) + 2u * (unsigned)excl;
  flatten_slot<0>(acc, (unsigned)lane * 4u, p, one);
  ORBFE_DCHECK(p <= (unsigned)__cvta_generic_to_shared(list) + 2u * FD_CAP);
}

// write the entries (row<<7 | column) whose flat index lies in [lo, lo+FD_CAP) into list[index - lo]; WINDOWED=false
// when the whole list fits (no range check).  Fully unrolled and predicated: no data-dependent loop.
template <bool WINDOWED>
__device__ __forceinline__ void flatten_window(const unsigned (&acc)[FD_ACC], int lane, int excl, int lo, uint16_t* list) {
  int off = excl - lo;
  const int col = lane * 4;
#pragma unroll
  for (int r = 0; r < FD_ROWS; ++r) {
    const unsigned a = acc[r >> 3];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (a & row_bit(r, k)) {
        ORBFE_DCHECK(WINDOWED || (unsigned)off < (unsigned)FD_CAP);
        if (!WINDOWED || (unsigned)off < (unsigned)FD_CAP) list[off] = (uint16_t)((r << 7) | (col + k));
        ++off;
      }
    }
  }
}

// compile-time: the iniTh launch carries none of the open-cell bookkeeping of the minTh launch
#ifndef ORBFE_FAST_MINB
#define ORBFE_FAST_MINB 28
#endif
template <bool SECOND>
__global__ void __launch_bounds__(FD_T, ORBFE_FAST_MINB) fast_detect_kernel(const __grid_constant__ PyrGeom g,
                                                           const uint8_t* __restrict__ pyr,
                                                           uint32_t* __restrict__ cand,
                                                           int32_t* __restrict__ cand_count,
                                                           uint8_t* cell_hi, int32_t* __restrict__ flags,
                                                           int t_run,
                                                           const CUtensorMap* __restrict__ maps, unsigned one) {
  constexpr bool second_pass = SECOND;
  extern __shared__ __align__(128) unsigned char fd_smem_raw[];
  FdSmem* s_all = reinterpret_cast<FdSmem*>(fd_smem_raw);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int item = blockIdx.x * FD_WARPS + wid;
  if (item >= g.fast_strips) return;  // warps are independent: no CTA-wide barrier below
  FdSmem& sm = s_all[wid];
  const int b = blockIdx.y;
  int lvl = 0;
  while (lvl + 1 < g.nlevels && item >= g.lv[lvl + 1].fast_strip_base) ++lvl;
  const LevelGeom& L = g.lv[lvl];
  item -= L.fast_strip_base;
  const int sy = item / L.fast_strips_x, sx = item - sy * L.fast_strips_x;
  const int w = L.w, h = L.h;
  constexpr int kEdge = kMinBorder + 3;             // detection domain: kEdge <= x < w-kEdge, kEdge <= y < h-kEdge
  const int cx0 = kMinBorder - 4 + kFastOwnCols * sx;  // first computed column (word aligned); owned = +4 .. +123
  const int cy0 = kEdge - 1 + kFastOwnRows * sy;       // first computed row; owned = +1 .. +30
  const int x_end = w - kEdge, y_end = h - kEdge;
  const int x4 = cx0 + lane * 4;  // first of the lane's four columns
  const int t_lo = t_run;
  const unsigned K = (unsigned)(127 - min(t_lo, 127)) * 0x01010101u;  // a weaker threshold only passes a superset
  const unsigned inv_w = 0xFFFFFFFFu / (unsigned)L.w_cell + 1u;  // n / d == umulhi(n, inv) for n < 2^32 / d
  const unsigned inv_h = 0xFFFFFFFFu / (unsigned)L.h_cell + 1u;
  uint8_t* hi = cell_hi + (size_t)b * g.total_cells + L.cell_base;
  unsigned open0 = 0xFFFFFFFFu, open1 = 0xFFFFFFFFu;  // pixels (bit 7 of each byte) still looked at: rows < / >= y_split
  int y_split = 0;
  if (second_pass) {
    // the minTh pass only matters for cells without an iniTh corner (:830-837): a strip whose owned pixels all lie
    // in flagged cells has nothing to add
    const int ox0 = max(cx0 + 4, kEdge), ox1 = min(cx0 + 4 + kFastOwnCols, x_end) - 1;
    const int oy0 = max(cy0 + 1, kEdge), oy1 = min(cy0 + 1 + kFastOwnRows, y_end) - 1;
    if (ox1 < ox0 || oy1 < oy0) return;
    const int cj0 = (int)__umulhi((unsigned)(ox0 - kEdge), inv_w), cj1 = (int)__umulhi((unsigned)(ox1 - kEdge), inv_w);
    const int ci0 = (int)__umulhi((unsigned)(oy0 - kEdge), inv_h), ci1 = (int)__umulhi((unsigned)(oy1 - kEdge), inv_h);
    const int ncj = cj1 - cj0 + 1, ncells = ncj * (ci1 - ci0 + 1);
    bool open = false;
    for (int q = lane; q < ncells; q += 32) {
      const int qi = q / ncj, qj = q - qi * ncj;
      open |= hi[(ci0 + qi) * L.n_cols + cj0 + qj] == 0;
    }
    if (!__any_sync(0xffffffffu, open)) return;  // uniform
    // ... and inside an open strip only the pixels of unflagged cells are looked at (the owned rows span at most two
    // cell rows when cells are at least as tall as a strip; otherwise every pixel stays)
    if (ci1 - ci0 <= 1) {
      open0 = open1 = 0u;
      y_split = kEdge + ci1 * L.h_cell;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cj = min((int)__umulhi((unsigned)max(x4 + k - kEdge, 0), inv_w), L.n_cols - 1);
        if (hi[ci0 * L.n_cols + cj] == 0) open0 |= 0x80u << (8 * k);
        if (hi[ci1 * L.n_cols + cj] == 0) open1 |= 0x80u << (8 * k);
      }
    }
  }
  unsigned colmask = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (x4 + k >= kEdge && x4 + k < x_end) colmask |= 0x80u << (8 * k);
  // TMA wants a 16-byte aligned start column: the box starts at the aligned column below cx0-4 (0 or 8 bytes earlier;
  // the 144-byte box still covers the 136 bytes the strip reads) and the tile is addressed from `timg`
  const int tma_x = (cx0 - 4) & ~15;
  const uint8_t* timg = sm.img + (cx0 - 4 - tma_x);
  const unsigned* tile_w = reinterpret_cast<const unsigned*>(timg);
  unsigned* score_w = reinterpret_cast<unsigned*>(sm.score);

  // ---- 0. stage rows cy0-3 .. cy0+18, columns tma_x .. tma_x+143 in shared memory: one TMA tile load per strip ----
  if (lane == 0) tma_load_tile(sm.img, maps + lvl, tma_x, cy0 - 3, b, &sm.bar);
  __syncwarp();
  mbar_wait_phase0(&sm.bar);

  // ---- 1. dense quick reject (branch-free) ----
  unsigned c[7];   // centre words of the last 7 rows
  unsigned hf[4];  // horizontal-pair flags of the last 4 rows
  unsigned acc[FD_ACC];  // survivor bits: pixel k of row r at bit 8 * k + (r & 7) of acc[r >> 3] (put_row)
#pragma unroll
  for (int q = 0; q < FD_ACC; ++q) acc[q] = 0u;
#pragma unroll
  for (int j = 0; j < FD_TROWS; ++j) {
    const unsigned w1 = tile_w[j * FD_TW + lane + 1];
    c[j % 7] = w1;
    if (j >= 3 && j < FD_ROWS + 3) {
      const unsigned w0 = tile_w[j * FD_TW + lane], w2 = tile_w[j * FD_TW + lane + 2];
      const unsigned pl = __funnelshift_r(w0, w1, 8);   // pixels x-3
      const unsigned pr = __funnelshift_r(w1, w2, 24);  // pixels x+3
      hf[j % 4] = swar_gt2(__vabsdiffu4(w1, pl), __vabsdiffu4(w1, pr), K, one);
    }
    if (j >= 6) {
      const int r = j - 6, y = cy0 + r;
      score_w[r * 32 + lane] = 0u;  // (four 16-byte stores per lane up front instead: 0.1594 vs 0.1544 ms)
      // (rows outside the detection domain are cleared after the loop: only a level's first / last strip row has any)
      const unsigned rm = second_pass ? (colmask & (y < y_split ? open0 : open1)) : colmask;
      const unsigned cc = c[(j - 3) % 7];
      const unsigned vf = swar_gt2(__vabsdiffu4(cc, c[(j - 6) % 7]), __vabsdiffu4(cc, c[j % 7]), K, one);
      put_row(acc, r, vf & hf[(j - 3) % 4] & rm);
    }
  }
  {
    // valid computed rows r: kEdge <= cy0 + r < y_end
    const int r_lo = max(kEdge - cy0, 0), r_hi = min(y_end - cy0, FD_ROWS);
#pragma unroll
    for (int q = 0; q < FD_ACC; ++q) {
      const int lo = min(max(r_lo - 8 * q, 0), 8), hi = min(max(r_hi - 8 * q, 0), 8);
      // bits lo .. hi-1 of every byte
      const unsigned m = hi > lo ? (((0xFFu >> (8 - hi)) & (0xFFu << lo)) & 0xFFu) * 0x01010101u : 0u;
      acc[q] &= m;
    }
  }
  int excl;
  int total = mask_scan(acc, lane, &excl);
  if (total == 0) return;  // uniform
  __syncwarp();

  uint32_t* out = cand + (size_t)b * g.cand_total + L.cand_off;
  int32_t* out_n = cand_count + b * kMaxLevels + lvl;
  const unsigned lt = (1u << lane) - 1u;

  // 3. cell-local 3x3 suppression of list[0..n) (owned pixels that have a score); the maxima compact in place, then
  // ONE atomicAdd reserves their slots in the level's candidate list
  auto suppress_emit = [&](int n) {
    int nmax = 0;
    for (int base = 0; base < n; base += 32) {
      const int i = base + lane;
      bool keep = false;
      int e = 0;
      if (i < n) {
        ORBFE_DCHECK(i < FD_CAP);
        e = sm.list[i];
        const int r = e >> 7, cx = e & 127;
        const int x = cx0 + cx, y = cy0 + r;
        ORBFE_DCHECK(r >= 1 && r <= FD_ROWS - 2 && cx >= 1 && cx <= 126);  // the 8 neighbours are inside the score tile
        ORBFE_DCHECK(x >= kEdge && x < x_end && y >= kEdge && y < y_end);
        const int cj = (int)__umulhi((unsigned)(x - kEdge), inv_w), ci = (int)__umulhi((unsigned)(y - kEdge), inv_h);
        ORBFE_DCHECK(cj >= 0 && cj < L.n_cols && ci >= 0 && ci < L.n_rows);
        // detection window of the pixel's cell (:803-818 minus FAST's 3-px margin)
        const int wx0 = kEdge + cj * L.w_cell, wx1 = min(wx0 + L.w_cell, x_end);
        const int wy0 = kEdge + ci * L.h_cell, wy1 = min(wy0 + L.h_cell, y_end);
        const bool xl = x > wx0, xr = x + 1 < wx1, yu = y > wy0, yd = y + 1 < wy1;
        const uint8_t* s = sm.score + r * 128 + cx;
        const int sc = s[0];
        // max over the neighbours inside the window (all 8 are inside the computed tile: owned pixels are inset)
        int nb = max(xl ? (int)s[-1] : 0, xr ? (int)s[1] : 0);
        const int u0 = s[-129], u1 = s[-128], u2 = s[-127], d0 = s[127], d1 = s[128], d2 = s[129];
        const int un = max(u1, max(xl ? u0 : 0, xr ? u2 : 0)), dn = max(d1, max(xl ? d0 : 0, xr ? d2 : 0));
        nb = max(nb, max(yu ? un : 0, yd ? dn : 0));
        keep = sc > nb;
        if (second_pass && keep) keep = hi[ci * L.n_cols + cj] == 0;  // flagged cells keep their iniTh set (first pass)
      }
      const unsigned mk = __ballot_sync(0xffffffffu, keep);
      __syncwarp();  // this chunk has been read by every lane before slots <= base+31 are overwritten
      ORBFE_DCHECK(!keep || nmax + __popc(mk & lt) <= i);  // in-place compaction never runs ahead of the reads
      if (keep) sm.list[nmax + __popc(mk & lt)] = (uint16_t)e;
      nmax += __popc(mk);
    }
    if (nmax == 0) return;
    __syncwarp();
    int ob = 0;
    if (lane == 0) ob = atomicAdd(out_n, nmax);
    ob = __shfl_sync(0xffffffffu, ob, 0);
    for (int i = lane; i < nmax; i += 32) {
      const int e = sm.list[i];
      const int r = e >> 7, cx = e & 127;
      const int x = cx0 + cx, y = cy0 + r;
      const int sc = sm.score[r * 128 + cx];
      if (ob + i < L.cand_cap) out[ob + i] = pack_cand(x - kMinBorder, y - kMinBorder, sc);
      else atomicOr(&flags[b], FLAG_CAND_OVERFLOW);
      if (!second_pass && sc >= g.ini_th) {
        const int cj = (int)__umulhi((unsigned)(x - kEdge), inv_w), ci = (int)__umulhi((unsigned)(y - kEdge), inv_h);
        hi[ci * L.n_cols + cj] = 1;
      }
    }
  };

  if (total <= FD_CAP) {
    // ---- 2. exact corner strength of the survivors -> score tile; owned corners compact in place ----
    // (a data-dependent walk over the set bits only, ffs + clear-lowest, measured slower in either launch:
    // 0.203 vs 0.195 ms in the iniTh launch, 0.177 vs 0.172 ms when used in the minTh launch only)
    flatten_all(acc, lane, excl, sm.list, one);
    __syncwarp();
    int nc = 0;
    for (int base = 0; base < total; base += 32) {
      const int i = base + lane;
      bool corner = false;
      int e = 0;
      if (i < total) {
        e = sm.list[i];
        const int r = e >> 7, cx = e & 127;
        ORBFE_DCHECK(i < FD_CAP && r >= 0 && r < FD_ROWS && cx >= 0 && cx < 128);
        ORBFE_DCHECK(timg + (r + 6) * FD_TS + cx + 7 < sm.img + sizeof(sm.img) && timg + r * FD_TS + cx + 1 >= sm.img);
        const int m = corner_strength(timg + (r + 3) * FD_TS + cx + 4);
        if (m > t_lo) {
          sm.score[r * 128 + cx] = (uint8_t)(m - 1);
          corner = m > 1 && r >= 1 && r <= kFastOwnRows && cx >= 4 && cx < 4 + kFastOwnCols;  // score 0 never wins
        }
      }
      const unsigned mk = __ballot_sync(0xffffffffu, corner);
      __syncwarp();  // this chunk has been read by every lane before slots <= base+31 are overwritten
      ORBFE_DCHECK(!corner || nc + __popc(mk & lt) <= i);
      if (corner) sm.list[nc + __popc(mk & lt)] = (uint16_t)e;
      nc += __popc(mk);
    }
    __syncwarp();
    suppress_emit(nc);
    return;
  }

  // ---- dense strips (more survivors than the list holds): same steps in windows of FD_CAP entries ----
  for (int lo = 0; lo < total; lo += FD_CAP) {
    flatten_window<true>(acc, lane, excl, lo, sm.list);
    __syncwarp();
    const int nwin = min(FD_CAP, total - lo);
    for (int i = lane; i < nwin; i += 32) {
      const int e = sm.list[i];
      const int r = e >> 7, cx = e & 127;
      const int m = corner_strength(timg + (r + 3) * FD_TS + cx + 4);
      if (m > t_lo) sm.score[r * 128 + cx] = (uint8_t)(m - 1);
    }
    __syncwarp();
  }
#pragma unroll
  for (int q = 0; q < FD_ACC; ++q) acc[q] = 0u;
  if (lane >= 1 && lane <= 30) {  // owned columns 4 .. 123
#pragma unroll
    for (int r = 1; r <= kFastOwnRows; ++r) {
      const unsigned v = score_w[r * 32 + lane];
      put_row(acc, r, ((v & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | v);  // bytes != 0
    }
  }
  total = mask_scan(acc, lane, &excl);
  for (int lo = 0; lo < total; lo += FD_CAP) {
    flatten_window<true>(acc, lane, excl, lo, sm.list);
    __syncwarp();
    suppress_emit(min(FD_CAP, total - lo));
    __syncwarp();
  }
}

int configure_fast_kernel() {
  const int bytes = (int)(FD_WARPS * sizeof(FdSmem));
  return (cudaFuncSetAttribute(fast_detect_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) == cudaSuccess &&
          cudaFuncSetAttribute(fast_detect_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) == cudaSuccess)
             ? 0 : -1;
}

void launch_fast_cells(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s) {
  if (g.fast_strips <= 0) return;
  cudaMemsetAsync(a.cell_hi, 0, (size_t)batch * g.total_cells, s);
  dim3 grid((g.fast_strips + FD_WARPS - 1) / FD_WARPS, batch);
  const CUtensorMap* maps = static_cast<const CUtensorMap*>(a.fast_maps);
  // cv::FAST(iniTh) everywhere, then cv::FAST(minTh) where a cell came back empty (:830-837): the second launch
  // leaves every strip whose cells all have an iniTh corner at once
  fast_detect_kernel<false><<<grid, FD_T, FD_WARPS * sizeof(FdSmem), s>>>(g, a.pyr, a.cand, a.cand_count, a.cell_hi,
                                                                          a.flags, g.ini_th, maps, 1u);
  if (g.min_th < g.ini_th)
    fast_detect_kernel<true><<<grid, FD_T, FD_WARPS * sizeof(FdSmem), s>>>(g, a.pyr, a.cand, a.cand_count, a.cell_hi,
                                                                           a.flags, g.min_th, maps, 1u);
}

}  // namespace orbfe
