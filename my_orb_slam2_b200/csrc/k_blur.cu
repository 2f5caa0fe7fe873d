// k_blur.cu — GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) of every pyramid level (reference:
// src/ORBextractor.cc:1107-1108; OpenCV 4.13 fixed-point semantics):
//   kernel (8.8 fixed point) [18 34 48 56 48 34 18], h = sum k*p (exact, <= 65280), v = sum k*h,
//   dst = (v + 32768) >> 16.
// The reference blurs whole levels; so does this kernel (2000 keypoints x 37x37 windows would blur MORE pixels than
// the whole pyramid has), but in one launch over all levels and images and without any intermediate buffer:
//   * a thread owns a 4-pixel-wide, 32-row-tall strip; per source row it loads the three aligned words that cover
//     columns x-4..x+7, forms the four 7-tap horizontal sums with funnel shifts + two DP4A each, and keeps the last
//     seven rows of sums in registers (sliding window) for the vertical pass;
//   * one 32-bit store per 4 output pixels; reads are served by L1/L2 (each source word is read by 3 lanes and each
//     source row by the strips above/below only at the 6 halo rows); the vertical pass works on PAIRS of rows (two
//     16-bit sums per register): three IDP.2A + one IMAD per pixel;
//   * reflect-101 at the level edges: rows by index reflection, columns by byte shuffles (left edge) or a byte-wise
//     slow path (last words of a row).
// HBM-bound by design: reads each level once, writes it once.
#include "kernels.h"

namespace orbfe {

constexpr int BL_TX = 32;  // a CTA is BL_TY warps stacked vertically and covers 128 x (BL_TY * BL_ROWS) pixels: 128 x 32 either as
                           // one warp of 32-row strips (throughput) or as four warps of 8-row strips (few images)
#ifndef ORBFE_BL_PF
#define ORBFE_BL_PF 3
#endif
constexpr int BL_PF = ORBFE_BL_PF;                  // rows of loads in flight per thread

__device__ __forceinline__ int reflect101i(int p, int len) {
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  return p;
}

// vertical pass of one pixel: sum k*h + 32768 (the rounding term rides in the multiply-add chain).  The ALU pipe
// (adds, shifts, permutes) is the busier one in this kernel, so only the pair sums are adds; everything else is IMAD.
__device__ __forceinline__ unsigned mad_u32(unsigned a, unsigned b, unsigned c) {
  unsigned d;
  asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));  // kept a multiply-add: no refactoring into adds
  return d;
}
__device__ __forceinline__ unsigned vsum(unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned a4, unsigned a5,
                                         unsigned a6) {
#ifndef BL_PAIRS
#define BL_PAIRS 2
#endif
  // BL_PAIRS of the three symmetric pairs are summed first (adds, ALU pipe); the others are two multiply-adds each
  // (FMA pipe).  Measured: 3 pairs 0.1088 ms, 2 pairs 0.1055 ms, 0 pairs 0.1079 ms
  unsigned t = mad_u32(a3, 56u, 32768u);
  if (BL_PAIRS >= 1) t = mad_u32(a2 + a4, 48u, t); else { t = mad_u32(a2, 48u, t); t = mad_u32(a4, 48u, t); }
  if (BL_PAIRS >= 2) t = mad_u32(a1 + a5, 34u, t); else { t = mad_u32(a1, 34u, t); t = mad_u32(a5, 34u, t); }
  if (BL_PAIRS >= 3) t = mad_u32(a0 + a6, 18u, t); else { t = mad_u32(a0, 18u, t); t = mad_u32(a6, 18u, t); }
  return t;
}
// dst = v >> 16 for four pixels: v < 2^24, so the result is byte 2 of v; three byte permutes pack the four
__device__ __forceinline__ unsigned pack_round(const unsigned (&v)[4]) {
  return __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
}

template <int BL_ROWS, int BL_TY>
__global__ void __launch_bounds__(BL_TX * BL_TY) blur_levels_kernel(const __grid_constant__ PyrGeom g,
                                                                    const uint8_t* __restrict__ pyr,
                                                                    uint8_t* __restrict__ blur) {
  int tile = blockIdx.x;
  int lvl = 0;
  constexpr bool kTall = BL_ROWS * BL_TY == 128;  // 128 x 128 tiles (geometry: blur_tile_base), else 128 x 32 (_small)
  const int n_tiles = kTall ? g.blur_tiles : g.blur_tiles_small;
  const int b = blockIdx.y;
  // The last one or two words of a row (x4 + 7 >= w) need column reflection: byte-wise loads, a rolled loop, ~5x the
  // instructions of the word path.  Left inside the 128-column tiles they would keep a whole warp busy for one or
  // two lanes; instead the CTAs past n_tiles take them with every lane on its own (row strip, word) job.
  const bool edge_cta = tile >= n_tiles;
  int x4, y0;
  if (!edge_cta) {
    while (lvl + 1 < g.nlevels && tile >= (kTall ? g.lv[lvl + 1].blur_tile_base : g.lv[lvl + 1].blur_tile_base_small)) ++lvl;
    const LevelGeom& L0 = g.lv[lvl];
    tile -= kTall ? L0.blur_tile_base : L0.blur_tile_base_small;
    const int tiles_x = (L0.w + 127) / 128;
    const int ty = tile / tiles_x, tx = tile - ty * tiles_x;
    x4 = (tx * BL_TX + threadIdx.x) * 4;
    y0 = (ty * BL_TY + threadIdx.y) * BL_ROWS;
    if (x4 >= L0.w || y0 >= L0.h || x4 + 7 >= L0.w) return;
  } else {
    int job = (tile - n_tiles) * (BL_TX * BL_TY) + threadIdx.y * BL_TX + threadIdx.x;
    for (; lvl < g.nlevels; ++lvl) {
      const int nj = ((g.lv[lvl].h + BL_ROWS - 1) / BL_ROWS) * 2;
      if (job < nj) break;
      job -= nj;
    }
    if (lvl >= g.nlevels) return;
    x4 = ((max(g.lv[lvl].w - 7, 0) + 3) >> 2) * 4 + 4 * (job & 1);  // the first word with x4 + 7 >= w, or the next one
    y0 = (job >> 1) * BL_ROWS;
    if (x4 >= g.lv[lvl].w) return;
  }
  const LevelGeom& L = g.lv[lvl];
  const uint8_t* src = pyr + L.lvl_base + (size_t)b * L.lvl_bytes;
  uint8_t* dst = blur + L.lvl_base + (size_t)b * L.lvl_bytes;
  const int w = L.w, h = L.h, pitch = L.pitch;
  const bool slow = edge_cta;
  const bool left = (x4 == 0);
  const unsigned K_LO = 0x38302212u;  // 18,34,48,56 (byte 0 first)
  const unsigned K_HI = 0x00122230u;  // 48,34,18,0

  const int rows = min(BL_ROWS, h - y0);  // warp-uniform
  if (slow) {
    // right edge of the level (the last one or two words of a row): byte-wise loads with column reflection, rolled
    // loop; block-uniform branch (edge CTAs only)
    unsigned win[7][4];
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) win[i][j] = 0;
#pragma unroll 1
    for (int r = 0; r < rows + 6; ++r) {
      const int y = reflect101i(y0 + r - 3, h);
      const uint8_t* row = src + (size_t)y * pitch;
      unsigned t[3] = {0, 0, 0};
#pragma unroll
      for (int i = 0; i < 12; ++i) {
        const int c = reflect101i(x4 - 4 + i, w);
        t[i >> 2] |= (unsigned)__ldg(row + c) << (8 * (i & 3));
      }
      const unsigned w0 = t[0], w1 = t[1], w2 = t[2];
      unsigned hs[4];
      hs[0] = __dp4a(__funnelshift_r(w0, w1, 8), K_LO, __dp4a(__funnelshift_r(w1, w2, 8), K_HI, 0u));
      hs[1] = __dp4a(__funnelshift_r(w0, w1, 16), K_LO, __dp4a(__funnelshift_r(w1, w2, 16), K_HI, 0u));
      hs[2] = __dp4a(__funnelshift_r(w0, w1, 24), K_LO, __dp4a(__funnelshift_r(w1, w2, 24), K_HI, 0u));
      hs[3] = __dp4a(w1, K_LO, __dp4a(w2, K_HI, 0u));
#pragma unroll
      for (int i = 0; i < 6; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) win[i][j] = win[i + 1][j];
#pragma unroll
      for (int j = 0; j < 4; ++j) win[6][j] = hs[j];
      if (r >= 6) {
        unsigned v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
          v[j] = vsum(win[0][j], win[1][j], win[2][j], win[3][j], win[4][j], win[5][j], win[6][j]);
        const unsigned out = pack_round(v);
        *reinterpret_cast<unsigned*>(dst + (size_t)(y0 + r - 6) * pitch + x4) = out;
      }
    }
    return;
  }
  // win[r % 7] holds the horizontal sums of loaded row r: with the row loop fully unrolled the window is renamed, not
  // moved (a rolled loop spends a quarter of its instructions on register moves); the words of row r + BL_PF are
  // requested before row r is processed so that several rows of loads are in flight per thread
  const int xl = left ? x4 : x4 - 4;  // at the left edge w0 is rebuilt from w1 (reflection of columns -1..-3)
  if (y0 >= 3 && y0 + BL_ROWS + 3 <= h) {
    // strip with all its 6 halo rows inside the level (warp-uniform): no row reflection, no row-count tests, the
    // source and destination rows are reached by bumping two pointers
    // vertical pass on PAIRS of rows: pr[r % 6][j] = h(r-1) | h(r) << 16 (h <= 65280 fits 16 bits; one multiply-add
    // per row and pixel), and v = 18 h(r-6) + 34 h(r-5) + 48 h(r-4) + 56 h(r-3) + 48 h(r-2) + 34 h(r-1) + 18 h(r) + 32768
    // is three 16-bit x 8-bit two-way dot products (IDP.2A) on the pairs ending at rows r-5, r-3, r-1 plus one multiply-add:
    // 5 instructions per pixel instead of 7 (2 adds + 5 multiply-adds)
    constexpr unsigned W_A = 18u | (34u << 8) | (48u << 16) | (56u << 24), W_B = 48u | (34u << 8);
    unsigned pr[6][4];
    unsigned hprev[4] = {0u, 0u, 0u, 0u};
    unsigned pf[BL_PF][3];
    const uint8_t* rp = src + (size_t)(y0 - 3) * pitch + x4;
    uint8_t* wp = dst + (size_t)y0 * pitch + x4;
    const int dl = xl - x4;  // 0 or -4
#pragma unroll
    for (int r = 0; r < BL_PF; ++r) {
      pf[r][0] = __ldg(reinterpret_cast<const unsigned*>(rp + dl));
      pf[r][1] = __ldg(reinterpret_cast<const unsigned*>(rp));
      pf[r][2] = __ldg(reinterpret_cast<const unsigned*>(rp + 4));
      rp += pitch;
    }
#pragma unroll
    for (int r = 0; r < BL_ROWS + 6; ++r) {
      const unsigned w1 = pf[r % BL_PF][1], w2 = pf[r % BL_PF][2];
      const unsigned w0 = left ? __byte_perm(w1, 0, 0x1230) : pf[r % BL_PF][0];
      if (r + BL_PF < BL_ROWS + 6) {
        pf[r % BL_PF][0] = __ldg(reinterpret_cast<const unsigned*>(rp + dl));
        pf[r % BL_PF][1] = __ldg(reinterpret_cast<const unsigned*>(rp));
        pf[r % BL_PF][2] = __ldg(reinterpret_cast<const unsigned*>(rp + 4));
        rp += pitch;
      }
      unsigned hs[4];
      hs[0] = __dp4a(__funnelshift_r(w0, w1, 8), K_LO, __dp4a(__funnelshift_r(w1, w2, 8), K_HI, 0u));
      hs[1] = __dp4a(__funnelshift_r(w0, w1, 16), K_LO, __dp4a(__funnelshift_r(w1, w2, 16), K_HI, 0u));
      hs[2] = __dp4a(__funnelshift_r(w0, w1, 24), K_LO, __dp4a(__funnelshift_r(w1, w2, 24), K_HI, 0u));
      hs[3] = __dp4a(w1, K_LO, __dp4a(w2, K_HI, 0u));
      if (r >= 6) {
        unsigned v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          unsigned t = mad_u32(hs[j], 18u, 32768u);
          t = __dp2a_lo(pr[(r + 1) % 6][j], W_A, t);  // pair ending at row r-5: 18 h(r-6) + 34 h(r-5)
          t = __dp2a_hi(pr[(r + 3) % 6][j], W_A, t);  // pair ending at row r-3: 48 h(r-4) + 56 h(r-3)
          v[j] = __dp2a_lo(pr[(r + 5) % 6][j], W_B, t);  // pair ending at row r-1: 48 h(r-2) + 34 h(r-1)
        }
        const unsigned out = pack_round(v);
        *reinterpret_cast<unsigned*>(wp) = out;
        wp += pitch;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (r >= 1) pr[r % 6][j] = mad_u32(hs[j], 65536u, hprev[j]);
        hprev[j] = hs[j];
      }
    }
    return;
  }
  // strips touching the top or bottom of the level: row reflection and a row count, window moved instead of renamed
  // (rolled loop: these are the minority of strips and a second unrolled body would double the kernel's code size);
  // the same pair form of the vertical pass: at row r, pr[i] = pair ending at row r-5+i
  {
    constexpr unsigned W_A = 18u | (34u << 8) | (48u << 16) | (56u << 24), W_B = 48u | (34u << 8);
    unsigned pr[5][4], hprev[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      hprev[j] = 0;
#pragma unroll
      for (int i = 0; i < 5; ++i) pr[i][j] = 0;
    }
    unsigned q[3], nq[3];
    auto load_row = [&](int r, unsigned (&o)[3]) {
      const int y = reflect101i(y0 + r - 3, h);
      const uint8_t* row = src + (size_t)y * pitch;
      o[0] = __ldg(reinterpret_cast<const unsigned*>(row + xl));
      o[1] = __ldg(reinterpret_cast<const unsigned*>(row + x4));
      o[2] = __ldg(reinterpret_cast<const unsigned*>(row + x4 + 4));
    };
    load_row(0, q);
#pragma unroll 2
    for (int r = 0; r < rows + 6; ++r) {
      load_row(min(r + 1, rows + 5), nq);
      const unsigned w1 = q[1], w2 = q[2];
      const unsigned w0 = left ? __byte_perm(w1, 0, 0x1230) : q[0];
      unsigned hs[4];
      hs[0] = __dp4a(__funnelshift_r(w0, w1, 8), K_LO, __dp4a(__funnelshift_r(w1, w2, 8), K_HI, 0u));
      hs[1] = __dp4a(__funnelshift_r(w0, w1, 16), K_LO, __dp4a(__funnelshift_r(w1, w2, 16), K_HI, 0u));
      hs[2] = __dp4a(__funnelshift_r(w0, w1, 24), K_LO, __dp4a(__funnelshift_r(w1, w2, 24), K_HI, 0u));
      hs[3] = __dp4a(w1, K_LO, __dp4a(w2, K_HI, 0u));
      if (r >= 6) {
        unsigned v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          unsigned t = mad_u32(hs[j], 18u, 32768u);
          t = __dp2a_lo(pr[0][j], W_A, t);
          t = __dp2a_hi(pr[2][j], W_A, t);
          v[j] = __dp2a_lo(pr[4][j], W_B, t);
        }
        const unsigned out = pack_round(v);
        *reinterpret_cast<unsigned*>(dst + (size_t)(y0 + r - 6) * pitch + x4) = out;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int i = 0; i < 4; ++i) pr[i][j] = pr[i + 1][j];
        pr[4][j] = mad_u32(hs[j], 65536u, hprev[j]);
        hprev[j] = hs[j];
      }
      q[0] = nq[0]; q[1] = nq[1]; q[2] = nq[2];
    }
  }
}

// CTAs that take the right-edge (row strip, word) jobs of one image: 2 jobs per strip row of every level
static int blur_edge_ctas(const PyrGeom& g, int rows, int threads) {
  int jobs = 0;
  for (int l = 0; l < g.nlevels; ++l) jobs += ((g.lv[l].h + rows - 1) / rows) * 2;
  return (jobs + threads - 1) / threads;
}

void launch_blur_levels(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s) {
  if ((long long)g.blur_tiles * batch >= 2 * 148 * 4) {
    // throughput: one-warp CTAs of 32-row strips (a slot frees as soon as its warp is done)
    blur_levels_kernel<32, 1><<<dim3(g.blur_tiles_small + blur_edge_ctas(g, 32, 32), batch), dim3(BL_TX, 1), 0, s>>>(
        g, a.pyr, a.blur);
  } else {  // few images: 8-row strips per thread, four warps per 128 x 32 tile
    blur_levels_kernel<8, 4><<<dim3(g.blur_tiles_small + blur_edge_ctas(g, 8, 128), batch), dim3(BL_TX, 4), 0, s>>>(
        g, a.pyr, a.blur);
  }
}

}  // namespace orbfe
