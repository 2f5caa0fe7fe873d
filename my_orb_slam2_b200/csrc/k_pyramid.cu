// k_pyramid.cu — ComputePyramid (reference: src/ORBextractor.cc:1129-1154).
//
// Level l is cv::resize(level l-1, INTER_LINEAR) in OpenCV's 11-bit fixed point
// (SURVEY.md A.1); the 19 px copyMakeBorder frame is not materialised (A.5).
// Coefficients come from per-level tables built on the host exactly as OpenCV
// builds them.
//
// pyramid_level_kernel (the normal path, scale < ~1.3): a thread produces 4
// adjacent destination pixels x 8 destination rows.  The horizontal pass of a
// source row is 3 aligned word loads (the 4 pixels' taps lie within 9 bytes of
// an aligned address), and per pixel one funnel shift that brings its two taps
// into the low bytes + one DP2A with the packed (a0, a1) weights; it is computed
// once per source row and reused by the destination rows that share it (the
// vertical pass walks down the column: consecutive rows share a source row).
// The vertical pass is two IMAD.HI per pixel.  One 32-bit store per 4 pixels.
// pyramid_level_generic_kernel: byte loads through the tables, any scale factor.
#include <cuda.h>

#include "kernels.h"

namespace orbfe {

__global__ void __launch_bounds__(256) pyramid_level_generic_kernel(const __grid_constant__ PyrGeom g, uint8_t* __restrict__ pyr,
                                                            const XTab* __restrict__ xtab,
                                                            const YTab* __restrict__ ytab, int lvl) {
  const LevelGeom& D = g.lv[lvl];
  const LevelGeom& S = g.lv[lvl - 1];
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int y = blockIdx.y * blockDim.y + threadIdx.y;
  const int b = blockIdx.z;
  if (x4 >= D.w || y >= D.h) return;
  const uint8_t* src = pyr + S.lvl_base + (size_t)b * S.lvl_bytes;
  uint8_t* dst = pyr + D.lvl_base + (size_t)b * D.lvl_bytes;
  const YTab yt = ytab[D.ytab_off + y];
  const uint8_t* r0 = src + (size_t)yt.sy0 * S.pitch;
  const uint8_t* r1 = src + (size_t)yt.sy1 * S.pitch;
  const int b0 = yt.b0, b1 = yt.b1;
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int x = x4 + i;
    if (x < D.w) {
      const XTab xt = xtab[D.xtab_off + x];
      const int h0 = (int)__ldg(r0 + xt.sx0) * xt.a0 + (int)__ldg(r0 + xt.sx1) * xt.a1;
      const int h1 = (int)__ldg(r1 + xt.sx0) * xt.a0 + (int)__ldg(r1 + xt.sx1) * xt.a1;
      const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
      out |= (uint32_t)(v & 0xFF) << (8 * i);
    }
  }
  uint8_t* d = dst + (size_t)y * D.pitch + x4;
  if (x4 + 3 < D.w || x4 + 3 < D.pitch) {
    *reinterpret_cast<uint32_t*>(d) = out;  // pitch is a multiple of 128: the padding bytes may be written
  } else {
    for (int i = 0; i < 4 && x4 + i < D.w; ++i) d[i] = (uint8_t)(out >> (8 * i));
  }
}

constexpr int PY_WARPS = 4;  // warps per CTA, stacked vertically: a CTA covers 128 x (4 * PY_ROWS) destination pixels
// PY_ROWS = destination rows per thread: 8 in throughput mode (source rows are reused by the rows that share them), 2
// when the launch would not fill the GPU (single frames: shorter dependent chains per thread, 4x more CTAs).
// TILE: the CTA's whole source window arrives in shared memory by ONE TMA tile load (cp.async.bulk.tensor, issued by
// one thread; round 1 used up to six 16-byte cp.async per thread, whose address arithmetic was a fifth of the kernel's
// instructions), so a thread waits for HBM once instead of once per source row (the row-by-row variant keeps one
// 154-byte row per warp in flight: 0.9 TB/s); needs the source window of the CTA to fit (LevelGeom::resize_tile,
// checked on the host).
constexpr int PT_STRIDE = kPyrTileW;      // bytes per staged source row
constexpr int PT_ROWS = kPyrTileH;        // staged source rows (32 destination rows)

template <int PY_ROWS, bool TILE>
__global__ void __launch_bounds__(PY_WARPS * 32) pyramid_level_kernel(const __grid_constant__ PyrGeom g,
                                                                      uint8_t* __restrict__ pyr,
                                                                      const XWord* __restrict__ xword,
                                                                      const YTab* __restrict__ ytab, int lvl,
                                                                      const CUtensorMap* __restrict__ maps) {
  static_assert(!TILE || PY_ROWS * PY_WARPS == 32, "PT_ROWS is sized for 32 destination rows");
  __shared__ __align__(128) uint8_t s_src[TILE ? PT_ROWS * PT_STRIDE : 16];
  __shared__ unsigned long long s_bar;
  const LevelGeom& D = g.lv[lvl];
  const LevelGeom& S = g.lv[lvl - 1];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int x4 = (blockIdx.x * 32 + lane) * 4;
  const int y0 = (blockIdx.y * PY_WARPS + wid) * PY_ROWS;
  const int b = blockIdx.z;
  const XWord* xw = xword + D.xw_off;
  const unsigned spitch = (unsigned)S.pitch, dpitch = (unsigned)D.pitch;
  const uint8_t* src = pyr + S.lvl_base + (size_t)b * S.lvl_bytes;
  int c0 = 0, s_first = 0;
  if (TILE) {
    // the CTA's source window: PT_STRIDE bytes from the 16-byte aligned column below its first tap, PT_ROWS rows from
    // its first source row — ONE TMA tile load by one thread (rows / columns past the level are zero-filled and only
    // feed pixels that are dropped); the table loads below overlap its flight
    const int xc = blockIdx.x * 128, yc = blockIdx.y * (PY_WARPS * PY_ROWS);
    c0 = __ldg(&xw[xc >> 2].ws) & ~15;
    s_first = __ldg(&ytab[D.ytab_off + yc].sy0);
    if (threadIdx.x == 0) {
      const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar);
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(bar) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(PT_ROWS * PT_STRIDE) : "memory");
      asm volatile(
          "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n" ::
              "r"((unsigned)__cvta_generic_to_shared(s_src)), "l"(maps + (lvl - 1)), "r"(c0), "r"(s_first), "r"(b), "r"(bar)
          : "memory");
    }
  }
  const bool active = x4 < D.w && y0 < D.h;
  // per-pixel taps: bytes o, o+1 of the 12-byte window at ws (o <= 7, checked on the host: LevelGeom::resize_fast).
  // Two byte permutes bring them into the low half-word: PRMT(w0, w1, selA) takes byte o and (o < 7) byte o+1 of the
  // first two words, PRMT(t, w2, selB) replaces byte 1 by the first byte of the third word when o == 7.
  // (selectors: only bits 0..7 of a PRMT selector register matter for the two bytes kept, and the packed neighbours
  // above them are valid selector nibbles, so the records' bytes are used after a plain shift)
  const uint4* rec = reinterpret_cast<const uint4*>(xw + min(x4, D.w - 1) / 4);
  const uint4 q0 = __ldg(rec), q1 = __ldg(rec + 1);
  const int ws = (int)q0.x;
  const unsigned a01[4] = {q1.x, q1.y, q1.z, q1.w};
  unsigned selA[4], selB[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    selA[k] = q0.y >> (8 * k);
    selB[k] = q0.z >> (8 * k);
  }
  if (TILE) {
    __syncthreads();  // the barrier's initialisation is visible to every waiter
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "PY_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n"
        "@p bra PY_DONE;\n"
        "bra PY_WAIT;\n"
        "PY_DONE:\n"
        "}\n" ::"r"((unsigned)__cvta_generic_to_shared(&s_bar))
        : "memory");
  }
  if (!active) return;
  const uint8_t* sbase = TILE ? s_src + (ws - c0) : src + ws;
  uint8_t* dp = pyr + D.lvl_base + (size_t)b * D.lvl_bytes + (size_t)y0 * dpitch + x4;
  if (!TILE) asm volatile("" : "+l"(sbase));  // keep the row base in a register pair: rebuilt per source row otherwise
  auto hrow = [&](int sy, unsigned (&h)[4]) {
    unsigned w0, w1, w2;
    if (TILE) {
      ORBFE_DCHECK(sy >= s_first && sy - s_first < PT_ROWS && ws - c0 >= 0 && ws - c0 + 12 <= PT_STRIDE);
      const unsigned* p = reinterpret_cast<const unsigned*>(sbase + (sy - s_first) * PT_STRIDE);
      w0 = p[0]; w1 = p[1]; w2 = p[2];
    } else {
      const unsigned* p = reinterpret_cast<const unsigned*>(sbase + (size_t)((unsigned)sy * spitch));
      w0 = __ldg(p); w1 = __ldg(p + 1); w2 = __ldg(p + 2);
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned taps = __byte_perm(__byte_perm(w0, w1, selA[k]), w2, selB[k]);  // bytes: S[sx0], S[sx0+1]
      h[k] = __dp2a_lo(a01[k], taps, 0u) >> 4;                                       // (S0*a0 + S1*a1) >> 4
    }
  };
  int cur0 = -1, cur1 = -1;
  unsigned h0[4] = {0u, 0u, 0u, 0u}, h1[4] = {0u, 0u, 0u, 0u};
  const int4* yt = reinterpret_cast<const int4*>(ytab + D.ytab_off + y0);
  const int rows = min(PY_ROWS, D.h - y0);
#pragma unroll 1
  for (int r = 0; r < rows; ++r) {
    const int4 t = __ldg(yt + r);  // sy0, sy1, b0 | b1 << 16, pad
    const int sy0 = t.x, sy1 = t.y;
    const unsigned b01 = (unsigned)t.z;
    // all conditions below are warp-uniform (they depend on y only)
    if (sy0 != cur0) {
      if (sy0 == cur1) {
#pragma unroll
        for (int k = 0; k < 4; ++k) h0[k] = h1[k];
      } else {
        hrow(sy0, h0);
      }
      cur0 = sy0;
    }
    if (sy1 != cur1) {
      if (sy1 == cur0) {
#pragma unroll
        for (int k = 0; k < 4; ++k) h1[k] = h0[k];
      } else {
        hrow(sy1, h1);
      }
      cur1 = sy1;
    }
    const unsigned b0s = b01 << 16, b1s = b01 & 0xFFFF0000u;  // b << 16: umulhi(b << 16, h) == (b * h) >> 16
    unsigned out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const unsigned v = (__umulhi(b0s, h0[k]) + __umulhi(b1s, h1[k]) + 2u) >> 2;
      out |= v << (8 * k);  // (two 16-bit fields per register + one PRMT was measured slower: 0.1032 vs 0.1008 ms)
    }
    *reinterpret_cast<unsigned*>(dp) = out;  // pitch is a multiple of 128: the padding bytes may be written
    dp += dpitch;
  }
}

void launch_pyramid_level(const PyrGeom& g, const ExtractorArena& a, int lvl, int batch, cudaStream_t s) {
  const LevelGeom& D = g.lv[lvl];
  const CUtensorMap* maps = static_cast<const CUtensorMap*>(a.pyr_maps);
  if (D.resize_fast) {
    const int cols = (D.w + 127) / 128;
    // 8 rows per thread + staged tiles as soon as the launch has two CTAs per SM (the 2-row variant is for single
    // frames: at 64 images it ran levels 6 and 7 in 12.4 / 10.1 us where level 5 took 11.8 us with the 8-row kernel)
    if ((long long)cols * ((D.h + 31) / 32) * batch >= 2 * 148) {
      dim3 grid(cols, (D.h + 8 * PY_WARPS - 1) / (8 * PY_WARPS), batch);
      if (D.resize_tile) pyramid_level_kernel<8, true><<<grid, PY_WARPS * 32, 0, s>>>(g, a.pyr, a.xword, a.ytab, lvl, maps);
      else pyramid_level_kernel<8, false><<<grid, PY_WARPS * 32, 0, s>>>(g, a.pyr, a.xword, a.ytab, lvl, maps);
    } else {
      dim3 grid(cols, (D.h + 2 * PY_WARPS - 1) / (2 * PY_WARPS), batch);
      pyramid_level_kernel<2, false><<<grid, PY_WARPS * 32, 0, s>>>(g, a.pyr, a.xword, a.ytab, lvl, maps);
    }
    return;
  }
  dim3 block(32, 8);
  dim3 grid((D.w + 127) / 128, (D.h + 7) / 8, batch);
  pyramid_level_generic_kernel<<<grid, block, 0, s>>>(g, a.pyr, a.xtab, a.ytab, lvl);
}

}  // namespace orbfe
