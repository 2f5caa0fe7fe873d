// k_describe.cu — everything after DistributeOctTree for the kept keypoints.  A CTA takes 32 output rows; after the
// row -> (level, kept record) lookup (one thread per keypoint, one CTA barrier) every warp owns 4 keypoints through all
// steps, with its global loads issued early (moments: the disc boxes of all 4 keypoints in flight at once; descriptor
// patches: two boxes in flight per warp; all of them TMA tile loads that bypass the LSU / L1), because the kernel is
// bound by memory latency, not by issue slots:
//   * the fix-up of src/ORBextractor.cc:859-870 (pt += minBorder, octave, size);
//   * computeOrientation -> IC_Angle (:472-479, :77-104): m10/m01 over the radius-15 disc of the UN-blurred level
//     (lane = row of the disc, DP4A against a weight table), warp-shuffle reduction, cv::fastAtan2 polynomial;
//   * computeOrbDescriptor (:108-147) on the blurred level (k_blur.cu): lane i builds descriptor byte i
//     (8 tests = 16 gathers) from its 32 fp32 pattern values (loaded once per warp, coalesced float4);
//   * the rescale by mvScaleFactor[level] and the level-order concatenation of operator() (:1117-1125).
// Kept keypoints are >= 19 px from every level edge (FAST margin 3 + minBorder 16), the disc has radius 15 and the
// rotated pattern reaches <= 18 px, so no access leaves the level image.
// Float steps that decide bits (angle, sin/cos, the rotated sample coordinates) use the explicitly rounded
// operations of include/orb_math_shared.h so that no FMA is formed.
#include <cuda.h>

#include "../../include/orb_math_shared.h"
#include "kernels.h"

namespace orbfe {

constexpr int DESC_WARPS = 8;
constexpr int DESC_KP = 32;  // keypoints per CTA (4 per warp)

// 256 test pairs x (x0,y0,x1,y1) as fp32 (the reference converts the int pattern to float in the rotation,
// src/ORBextractor.cc:123-131), transposed so that the 32 lanes of a warp read consecutive float4: element
// [q][lane] holds values 4q..4q+3 of lane's 32 values (= tests 8*lane .. 8*lane+7).
struct PatternT { float4 v[8][32]; };
static PatternT make_pattern() {
  static const int p[1024] = {
#include "../../include/orb_pattern_31.inc"
  };
  PatternT t;
  for (int lane = 0; lane < 32; ++lane)
    for (int q = 0; q < 8; ++q) {
      const int* s = p + lane * 32 + q * 4;
      t.v[q][lane] = make_float4((float)s[0], (float)s[1], (float)s[2], (float)s[3]);
    }
  return t;
}
__device__ PatternT d_pattern_t;
// end of each row of the radius-15 disc (umax, :454-469); the disc is symmetric: |u| <= umax[|v|] <=> |v| <= umax[|u|]
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

// IC_Angle as byte dot products: the disc rows are read as aligned words, so a word's four weights depend on the
// keypoint's alignment a = (kx - 15) & 3, the word j of the row and the row |v|: byte t of word j is column
// u = 4j + t - 15 - a, inside the disc iff |u| <= umax[|v|].  x = the four u (s8, 0 outside), y = the four 0/1 flags.
constexpr int IC_PW = 9;  // words per staged disc row: 31 bytes at any alignment fit 36; odd => conflict-free columns
struct IcTab { uint2 w[4][IC_PW][16]; };
__device__ IcTab d_ic_tab;
static IcTab make_ic_tab() {
  static const int umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
  IcTab t;
  for (int a = 0; a < 4; ++a)
    for (int j = 0; j < IC_PW; ++j)
      for (int av = 0; av < 16; ++av) {
        unsigned wu = 0, w1 = 0;
        for (int k = 0; k < 4; ++k) {
          const int u = 4 * j + k - 15 - a;
          if (u >= -umax[av] && u <= umax[av]) {
            wu |= (unsigned)(u & 0xFF) << (8 * k);
            w1 |= 1u << (8 * k);
          }
        }
        t.w[a][j][av] = make_uint2(wu, w1);
      }
  return t;
}

int upload_pattern() {
  const PatternT t = make_pattern();
  const IcTab ic = make_ic_tab();
  return (cudaMemcpyToSymbol(d_pattern_t, &t, sizeof(t)) == cudaSuccess &&
          cudaMemcpyToSymbol(d_ic_tab, &ic, sizeof(ic)) == cudaSuccess) ? 0 : -1;
}

// sum over the 4 bytes of a (unsigned) * b (signed) + c
__device__ __forceinline__ int dp4a_us(unsigned a, unsigned b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)),
               "l"(gmem_src)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

// ---- TMA: the blurred 37-row patch of a keypoint arrives as ONE cp.async.bulk.tensor box (kDescBoxW x kDescBoxH
// bytes from the 16-byte aligned column at or below kx-18), completing on an mbarrier; it does not pass through the
// LSU / L1, which bound this kernel when the patch came in as 13 four-byte cp.async per lane ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_box(void* dst, const void* tmap, int x, int y, int z, unsigned long long* bar,
                                             int bytes) {
  const unsigned b = smem_u32(bar);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(b), "r"(bytes) : "memory");
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n" ::
          "r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(b)
      : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, int bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_issue_box(void* dst, const void* tmap, int x, int y, int z, unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];\n" ::
          "r"(smem_u32(dst)), "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "DS_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DS_DONE;\n"
      "bra DS_WAIT;\n"
      "DS_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// one disc row (12 staged words, the disc's 9 words start at word WO) against the weight table: sum(u * I), sum(I)
template <int WO>
__device__ __forceinline__ void ic_row(const unsigned (&w)[12], const uint2* wt, int& a, int& s1) {
#pragma unroll
  for (int j = 0; j < IC_PW; ++j) {
    const uint2 wgt = wt[j * 16];
    a = dp4a_us(w[WO + j], wgt.x, a);
    s1 = (int)__dp4a(w[WO + j], wgt.y, (unsigned)s1);
  }
}

struct KpSlot {
  int kx, ky, lvl, score;  // level coordinates (minBorder added, :865-866); lvl < 0: no keypoint in this slot
  int m01, m10;
  float ca, sb;            // cos, sin of the orientation
};

constexpr int PBUF = 3072 / 4;  // words per patch buffer: the 64 x 37 blurred box (2368 bytes), a multiple of 512 bytes; a
                                // warp's two buffers also hold its four disc boxes (4 x 1536 bytes, 128-byte aligned)
static_assert(PBUF * 4 >= kDescBoxW * kDescBoxH && PBUF * 4 % 512 == 0, "patch buffer");
constexpr int DISC_STRIDE = 1536;  // bytes between the disc boxes of a warp (kDiscBoxW * kDiscBoxH = 1488, TMA wants 128)
static_assert(DISC_STRIDE >= kDiscBoxW * kDiscBoxH && DISC_STRIDE % 128 == 0, "disc box");
struct __align__(1024) DescSmem {
  unsigned patch_all[DESC_WARPS][2][PBUF];
  uint2 ic_tab[4 * IC_PW * 16];
  KpSlot slot[DESC_KP];
  unsigned long long patch_bar[DESC_WARPS][2];
  unsigned long long disc_bar[DESC_WARPS];
};

__global__ void __launch_bounds__(DESC_WARPS * 32) describe_kernel(const __grid_constant__ PyrGeom g,
                                                                   const uint8_t* __restrict__ pyr,
                                                                   const uint8_t* __restrict__ blur,
                                                                   const uint32_t* __restrict__ kept,
                                                                   const int32_t* __restrict__ kept_count,
                                                                   orbfe_keypoint* __restrict__ out_kp,
                                                                   uint8_t* __restrict__ out_desc,
                                                                   int32_t* __restrict__ out_count,
                                                                   int32_t* __restrict__ flags,
                                                                   const CUtensorMap* __restrict__ blur_maps,
                                                                   const CUtensorMap* __restrict__ disc_maps) {
  // dynamic shared memory (DescSmem, > 48 KB in all): patch / disc buffers first (512-byte aligned for the swizzle)
  extern __shared__ __align__(1024) unsigned char desc_smem_raw[];
  DescSmem& sm = *reinterpret_cast<DescSmem*>(desc_smem_raw);
  KpSlot* slot = sm.slot;
  uint2* ic_tab = sm.ic_tab;
  auto& patch_all = sm.patch_all;
  auto& patch_bar = sm.patch_bar;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.y;
  const int oi0 = blockIdx.x * DESC_KP;  // first output row of this CTA

  // ---- 0. one thread per keypoint: which level does output row oi belong to (levels are concatenated in order
  // 0..L-1, :1096-1126), and its kept record ----
  for (int i = tid; i < 4 * IC_PW * 16; i += DESC_WARPS * 32) ic_tab[i] = (&d_ic_tab.w[0][0][0])[i];
  if (tid < DESC_KP) {
    const int oi = oi0 + tid;
    const int32_t* kc = kept_count + b * kMaxLevels;
    int lvl = -1, first = 0, acc = 0;
    for (int l = 0; l < g.nlevels; ++l) {
      const int c = kc[l];
      if (lvl < 0 && oi < acc + c) { lvl = l; first = acc; }
      acc += c;
    }
    if (oi == 0) {
      out_count[b] = min(acc, g.out_cap);
      if (acc > g.out_cap) atomicOr(&flags[b], FLAG_OUT_OVERFLOW);
    }
    if (oi >= g.out_cap) lvl = -1;
    KpSlot k;
    k.lvl = lvl;
    k.kx = k.ky = k.score = k.m01 = k.m10 = 0;
    k.ca = k.sb = 0.f;
    if (lvl >= 0) {
      const uint32_t c = kept[(size_t)b * g.kept_total + g.lv[lvl].kept_off + (oi - first)];
      k.kx = cand_x(c) + kMinBorder;  // :865-866
      k.ky = cand_y(c) + kMinBorder;
      k.score = cand_score(c);
    }
    slot[tid] = k;
  }
  __syncthreads();
  if (slot[0].lvl < 0) return;  // output rows are dense: an empty first slot means an empty CTA

  // ---- 1. IC_Angle moments (:77-104).  A warp owns keypoints warp, warp+8, warp+16, warp+24 of the CTA through all
  // remaining steps (no CTA barrier after this point).  The 31 disc rows of ALL FOUR keypoints are requested at once
  // with coalesced cp.async word copies (3 rows per warp instruction; one commit group per keypoint), so the warp
  // waits for HBM / L2 once instead of once per keypoint (the kernel is bound by memory latency: ncu showed 65 % of
  // the samples on the waits of this step with one keypoint in flight).  Then lane = ROW v of the disc: 9 words x two
  // DP4A against the weight table give sum(u * I) and sum(I) over the row's part of the disc; m01 = v * sum(I).
  // (Reading the rows straight from global memory, one row per lane, was measured 2x slower: 31 sectors per load.) ----
  constexpr int KPW = DESC_KP / DESC_WARPS;                        // keypoints per warp
  // two patch buffers per warp, 512-byte aligned: the TMA writes the box with the 64-byte swizzle (16-byte chunk
  // index ^= bits 7-8 of the shared address = (row >> 1) & 3), so that the 512 scattered byte gathers of a keypoint
  // spread over all 32 banks: with plain 64-byte rows a warp's 32 gathers cost 4.8 shared-memory wavefronts on
  // average, swizzled 3.1 (simulated over the pattern, all alignments and angles; 3.0 is what 32 random words do)
  // rows ky-18 .. ky+18 of the blurred level from the 16-byte aligned column at or below kx-18: one TMA box
  auto stage_patch = [&](int i, int buf) {
    const KpSlot k = slot[i];
    if (lane == 0 && k.lvl >= 0)
      tma_load_box(patch_all[warp][buf], blur_maps + k.lvl, (k.kx - 18) & ~15, k.ky - 18, b, &patch_bar[warp][buf],
                   kDescBoxW * kDescBoxH);
  };
  if (lane == 0) {
    mbar_init(&patch_bar[warp][0]);
    mbar_init(&patch_bar[warp][1]);
    mbar_init(&sm.disc_bar[warp]);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  }
  // rows ky-15 .. ky+15 of the un-blurred level from the 16-byte aligned column at or below kx-15: one TMA box of
  // 48 x 31 bytes per keypoint, all four of the warp in flight on one mbarrier (round 1 used 11 four-byte cp.async per
  // lane and keypoint: as many L1 wavefronts as the descriptor gathers themselves).  The warp's 4 boxes share its two
  // patch buffers, which the descriptor patches only need afterwards.
  unsigned char* disc_all = reinterpret_cast<unsigned char*>(patch_all[warp][0]);
  if (lane == 0) {
    int nvalid = 0;
#pragma unroll
    for (int q = 0; q < KPW; ++q) nvalid += slot[warp + q * DESC_WARPS].lvl >= 0;
    mbar_expect_tx(&sm.disc_bar[warp], nvalid * (kDiscBoxW * kDiscBoxH));  // nvalid >= 1 is not needed: see the wait
#pragma unroll
    for (int q = 0; q < KPW; ++q) {
      const KpSlot k = slot[warp + q * DESC_WARPS];
      if (k.lvl >= 0)
        tma_issue_box(disc_all + q * DISC_STRIDE, disc_maps + k.lvl, (k.kx - 15) & ~15, k.ky - 15, b, &sm.disc_bar[warp]);
    }
  }
  {
    const int v = lane - 15;
    const int av = lane < 31 ? (v < 0 ? -v : v) : 15;
    int m10[KPW], m01[KPW];
    __syncwarp();
    mbar_wait(&sm.disc_bar[warp], 0u);  // (with no valid keypoint the arrive alone completes the phase)
#pragma unroll
    for (int q = 0; q < KPW; ++q) {
      // lane = ROW v of the disc: its 48 staged bytes are three 16-byte loads (rows are 48 bytes apart: a quarter
      // warp's eight 16-byte chunks fall into eight different bank groups), the disc's 9 words start at word `wo`
      const int off = (slot[warp + q * DESC_WARPS].kx - 15) & 15;
      const int al = off & 3, wo = off >> 2;  // warp-uniform
      ORBFE_DCHECK(av >= 0 && av < 16);
      const uint2* wt = ic_tab + al * (IC_PW * 16) + av;
      int a = 0, s1 = 0;
      if (lane < 31) {
        const uint4* row = reinterpret_cast<const uint4*>(disc_all + q * DISC_STRIDE + lane * kDiscBoxW);
        const uint4 r0 = row[0], r1 = row[1], r2 = row[2];
        const unsigned w[12] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w, r2.x, r2.y, r2.z, r2.w};
        if (wo == 0) ic_row<0>(w, wt, a, s1);
        else if (wo == 1) ic_row<1>(w, wt, a, s1);
        else if (wo == 2) ic_row<2>(w, wt, a, s1);
        else ic_row<3>(w, wt, a, s1);
      }
      m10[q] = a;
      m01[q] = v * s1;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int q = 0; q < KPW; ++q) {
        m10[q] += __shfl_xor_sync(0xffffffffu, m10[q], o);
        m01[q] += __shfl_xor_sync(0xffffffffu, m01[q], o);
      }
    }
    __syncwarp();  // every lane has read the discs
    // the first two descriptor patches travel while the angles are computed; the buffers were last READ through the
    // generic proxy, the TMA writes through the async proxy
    if (lane == 0) asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    stage_patch(warp, 0);
    stage_patch(warp + DESC_WARPS, 1);
    // ---- 2. lane q of the warp finishes keypoint q: angle (cv::fastAtan2, :103), its cos / sin (:112-113, documented
    // deviation 3), and the keypoint record with the rescale of operator() (:1117-1125) ----
    int my10 = 0, my01 = 0;
#pragma unroll
    for (int q = 0; q < KPW; ++q)
      if (lane == q) { my10 = m10[q]; my01 = m01[q]; }
    if (lane < KPW) {
      const int i = warp + lane * DESC_WARPS;
      const KpSlot k = slot[i];
      if (k.lvl >= 0) {
        const LevelGeom& L = g.lv[k.lvl];
        const float angle = orb_fast_atan2_deg((float)my01, (float)my10);
        float ca, sb;
        orb_sincos_deg(angle, &ca, &sb);
        slot[i].ca = ca;
        slot[i].sb = sb;
        orbfe_keypoint o;
        o.x = (float)k.kx;
        o.y = (float)k.ky;
        if (k.lvl != 0) { o.x = __fmul_rn(o.x, L.scale); o.y = __fmul_rn(o.y, L.scale); }
        o.size = L.kp_size;
        o.angle = angle;
        o.response = (float)k.score;
        o.octave = k.lvl;
        o.class_id = -1;
        out_kp[(size_t)b * g.out_cap + oi0 + i] = o;
      }
    }
    __syncwarp();  // slot[].ca / sb of the warp's keypoints are visible to all its lanes
  }

  // ---- 3. computeOrbDescriptor (:108-147) on the blurred level: lane -> descriptor byte (8 tests = 16 gathers).
  // The 37 x 37 patch the rotated pattern can reach is first copied into shared memory with coalesced word copies
  // (3 rows per warp instruction, cp.async, the next keypoint's patch in flight): a gather of 32 scattered bytes then
  // costs ~3 bank-conflict wavefronts instead of ~20 L1 sectors.  The lane's 32 pattern values stay in registers for
  // the warp's 4 keypoints. ----
  float pat[32];
#pragma unroll
  for (int q = 0; q < 8; ++q) {
    const float4 t = d_pattern_t.v[q][lane];
    pat[4 * q] = t.x; pat[4 * q + 1] = t.y; pat[4 * q + 2] = t.z; pat[4 * q + 3] = t.w;
  }
  int buf = 0, n = 0;
#pragma unroll 1
  for (int i = warp; i < DESC_KP; i += DESC_WARPS, buf ^= 1, ++n) {
    const KpSlot k = slot[i];
    if (k.lvl < 0) break;
    mbar_wait(&patch_bar[warp][buf], (unsigned)(n >> 1) & 1u);  // this buffer's (n / 2)-th load
    // Rounding on the FMA pipe instead of 1024 F2I per keypoint on the XU pipe: for |v| < 2^22, v + M (M = 1.5 * 2^23 +
    // 18, an EVEN integer, so ties still go to the even neighbour of v) is exactly M + rint(v), and its float bits are
    // 0x4B400012 + rint(v): the low bits of the sum ARE the patch row / column (18 + r in 0..36).  Byte address of
    // sample (r, c) in the swizzled patch: row * 64 + (col ^ ((row & 6) << 3)), col = 18 + c + (kx - 18) % 16 < 64.
    // Everything constant is folded into `kc` (multiples of 64 except the column shift, so the XOR of bits 4-5 is not
    // disturbed): 32-bit shared-window address arithmetic, wrapping mod 2^32.
    constexpr float kMagic = 12582912.f + 18.f;
    constexpr unsigned kBits = 0x4B400000u;
    const unsigned kc = smem_u32(patch_all[warp][buf]) - kBits * 64u - kBits + (unsigned)((k.kx - 18) & 15);
    const float ca = k.ca, sb = k.sb;
    int val = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x0 = pat[4 * j], y0 = pat[4 * j + 1], x1 = pat[4 * j + 2], y1 = pat[4 * j + 3];
      const unsigned r0 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(x0, sb), __fmul_rn(y0, ca)), kMagic));
      const unsigned c0 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sb)), kMagic));
      const unsigned r1 = __float_as_uint(__fadd_rn(__fadd_rn(__fmul_rn(x1, sb), __fmul_rn(y1, ca)), kMagic));
      const unsigned c1 = __float_as_uint(__fadd_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sb)), kMagic));
      const unsigned a0 = r0 * 64u + ((c0 + kc) ^ ((r0 << 3) & 0x30u));
      const unsigned a1 = r1 * 64u + ((c1 + kc) ^ ((r1 << 3) & 0x30u));
      ORBFE_DCHECK(a0 - smem_u32(patch_all[warp][buf]) < (unsigned)(kDescBoxW * kDescBoxH) &&
                   a1 - smem_u32(patch_all[warp][buf]) < (unsigned)(kDescBoxW * kDescBoxH));
      unsigned t0, t1;
      asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t0) : "r"(a0));
      asm volatile("ld.shared.u8 %0, [%1];" : "=r"(t1) : "r"(a1));
      val |= (t0 < t1) << j;
    }
    out_desc[((size_t)b * g.out_cap + oi0 + i) * 32 + lane] = (uint8_t)val;
    __syncwarp();  // every lane has read this buffer: the keypoint after next may land in it
    if (i + 2 * DESC_WARPS < DESC_KP) {
      if (lane == 0) asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // generic reads before async writes
      stage_patch(i + 2 * DESC_WARPS, buf);
    }
  }
}

void launch_describe(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s) {
  dim3 grid((g.out_cap + DESC_KP - 1) / DESC_KP, batch);
  describe_kernel<<<grid, DESC_WARPS * 32, sizeof(DescSmem), s>>>(
      g, a.pyr, a.blur, a.kept, a.kept_count, a.out_kp, a.out_desc, a.out_count, a.flags,
      static_cast<const CUtensorMap*>(a.desc_maps), static_cast<const CUtensorMap*>(a.disc_maps));
}

int configure_describe_kernel() {
  return cudaFuncSetAttribute(describe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DescSmem)) ==
                 cudaSuccess ? 0 : -1;
}

}  // namespace orbfe
