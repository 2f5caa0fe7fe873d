// k_describe.cu — everything after DistributeOctTree for one kept keypoint, fused,
// one warp per keypoint:
//   * the fix-up of src/ORBextractor.cc:859-870 (pt += minBorder, octave, size);
//   * computeOrientation -> IC_Angle (:472-479, :77-104): m10/m01 over the radius-15
//     disc of the UN-blurred level, warp-shuffle reduction, cv::fastAtan2 polynomial;
//   * GaussianBlur(7x7, sigma 2, REFLECT_101) (:1107-1108) — only the 37x37 window the
//     rotated test pattern can reach is blurred (OpenCV 4.13 fixed point),
//     from a 43x43 patch staged in shared memory, with reflect-101 at the LEVEL image
//     edges exactly as the full-image blur has it (A.5);
//   * computeOrbDescriptor (:108-147): lane i builds descriptor byte i (8 tests);
//   * the rescale by mvScaleFactor[level] and the level-order concatenation of
//     operator() (:1117-1125).
// Float steps that decide bits (angle, sin/cos, the rotated sample coordinates) use the
// explicitly rounded operations of include/orb_math_shared.h so that no FMA is formed.
#include "../../include/orb_math_shared.h"
#include "kernels.h"

namespace orbfe {

constexpr int DESC_WARPS = 4;
constexpr int PR = 21;              // patch radius: 18 (pattern reach) + 3 (blur taps)
constexpr int PW = 2 * PR + 1;      // 43
constexpr int PS = 44;              // patch row stride
constexpr int BRAD = 18;
constexpr int BW = 2 * BRAD + 1;    // 37
constexpr int BS = 40;              // blurred row stride

__constant__ int8_t c_pattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};
// end of each row of the radius-15 disc (umax, :454-469)
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

struct WarpSmem {
  uint8_t patch[PW * PS];     // raw level pixels, rows/cols -21..21 around the keypoint
  uint16_t hbuf[PW * BW];     // horizontal pass, 8.8 fixed point
  uint8_t blur[BW * BS];      // blurred window, rows/cols -18..18
};

__device__ __forceinline__ int reflect101(int p, int len) {
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  return p;
}

__global__ void __launch_bounds__(DESC_WARPS * 32) describe_kernel(const __grid_constant__ PyrGeom g,
                                                                   const uint8_t* __restrict__ pyr,
                                                                   const uint32_t* __restrict__ kept,
                                                                   const int32_t* __restrict__ kept_count,
                                                                   orbfe_keypoint* __restrict__ out_kp,
                                                                   uint8_t* __restrict__ out_desc,
                                                                   int32_t* __restrict__ out_count,
                                                                   int32_t* __restrict__ flags) {
  __shared__ WarpSmem smem[DESC_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int oi = blockIdx.x * DESC_WARPS + warp;  // output row of this keypoint
  // level of this output row: levels are concatenated in order 0..L-1 (:1096-1126)
  const int32_t* kc = kept_count + b * kMaxLevels;
  int lvl = 0, first = 0, total = 0;
  {
    int acc = 0;
    bool found = false;
    for (int l = 0; l < g.nlevels; ++l) {
      const int c = kc[l];
      if (!found && oi < acc + c) { lvl = l; first = acc; found = true; }
      acc += c;
    }
    total = acc;
    if (oi == 0 && lane == 0) {
      out_count[b] = min(total, g.out_cap);
      if (total > g.out_cap) atomicOr(&flags[b], FLAG_OUT_OVERFLOW);
    }
    if (!found || oi >= g.out_cap) return;
  }
  const LevelGeom& L = g.lv[lvl];
  const uint32_t c = kept[(size_t)b * g.kept_total + L.kept_off + (oi - first)];
  const int kx = cand_x(c) + kMinBorder, ky = cand_y(c) + kMinBorder;  // :865-866
  const uint8_t* img = pyr + L.lvl_base + (size_t)b * L.lvl_bytes;
  WarpSmem& S = smem[warp];

  // ---- stage the 43x43 patch (reflect-101 only matters within 2 px of the level edge) ----
  const bool inside = kx - PR >= 0 && kx + PR < L.w && ky - PR >= 0 && ky + PR < L.h;
  if (inside) {
    const uint8_t* src = img + (size_t)(ky - PR) * L.pitch + (kx - PR);
    for (int i = lane; i < PW * PW; i += 32) {
      const int r = i / PW, cc = i - r * PW;
      S.patch[r * PS + cc] = __ldg(src + (size_t)r * L.pitch + cc);
    }
  } else {
    for (int i = lane; i < PW * PW; i += 32) {
      const int r = i / PW, cc = i - r * PW;
      const int yy = reflect101(ky - PR + r, L.h), xx = reflect101(kx - PR + cc, L.w);
      S.patch[r * PS + cc] = __ldg(img + (size_t)yy * L.pitch + xx);
    }
  }
  __syncwarp();

  // ---- IC_Angle (:77-104) ----
  int m01 = 0, m10 = 0;
  {
    const int u = lane - 15;
    if (lane < 31) {
      const int au = u < 0 ? -u : u;
#pragma unroll 1
      for (int v = -15; v <= 15; ++v) {
        const int av = v < 0 ? -v : v;
        if (au <= c_umax[av]) {
          const int val = S.patch[(PR + v) * PS + PR + u];
          m10 += u * val;
          m01 += v * val;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m10 += __shfl_xor_sync(0xffffffffu, m10, o);
      m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
  }
  const float angle = orb_fast_atan2_deg((float)m01, (float)m10);  // :103

  // ---- GaussianBlur 7x7 sigma 2, 8.8 fixed point kernel [18 34 48 56 48 34 18] ----
  for (int i = lane; i < PW * BW; i += 32) {
    const int r = i / BW, cc = i - r * BW;
    const uint8_t* p = S.patch + r * PS + cc;  // taps cc..cc+6  <->  x offsets (cc-18)-3 .. +3
    const int h = 18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3];
    S.hbuf[i] = (uint16_t)h;
  }
  __syncwarp();
  for (int i = lane; i < BW * BW; i += 32) {
    const int r = i / BW, cc = i - r * BW;
    const uint16_t* p = S.hbuf + r * BW + cc;  // rows r..r+6
    const uint32_t v = 18u * (p[0] + p[6 * BW]) + 34u * (p[BW] + p[5 * BW]) + 48u * (p[2 * BW] + p[4 * BW]) +
                       56u * p[3 * BW];
    S.blur[r * BS + cc] = (uint8_t)((v + 32768u) >> 16);
  }
  __syncwarp();

  // ---- computeOrbDescriptor (:108-147): lane -> descriptor byte ----
  float ca, sb;
  orb_sincos_deg(angle, &ca, &sb);  // a = cos, b = sin (:112-113, documented deviation 3)
  const uint8_t* center = S.blur + BRAD * BS + BRAD;
  int val = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int8_t* pt = c_pattern + (lane * 16 + 2 * j) * 2;
    const float x0 = (float)pt[0], y0 = (float)pt[1], x1 = (float)pt[2], y1 = (float)pt[3];
    const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sb), __fmul_rn(y0, ca)));
    const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sb)));
    const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sb), __fmul_rn(y1, ca)));
    const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sb)));
    const int t0 = center[r0 * BS + c0], t1 = center[r1 * BS + c1];
    val |= (t0 < t1) << j;
  }
  out_desc[((size_t)b * g.out_cap + oi) * 32 + lane] = (uint8_t)val;

  if (lane == 0) {
    orbfe_keypoint k;
    k.x = (float)kx;
    k.y = (float)ky;
    if (lvl != 0) { k.x = __fmul_rn(k.x, L.scale); k.y = __fmul_rn(k.y, L.scale); }  // :1117-1123
    k.size = L.kp_size;
    k.angle = angle;
    k.response = (float)cand_score(c);
    k.octave = lvl;
    k.class_id = -1;
    out_kp[(size_t)b * g.out_cap + oi] = k;
  }
}

void launch_describe(const PyrGeom& g, const ExtractorArena& a, int batch, cudaStream_t s) {
  dim3 grid((g.out_cap + DESC_WARPS - 1) / DESC_WARPS, batch);
  describe_kernel<<<grid, DESC_WARPS * 32, 0, s>>>(g, a.pyr, a.kept, a.kept_count, a.out_kp, a.out_desc, a.out_count,
                                                   a.flags);
}

}  // namespace orbfe
