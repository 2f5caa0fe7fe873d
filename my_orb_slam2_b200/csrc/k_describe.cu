// k_describe.cu — everything after DistributeOctTree for one kept keypoint, one warp per keypoint:
//   * the fix-up of src/ORBextractor.cc:859-870 (pt += minBorder, octave, size);
//   * computeOrientation -> IC_Angle (:472-479, :77-104): m10/m01 over the radius-15 disc of the UN-blurred level
//     (lane u = column of the disc, loop over its 31 rows), warp-shuffle reduction, cv::fastAtan2 polynomial;
//   * computeOrbDescriptor (:108-147) on the blurred level (k_blur.cu): lane i builds descriptor byte i
//     (8 tests = 16 gathers), its 32 pattern bytes read with two coalesced 16-byte loads;
//   * the rescale by mvScaleFactor[level] and the level-order concatenation of operator() (:1117-1125).
// Kept keypoints are >= 19 px from every level edge (FAST margin 3 + minBorder 16), the disc has radius 15 and the
// rotated pattern reaches <= 18 px, so no access leaves the level image.
// Float steps that decide bits (angle, sin/cos, the rotated sample coordinates) use the explicitly rounded
// operations of include/orb_math_shared.h so that no FMA is formed.
#include "../../include/orb_math_shared.h"
#include "kernels.h"

namespace orbfe {

constexpr int DESC_WARPS = 8;

// 256 test pairs x (x0,y0,x1,y1) as int8.  In GLOBAL memory, not __constant__: lane i reads ITS 32 bytes (two
// coalesced 16-byte loads); a lane-indexed read of constant memory would be serialised 32 ways.
__device__ __align__(16) int8_t d_pattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};
// end of each row of the radius-15 disc (umax, :454-469)
__constant__ int c_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};

__global__ void __launch_bounds__(DESC_WARPS * 32) describe_kernel(const __grid_constant__ PyrGeom g,
                                                                   const uint8_t* __restrict__ pyr,
                                                                   const uint8_t* __restrict__ blur,
                                                                   const uint32_t* __restrict__ kept,
                                                                   const int32_t* __restrict__ kept_count,
                                                                   orbfe_keypoint* __restrict__ out_kp,
                                                                   uint8_t* __restrict__ out_desc,
                                                                   int32_t* __restrict__ out_count,
                                                                   int32_t* __restrict__ flags) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.y;
  const int oi = blockIdx.x * DESC_WARPS + warp;  // output row of this keypoint
  // level of this output row: levels are concatenated in order 0..L-1 (:1096-1126)
  const int32_t* kc = kept_count + b * kMaxLevels;
  int lvl = 0, first = 0;
  {
    int acc = 0;
    bool found = false;
    for (int l = 0; l < g.nlevels; ++l) {
      const int c = kc[l];
      if (!found && oi < acc + c) { lvl = l; first = acc; found = true; }
      acc += c;
    }
    if (oi == 0 && lane == 0) {
      out_count[b] = min(acc, g.out_cap);
      if (acc > g.out_cap) atomicOr(&flags[b], FLAG_OUT_OVERFLOW);
    }
    if (!found || oi >= g.out_cap) return;
  }
  const LevelGeom& L = g.lv[lvl];
  const uint32_t c = kept[(size_t)b * g.kept_total + L.kept_off + (oi - first)];
  const int kx = cand_x(c) + kMinBorder, ky = cand_y(c) + kMinBorder;  // :865-866
  const size_t lvl_off = L.lvl_base + (size_t)b * L.lvl_bytes + (size_t)ky * L.pitch + kx;

  // ---- IC_Angle (:77-104) ----
  int m01 = 0, m10 = 0;
  {
    const uint8_t* center = pyr + lvl_off;
    const int u = lane - 15;
    const int au = u < 0 ? -u : u;
    if (lane < 31) {
#pragma unroll
      for (int v = -15; v <= 15; ++v) {
        const int av = v < 0 ? -v : v;
        if (au <= c_umax[av]) {
          const int val = __ldg(center + v * L.pitch + u);
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

  // ---- computeOrbDescriptor (:108-147): lane -> descriptor byte ----
  float ca, sb;
  orb_sincos_deg(angle, &ca, &sb);  // a = cos, b = sin (:112-113, documented deviation 3)
  const uint8_t* center = blur + lvl_off;
  const int step = L.pitch;
  int val = 0;
  const uint4 p_lo = __ldg(reinterpret_cast<const uint4*>(d_pattern) + 2 * lane);
  const uint4 p_hi = __ldg(reinterpret_cast<const uint4*>(d_pattern) + 2 * lane + 1);
  const unsigned pw[8] = {p_lo.x, p_lo.y, p_lo.z, p_lo.w, p_hi.x, p_hi.y, p_hi.z, p_hi.w};
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const unsigned q = pw[j];  // bytes: x0, y0, x1, y1 of test 8*lane + j
    const float x0 = (float)(int8_t)(q & 0xFF), y0 = (float)(int8_t)((q >> 8) & 0xFF);
    const float x1 = (float)(int8_t)((q >> 16) & 0xFF), y1 = (float)(int8_t)(q >> 24);
    const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sb), __fmul_rn(y0, ca)));
    const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, ca), __fmul_rn(y0, sb)));
    const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sb), __fmul_rn(y1, ca)));
    const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, ca), __fmul_rn(y1, sb)));
    const int t0 = __ldg(center + r0 * step + c0), t1 = __ldg(center + r1 * step + c1);
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
  describe_kernel<<<grid, DESC_WARPS * 32, 0, s>>>(g, a.pyr, a.blur, a.kept, a.kept_count, a.out_kp, a.out_desc,
                                                   a.out_count, a.flags);
}

}  // namespace orbfe
