// k_pyramid.cu — ComputePyramid (reference: src/ORBextractor.cc:1129-1154).
//
// Level l is cv::resize(level l-1, INTER_LINEAR) in OpenCV's 11-bit fixed point
// (SURVEY.md A.1); the 19 px copyMakeBorder frame is not materialised (A.5).
// HBM-bound integer stencil: each thread produces 4 horizontally adjacent
// destination pixels (one uchar4 store); the two source rows it needs are read
// through the read-only path.  Coefficients come from per-level tables built on
// the host exactly as OpenCV builds them.
#include "kernels.h"

namespace orbfe {

__global__ void __launch_bounds__(256) pyramid_level_kernel(const __grid_constant__ PyrGeom g, uint8_t* __restrict__ pyr,
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

void launch_pyramid_level(const PyrGeom& g, const ExtractorArena& a, int lvl, int batch, cudaStream_t s) {
  const LevelGeom& D = g.lv[lvl];
  dim3 block(32, 8);
  dim3 grid((D.w + 127) / 128, (D.h + 7) / 8, batch);
  pyramid_level_kernel<<<grid, block, 0, s>>>(g, a.pyr, a.xtab, a.ytab, lvl);
}

}  // namespace orbfe
