// k_input.cu — the input side of the pipeline on the device, so raw camera frames can be handed to
// the extractor: cv::cvtColor(..., CV_{RGB,BGR}[A]2GRAY) as Tracking::GrabImage* applies it (src/Tracking.cc:189-214)
// and cv::remap(..., INTER_LINEAR) with the CV_32F maps of initUndistortRectifyMap as the EuRoC stereo example applies
// it (Examples/Stereo/stereo_euroc.cc:97-98, :136-137).  Both write level 0 of the pyramid arena directly; integer
// arithmetic identical to OpenCV 4.13 (tests/golden/input_side.npz).
#include "kernels.h"

namespace orbfe {

// gray = (R*9798 + G*19235 + B*3735 + 2^14) >> 15; one thread per 4 output pixels (one word store)
template <int CH>
__global__ void __launch_bounds__(256) gray_kernel(const uint8_t* __restrict__ src, size_t image_stride, int row_stride,
                                                   int rgb, uint8_t* __restrict__ dst, int pitch, size_t dst_image, int w,
                                                   int h) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int y = blockIdx.y, b = blockIdx.z;
  if (x4 >= w) return;
  const uint8_t* s = src + (size_t)b * image_stride + (size_t)y * row_stride + (size_t)x4 * CH;
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (x4 + i < w) {
      const int c0 = s[i * CH], c1 = s[i * CH + 1], c2 = s[i * CH + 2];
      const int r = rgb ? c0 : c2, bl = rgb ? c2 : c0;
      out |= (uint32_t)((r * 9798 + c1 * 19235 + bl * 3735 + (1 << 14)) >> 15) << (8 * i);
    }
  }
  uint8_t* d = dst + (size_t)b * dst_image + (size_t)y * pitch + x4;
  if (x4 + 3 < w) {
    *reinterpret_cast<uint32_t*>(d) = out;  // pitch is a multiple of 128
  } else {
    for (int i = 0; x4 + i < w; ++i) d[i] = (uint8_t)(out >> (8 * i));
  }
}

// one thread per output pixel: 1/32-pixel coordinates, 15-bit bilinear weights, constant border 0
__global__ void __launch_bounds__(256) remap_kernel(const uint8_t* __restrict__ src, size_t image_stride, int row_stride,
                                                    int sw, int sh, const float* __restrict__ mapx,
                                                    const float* __restrict__ mapy, uint8_t* __restrict__ dst, int pitch,
                                                    size_t dst_image, int w, int h) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x;
  const int y = blockIdx.y, b = blockIdx.z;
  if (x >= w) return;
  const size_t mi = (size_t)y * w + x;
  const int sx = __float2int_rn(__fmul_rn(__ldg(mapx + mi), 32.0f));  // cvRound(map * INTER_TAB_SIZE)
  const int sy = __float2int_rn(__fmul_rn(__ldg(mapy + mi), 32.0f));
  const int ix = max(-32768, min(32767, sx >> 5)), iy = max(-32768, min(32767, sy >> 5));  // saturate_cast<short>
  const int ax = sx & 31, ay = sy & 31;
  const uint8_t* s = src + (size_t)b * image_stride;
  const bool x0 = ix >= 0 && ix < sw, x1 = ix + 1 >= 0 && ix + 1 < sw;
  const bool y0 = iy >= 0 && iy < sh, y1 = iy + 1 >= 0 && iy + 1 < sh;
  const int p00 = (x0 && y0) ? s[(size_t)iy * row_stride + ix] : 0;
  const int p01 = (x1 && y0) ? s[(size_t)iy * row_stride + ix + 1] : 0;
  const int p10 = (x0 && y1) ? s[(size_t)(iy + 1) * row_stride + ix] : 0;
  const int p11 = (x1 && y1) ? s[(size_t)(iy + 1) * row_stride + ix + 1] : 0;
  const int v = p00 * ((32 - ay) * (32 - ax) * 32) + p01 * ((32 - ay) * ax * 32) + p10 * (ay * (32 - ax) * 32) +
                p11 * (ay * ax * 32);
  dst[(size_t)b * dst_image + (size_t)y * pitch + x] = (uint8_t)((v + (1 << 14)) >> 15);
}

void launch_gray(const uint8_t* src, size_t image_stride, int row_stride, int channels, int rgb, uint8_t* dst, int pitch,
                 size_t dst_image, int w, int h, int batch, cudaStream_t s) {
  const dim3 grid(((w + 3) / 4 + 255) / 256, h, batch);
  if (channels == 3)
    gray_kernel<3><<<grid, 256, 0, s>>>(src, image_stride, row_stride, rgb, dst, pitch, dst_image, w, h);
  else
    gray_kernel<4><<<grid, 256, 0, s>>>(src, image_stride, row_stride, rgb, dst, pitch, dst_image, w, h);
}

void launch_remap(const uint8_t* src, size_t image_stride, int row_stride, int sw, int sh, const float* mapx,
                  const float* mapy, uint8_t* dst, int pitch, size_t dst_image, int w, int h, int batch,
                  cudaStream_t s) {
  const dim3 grid((w + 255) / 256, h, batch);
  remap_kernel<<<grid, 256, 0, s>>>(src, image_stride, row_stride, sw, sh, mapx, mapy, dst, pitch, dst_image, w, h);
}

}  // namespace orbfe
