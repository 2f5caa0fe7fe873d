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

// ---- level 0 from a device buffer: rows of `w` bytes at ANY byte alignment (1241-byte KITTI rows) into the pitched,
// 16-byte aligned rows of the arena.  One warp per row, a lane per 16 destination bytes: two 16-byte aligned loads
// (the row's alignment is the same for all its chunks, so the shift is warp-uniform), byte funnel shifts, one 16-byte
// store.  cudaMemcpy2DAsync(DeviceToDevice) does this job at ~2.6 TB/s (22.6 us for 64 KITTI images, 5 % of a step).
// A source chunk is only loaded if it holds at least one byte of the row, so nothing outside the 16-byte aligned
// hull of the caller's rows is touched; destination bytes past `w` in the last chunk are written as 0. ----
template <int Q>
__device__ __forceinline__ uint4 repitch_shift(const unsigned (&v)[8], unsigned r8) {
  return make_uint4(__funnelshift_r(v[Q], v[Q + 1], r8), __funnelshift_r(v[Q + 1], v[Q + 2], r8),
                    __funnelshift_r(v[Q + 2], v[Q + 3], r8), __funnelshift_r(v[Q + 3], v[Q + 4], r8));
}
constexpr int RP_WARPS = 8;
__global__ void __launch_bounds__(RP_WARPS * 32) repitch_kernel(const uint8_t* __restrict__ src, size_t image_stride,
                                                                size_t row_stride, uint8_t* __restrict__ dst, int pitch,
                                                                size_t dst_image, int w, int h, long long total_rows) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * RP_WARPS + (threadIdx.x >> 5);
  if (row >= total_rows) return;
  const long long b = row / h;
  const int y = (int)(row - b * h);
  const uint8_t* s = src + (size_t)b * image_stride + (size_t)y * row_stride;
  uint8_t* d = dst + (size_t)b * dst_image + (size_t)y * pitch;
  const unsigned sh = (unsigned)(reinterpret_cast<size_t>(s) & 15);  // warp-uniform
  const uint4* sa = reinterpret_cast<const uint4*>(s - sh);
  const unsigned q = sh >> 2, r8 = 8u * (sh & 3u);
  const int nchunks = (w + 15) >> 4;
  for (int c = lane; c < nchunks; c += 32) {
    // destination bytes 16c .. 16c+15 = bytes sh .. sh+15 of the aligned source chunks c, c+1
    const uint4 a = __ldg(sa + c);
    uint4 n = make_uint4(0u, 0u, 0u, 0u);
    if (sh != 0 && 16 * (c + 1) - (int)sh < w) n = __ldg(sa + c + 1);  // holds a byte of this row
    const unsigned v[8] = {a.x, a.y, a.z, a.w, n.x, n.y, n.z, n.w};
    uint4 o;
    if (q == 0) o = repitch_shift<0>(v, r8);
    else if (q == 1) o = repitch_shift<1>(v, r8);
    else if (q == 2) o = repitch_shift<2>(v, r8);
    else o = repitch_shift<3>(v, r8);
    const int left = w - 16 * c;  // valid bytes of this chunk (>= 1)
    if (left < 16) {
      unsigned* ow = reinterpret_cast<unsigned*>(&o);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int nb = left - 4 * k;  // valid bytes of word k
        ow[k] = nb >= 4 ? ow[k] : (nb <= 0 ? 0u : (ow[k] & (0xFFFFFFFFu >> (8 * (4 - nb)))));
      }
    }
    *reinterpret_cast<uint4*>(d + 16 * c) = o;
  }
}

void launch_repitch(const uint8_t* src, size_t image_stride, size_t row_stride, uint8_t* dst, int pitch, size_t dst_image,
                    int w, int h, int batch, cudaStream_t s) {
  const long long rows = (long long)h * batch;
  repitch_kernel<<<(unsigned)((rows + RP_WARPS - 1) / RP_WARPS), RP_WARPS * 32, 0, s>>>(src, image_stride, row_stride, dst,
                                                                                       pitch, dst_image, w, h, rows);
}

}  // namespace orbfe
