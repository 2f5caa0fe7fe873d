// frame.cu — what the mono and RGB-D Frame constructors do after ExtractORB (reference: src/Frame.cc:118-228):
//   * Frame::UndistortKeyPoints (:429-460): cv::undistortPoints on every keypoint (five fixed-point iterations in
//     double, include/orb_math_shared.h: orb_undistort_point, bit-equal to OpenCV 4.13), all other fields copied;
//   * Frame::ComputeImageBounds (:462-494): the same function on the four image corners (host: four points);
//   * Frame::ComputeStereoFromRGBD (:688-712): a depth lookup per keypoint at its DISTORTED position and
//     mvuRight = undistorted x - mbf / depth.
// (AssignFeaturesToGrid, the remaining step, is build_grid_kernel in k_search.cu: it runs where the grid is consumed.)
// One thread per keypoint; the work is a few hundred double operations per keypoint, so these are latency-sized kernels.
#include <algorithm>
#include <cstring>

#include "../../include/orb_math_shared.h"
#include "common.cuh"
#include "extractor.h"
#include "kernels.h"

namespace orbfe {

__global__ void __launch_bounds__(128) undistort_kernel(const __grid_constant__ PyrGeom g,
                                                        const orbfe_keypoint* __restrict__ kp,
                                                        const int32_t* __restrict__ count, const orbfe_camera cam,
                                                        orbfe_keypoint* __restrict__ kp_un) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count[b]) return;
  orbfe_keypoint k = kp[(size_t)b * g.out_cap + i];
  if (cam.k[0] != 0.f) {  // mDistCoef.at<float>(0) == 0.0 -> mvKeysUn = mvKeys (:431-435)
    float xu, yu;
    orb_undistort_point(k.x, k.y, cam.fx, cam.fy, cam.cx, cam.cy, cam.k, &xu, &yu);
    k.x = xu;
    k.y = yu;
  }
  kp_un[(size_t)b * g.out_cap + i] = k;
}

__global__ void __launch_bounds__(128) rgbd_kernel(const __grid_constant__ PyrGeom g, const orbfe_keypoint* __restrict__ kp,
                                                   const orbfe_keypoint* __restrict__ kp_un,
                                                   const int32_t* __restrict__ count, const float* __restrict__ depth_img,
                                                   int w, int h, float mbf, float* __restrict__ u_right,
                                                   float* __restrict__ depth) {
  const int b = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= g.out_cap) return;
  const size_t o = (size_t)b * g.out_cap + i;
  float ur = -1.f, dz = -1.f;  // mvuRight = mvDepth = vector<float>(N, -1) (:690-691)
  if (i < count[b]) {
    const int u = (int)kp[o].x, v = (int)kp[o].y;  // imDepth.at<float>(v, u) with float arguments: truncation (:701)
    if (u >= 0 && u < w && v >= 0 && v < h) {      // memory guard (keypoints lie >= 19 px inside the image)
      const float d = depth_img[((size_t)b * h + v) * w + u];
      if (d > 0.f) {
        dz = d;
        ur = __fsub_rn(kp_un[o].x, __fdiv_rn(mbf, d));  // kpU.pt.x - mbf / d (:706)
      }
    }
  }
  u_right[o] = ur;
  depth[o] = dz;
}

void launch_undistort(const PyrGeom& g, const ExtractorArena& a, int batch, const orbfe_camera& cam, cudaStream_t s) {
  dim3 grid((g.out_cap + 127) / 128, batch);
  undistort_kernel<<<grid, 128, 0, s>>>(g, a.out_kp, a.out_count, cam, a.kp_un);
}

void launch_stereo_from_rgbd(const PyrGeom& g, const ExtractorArena& a, int batch, int w, int h, float mbf, cudaStream_t s) {
  dim3 grid((g.out_cap + 127) / 128, batch);
  rgbd_kernel<<<grid, 128, 0, s>>>(g, a.out_kp, a.kp_un, a.out_count, a.depth_img, w, h, mbf, a.st_uright, a.st_depth);
}

}  // namespace orbfe

using namespace orbfe;

extern "C" {

int orbfe_undistort_keypoints(orbfe_extractor* h, int batch, const orbfe_camera* cam, orbfe_keypoint* keys_un,
                              int capacity) {
  ORBFE_ARG(h && cam && capacity >= 0);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(batch >= 1 && batch <= e->max_batch);
  if (e->last_batch < batch) {
    set_error("orbfe_undistort_keypoints: run the extractor on %d images first", batch);
    return ORBFE_ERR_STATE;
  }
  DeviceGuard guard(e->device);
  const PyrGeom& g = e->geom;
  ExtractorArena& a = e->arena;
  if (!a.kp_un)
    ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&a.kp_un), (size_t)e->max_batch * g.out_cap * sizeof(orbfe_keypoint)));
  launch_undistort(g, a, batch, *cam, e->stream);
  a.kp_un_valid = true;
  e->launches += 1;
  if (keys_un) {
    const int ncopy = std::min(capacity, g.out_cap);
    if (ncopy > 0)
      ORBFE_CUDA(cudaMemcpy2DAsync(keys_un, (size_t)capacity * sizeof(orbfe_keypoint), a.kp_un,
                                   (size_t)g.out_cap * sizeof(orbfe_keypoint), (size_t)ncopy * sizeof(orbfe_keypoint),
                                   batch, cudaMemcpyDeviceToHost, e->stream));
  }
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_image_bounds(const orbfe_camera* cam, int width, int height, float* bounds) {
  ORBFE_ARG(cam && bounds && width > 0 && height > 0);
  if (cam->k[0] == 0.f) {  // :487-493
    bounds[0] = 0.f; bounds[1] = (float)width; bounds[2] = 0.f; bounds[3] = (float)height;
    return ORBFE_OK;
  }
  // corners (0,0), (cols,0), (0,rows), (cols,rows) (:466-470); this translation unit is compiled with
  // -ffp-contract=off, so the host evaluation of orb_undistort_point gives the device's (and OpenCV's) bits
  const float cx[4] = {0.f, (float)width, 0.f, (float)width}, cy[4] = {0.f, 0.f, (float)height, (float)height};
  float x[4], y[4];
  for (int i = 0; i < 4; ++i) orb_undistort_point(cx[i], cy[i], cam->fx, cam->fy, cam->cx, cam->cy, cam->k, &x[i], &y[i]);
  bounds[0] = std::min(x[0], x[2]);  // :477-480
  bounds[1] = std::max(x[1], x[3]);
  bounds[2] = std::min(y[0], y[1]);
  bounds[3] = std::max(y[2], y[3]);
  return ORBFE_OK;
}

int orbfe_stereo_from_rgbd(orbfe_extractor* h, int batch, const float* depth, size_t image_stride, int row_stride,
                           float mbf, float* u_right, float* depth_out, int capacity) {
  ORBFE_ARG(h && depth && capacity >= 0);
  Extractor* e = reinterpret_cast<Extractor*>(h);
  ORBFE_ARG(batch >= 1 && batch <= e->max_batch && row_stride >= e->width);
  if (e->last_batch < batch || !e->arena.kp_un_valid) {
    set_error("orbfe_stereo_from_rgbd: run the extractor and orbfe_undistort_keypoints on %d images first", batch);
    return ORBFE_ERR_STATE;
  }
  DeviceGuard guard(e->device);
  const PyrGeom& g = e->geom;
  ExtractorArena& a = e->arena;
  const int w = e->width, hh = e->height;
  if (!a.depth_img)
    ORBFE_CUDA(cudaMalloc(reinterpret_cast<void**>(&a.depth_img), (size_t)e->max_batch * w * hh * sizeof(float)));
  for (int b = 0; b < batch; ++b)
    ORBFE_CUDA(cudaMemcpy2DAsync(a.depth_img + (size_t)b * w * hh, (size_t)w * 4, depth + (size_t)b * image_stride,
                                 (size_t)row_stride * 4, (size_t)w * 4, hh, cudaMemcpyHostToDevice, e->stream));
  launch_stereo_from_rgbd(g, a, batch, w, hh, mbf, e->stream);
  e->launches += 1;
  const int ncopy = std::min(capacity, g.out_cap);
  if (u_right && ncopy > 0)
    ORBFE_CUDA(cudaMemcpy2DAsync(u_right, (size_t)capacity * 4, a.st_uright, (size_t)g.out_cap * 4, (size_t)ncopy * 4, batch,
                                 cudaMemcpyDeviceToHost, e->stream));
  if (depth_out && ncopy > 0)
    ORBFE_CUDA(cudaMemcpy2DAsync(depth_out, (size_t)capacity * 4, a.st_depth, (size_t)g.out_cap * 4, (size_t)ncopy * 4, batch,
                                 cudaMemcpyDeviceToHost, e->stream));
  ORBFE_CUDA(cudaStreamSynchronize(e->stream));
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

}  // extern "C"
