// stereo.cu — host side of orbfe_stereo_match*: Frame::ComputeStereoMatches
// (reference: src/Frame.cc:496-686) on the device-resident results of two extractor
// handles.  The kernels are in k_stereo.cu.
#include "common.cuh"
#include "extractor.h"
#include "kernels.h"

using namespace orbfe;

extern "C" {

int orbfe_stereo_match_run(orbfe_extractor* hl, orbfe_extractor* hr, int batch, float mbf, float mb) {
  ORBFE_ARG(hl && hr);
  Extractor* l = reinterpret_cast<Extractor*>(hl);
  Extractor* r = reinterpret_cast<Extractor*>(hr);
  ORBFE_ARG(l->device == r->device && l->width == r->width && l->height == r->height);
  ORBFE_ARG(l->geom.nlevels == r->geom.nlevels && l->geom.out_cap == r->geom.out_cap);
  ORBFE_ARG(l->params.scale_factor == r->params.scale_factor);
  ORBFE_ARG(batch >= 1 && batch <= l->max_batch && batch <= r->max_batch);
  if (l->last_batch < batch || r->last_batch < batch) {
    set_error("orbfe_stereo_match: run the extractors on %d images first", batch);
    return ORBFE_ERR_STATE;
  }
  ORBFE_ARG(mb > 0.f);
  DeviceGuard guard(l->device);
  if (l != r && l->stream != r->stream) {  // order left's stream after right's queued work (the join at src/Frame.cc:91-92)
    ORBFE_CUDA(cudaEventRecord(r->done, r->stream));
    ORBFE_CUDA(cudaStreamWaitEvent(l->stream, r->done, 0));
  }
  cudaEvent_t* ev = nullptr;
  if (l->profiling && l->prof_ev && l->prof_stereo_runs < Extractor::kProfRuns)
    ev = l->prof_ev + 6 * Extractor::kProfRuns + 3 * l->prof_stereo_runs++;
  if (ev || !l->use_graph) {
    launch_stereo(l->geom, l->arena, r->arena, batch, mbf, mb, l->stream, ev);
  } else {
    // the three stereo kernels as one graph launch (keyed by everything baked into the kernel parameters)
    if (!l->stereo_graph || l->stereo_graph_batch != batch || l->stereo_graph_right != r || l->stereo_graph_mbf != mbf ||
        l->stereo_graph_mb != mb) {
      if (l->stereo_graph) { cudaGraphExecDestroy(l->stereo_graph); l->stereo_graph = nullptr; }
      cudaGraph_t graph = nullptr;
      ORBFE_CUDA(cudaStreamBeginCapture(l->stream, cudaStreamCaptureModeThreadLocal));
      launch_stereo(l->geom, l->arena, r->arena, batch, mbf, mb, l->stream, nullptr);
      ORBFE_CUDA(cudaStreamEndCapture(l->stream, &graph));
      ORBFE_CUDA(cudaGraphInstantiate(&l->stereo_graph, graph, 0));
      cudaGraphDestroy(graph);
      l->stereo_graph_batch = batch;
      l->stereo_graph_right = r;
      l->stereo_graph_mbf = mbf;
      l->stereo_graph_mb = mb;
    }
    ORBFE_CUDA(cudaGraphLaunch(l->stereo_graph, l->stream));
  }
  l->launches += kLaunchesPerStereo;
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

int orbfe_stereo_match_download_async(orbfe_extractor* hl, int batch, float* u_right, float* depth, int capacity,
                                      int32_t* n_matched) {
  ORBFE_ARG(hl && capacity >= 0);
  Extractor* l = reinterpret_cast<Extractor*>(hl);
  ORBFE_ARG(batch >= 1 && batch <= l->max_batch);
  DeviceGuard guard(l->device);
  const int cap = l->geom.out_cap;
  const int ncopy = capacity < cap ? capacity : cap;
  if (u_right && ncopy > 0)
    ORBFE_CUDA(cudaMemcpy2DAsync(u_right, (size_t)capacity * 4, l->arena.st_uright, (size_t)cap * 4, (size_t)ncopy * 4,
                                 batch, cudaMemcpyDeviceToHost, l->stream));
  if (depth && ncopy > 0)
    ORBFE_CUDA(cudaMemcpy2DAsync(depth, (size_t)capacity * 4, l->arena.st_depth, (size_t)cap * 4, (size_t)ncopy * 4,
                                 batch, cudaMemcpyDeviceToHost, l->stream));
  if (n_matched)
    ORBFE_CUDA(cudaMemcpyAsync(n_matched, l->arena.st_count, (size_t)batch * 4, cudaMemcpyDeviceToHost, l->stream));
  return ORBFE_OK;
}

int orbfe_stereo_match_download(orbfe_extractor* hl, int batch, float* u_right, float* depth, int capacity,
                                int32_t* n_matched) {
  const int rc = orbfe_stereo_match_download_async(hl, batch, u_right, depth, capacity, n_matched);
  if (rc != ORBFE_OK) return rc;
  Extractor* l = reinterpret_cast<Extractor*>(hl);
  DeviceGuard guard(l->device);
  ORBFE_CUDA(cudaStreamSynchronize(l->stream));
  return ORBFE_OK;
}

int orbfe_stereo_frame(orbfe_extractor* hl, orbfe_extractor* hr, const uint8_t* left_images,
                       const uint8_t* right_images, size_t image_stride, int row_stride, int batch, float mbf, float mb,
                       orbfe_keypoint* kp_left, uint8_t* desc_left, int32_t* n_left, orbfe_keypoint* kp_right,
                       uint8_t* desc_right, int32_t* n_right, float* u_right, float* depth, int32_t* n_matched,
                       int capacity) {
  ORBFE_ARG(hl && hr && hl != hr && left_images && right_images && n_left && n_right);
  int rc = orbfe_extract_upload(hl, left_images, image_stride, row_stride, batch, 0);
  if (rc == ORBFE_OK) rc = orbfe_extract_upload(hr, right_images, image_stride, row_stride, batch, 0);
  if (rc == ORBFE_OK) rc = orbfe_extract_run(hl, batch);
  if (rc == ORBFE_OK) rc = orbfe_extract_run(hr, batch);
  if (rc == ORBFE_OK) rc = orbfe_stereo_match_run(hl, hr, batch, mbf, mb);
  // right results leave while the matcher runs on left's stream; left results and the matches follow it
  if (rc == ORBFE_OK) rc = orbfe_extract_download_async(hr, batch, kp_right, desc_right, capacity, n_right);
  if (rc == ORBFE_OK) rc = orbfe_extract_download_async(hl, batch, kp_left, desc_left, capacity, n_left);
  if (rc == ORBFE_OK) rc = orbfe_stereo_match_download_async(hl, batch, u_right, depth, capacity, n_matched);
  const int rr = orbfe_extract_finish(hr), rl = orbfe_extract_finish(hl);  // always drain both streams
  if (rc != ORBFE_OK) return rc;
  return rr != ORBFE_OK ? rr : rl;
}

int orbfe_stereo_match(orbfe_extractor* hl, orbfe_extractor* hr, int batch, float mbf, float mb, float* u_right,
                       float* depth, int capacity, int32_t* n_matched) {
  int rc = orbfe_stereo_match_run(hl, hr, batch, mbf, mb);
  if (rc != ORBFE_OK) return rc;
  return orbfe_stereo_match_download(hl, batch, u_right, depth, capacity, n_matched);
}

}  // extern "C"
