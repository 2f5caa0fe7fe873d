// stereo.cu — host side of orbfe_stereo_match*: Frame::ComputeStereoMatches
// (reference: src/Frame.cc:496-686) on the device-resident results of two extractor
// handles.  The kernels are in k_stereo.cu.
#include <cstring>

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
  // the three stereo kernels as one graph launch (keyed by everything baked into the kernel parameters)
  if (!ev && l->use_graph &&
      (!l->stereo_graph || l->stereo_graph_batch != batch || l->stereo_graph_right != r || l->stereo_graph_mbf != mbf ||
       l->stereo_graph_mb != mb)) {
    if (l->stereo_graph) { cudaGraphExecDestroy(l->stereo_graph); l->stereo_graph = nullptr; }
    const bool ok = capture_graph(l->stream, &l->stereo_graph, [&] {
      launch_stereo(l->geom, l->arena, r->arena, batch, mbf, mb, l->stream, nullptr);
    });
    if (ok) {
      l->stereo_graph_batch = batch;
      l->stereo_graph_right = r;
      l->stereo_graph_mbf = mbf;
      l->stereo_graph_mb = mb;
    } else {
      l->use_graph = false;  // plain launches from now on
    }
  }
  if (ev || !l->use_graph) {
    launch_stereo(l->geom, l->arena, r->arena, batch, mbf, mb, l->stream, ev);
  } else {
    ORBFE_CUDA(cudaGraphLaunch(l->stereo_graph, l->stream));
  }
  if (l != r && l->stream != r->stream) {
    // write-after-read: the matcher reads the right handle's pyramid, keypoints and descriptors on left's stream;
    // the right handle's NEXT upload / run (on its own stream) must not start before these reads are done
    ORBFE_CUDA(cudaEventRecord(l->stereo_done, l->stream));
    ORBFE_CUDA(cudaStreamWaitEvent(r->stream, l->stereo_done, 0));
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
  Extractor* l = reinterpret_cast<Extractor*>(hl);
  Extractor* r = reinterpret_cast<Extractor*>(hr);
  if (rc == ORBFE_OK && batch == l->max_batch && batch == r->max_batch && l->h_result && r->h_result) {
    // single frames / small batches: each handle's results are one device block (extractor.h), so they leave with
    // ONE copy per handle (11 separate copies otherwise, ~4 us of latency each) and are scattered on the host
    DeviceGuard guard(l->device);
    ORBFE_CUDA(cudaMemcpyAsync(r->h_result, r->result_block, r->result_bytes_extract, cudaMemcpyDeviceToHost, r->stream));
    ORBFE_CUDA(cudaMemcpyAsync(l->h_result, l->result_block, l->result_bytes, cudaMemcpyDeviceToHost, l->stream));
    ORBFE_CUDA(cudaStreamSynchronize(r->stream));
    ORBFE_CUDA(cudaStreamSynchronize(l->stream));
    const int cap = l->geom.out_cap;
    auto scatter = [&](const Extractor* e, orbfe_keypoint* kp, uint8_t* desc, int32_t* n_out) -> int {
      const int32_t* cnt = reinterpret_cast<const int32_t*>(e->h_result + e->off_count);
      const int32_t* flg = reinterpret_cast<const int32_t*>(e->h_result + e->off_flags);
      for (int b = 0; b < batch; ++b) {
        if (flg[b]) { set_error("image %d: internal slab overflow (flags=%d)", b, flg[b]); return ORBFE_ERR_CAPACITY; }
        if (cnt[b] > capacity) { set_error("image %d: %d keypoints > capacity %d", b, cnt[b], capacity); return ORBFE_ERR_CAPACITY; }
        n_out[b] = cnt[b];
        if (kp) memcpy(kp + (size_t)b * capacity, e->h_result + (size_t)b * cap * sizeof(orbfe_keypoint),
                       (size_t)cnt[b] * sizeof(orbfe_keypoint));
        if (desc) memcpy(desc + (size_t)b * capacity * 32, e->h_result + e->off_desc + (size_t)b * cap * 32, (size_t)cnt[b] * 32);
      }
      return ORBFE_OK;
    };
    rc = scatter(r, kp_right, desc_right, n_right);
    if (rc == ORBFE_OK) rc = scatter(l, kp_left, desc_left, n_left);
    if (rc != ORBFE_OK) return rc;
    for (int b = 0; b < batch; ++b) {
      const size_t n = (size_t)n_left[b];
      if (u_right) memcpy(u_right + (size_t)b * capacity, l->h_result + l->off_uright + (size_t)b * cap * 4, n * 4);
      if (depth) memcpy(depth + (size_t)b * capacity, l->h_result + l->off_depth + (size_t)b * cap * 4, n * 4);
      if (n_matched) n_matched[b] = reinterpret_cast<const int32_t*>(l->h_result + l->off_stcount)[b];
    }
    return ORBFE_OK;
  }
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
