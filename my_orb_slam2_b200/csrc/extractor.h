// extractor.h — the object behind an orbfe_extractor handle.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "geom.h"
#include "kernels.h"
#include "../../include/orbfe.h"

namespace orbfe {

struct Extractor {
  orbfe_extractor_params params{};
  orbfe_extractor_info info{};
  int width = 0, height = 0, max_batch = 0, device = 0;
  PyrGeom geom{};
  ExtractorArena arena{};
  size_t pyr_bytes = 0;
  cudaStream_t stream = nullptr;
  bool owns_stream = true;
  cudaEvent_t done = nullptr;   // recorded after the last queued work (cross-handle ordering for stereo)
  // recorded on THIS (left) handle's stream after the stereo kernels; the right handle's stream waits on it, so that
  // the right handle's next upload / run cannot overwrite the pyramid, keypoints and descriptors the matcher reads
  cudaEvent_t stereo_done = nullptr;
  int32_t* h_flags = nullptr;   // pinned
  int64_t launches = 0;
  int last_batch = 0;
  // orbfe_extract_run as one CUDA graph (captured lazily per batch size; ORBFE_NO_GRAPHS=1 disables)
  bool use_graph = true;
  cudaGraphExec_t run_graph = nullptr;
  int run_graph_batch = 0;
  // the captured schedule forks after the pyramid (blur beside FAST + quadtree); ORBFE_NO_FORK=1 keeps it linear
  bool fork_graph = true;
  cudaStream_t side = nullptr;
  cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
  cudaGraphExec_t stereo_graph = nullptr;  // orbfe_stereo_match_run with this handle as `left`
  int stereo_graph_batch = 0;
  const void* stereo_graph_right = nullptr;
  float stereo_graph_mbf = 0.f, stereo_graph_mb = 0.f;
  // host side of an orbfe_extract_download_async not yet finished (counts are checked by orbfe_extract_finish)
  const int32_t* pending_counts = nullptr;
  int pending_batch = 0, pending_capacity = 0;
  // the results of a handle live in ONE device block [keypoints | descriptors | counts | flags | uRight | depth |
  // stereo count] (the arena pointers point into it), so that a small batch can leave with a single D2H copy into
  // h_result (pinned; only allocated when the block is small) and be scattered to the caller's arrays on the host
  uint8_t* result_block = nullptr;
  uint8_t* h_result = nullptr;
  size_t off_desc = 0, off_count = 0, off_flags = 0, off_uright = 0, off_depth = 0, off_stcount = 0;
  size_t result_bytes = 0, result_bytes_extract = 0;  // whole block / the prefix without the stereo arrays
  // input side (k_input.cu): raw-frame staging and the rectification maps of orbfe_extractor_set_rectify_maps
  uint8_t* d_raw = nullptr;
  size_t raw_cap = 0;
  float* d_map_x = nullptr;
  float* d_map_y = nullptr;
  int src_w = 0, src_h = 0;
  // optional per-stage CUDA-event timing (bench.py roofline): ring of runs x (kStages+1) events
  static constexpr int kStages = 7;   // pyramid, fast, quadtree, blur, describe, stereo match, stereo median
  static constexpr int kProfRuns = 256;
  bool profiling = false;
  cudaEvent_t* prof_ev = nullptr;     // [kProfRuns][6] extractor stage events; [kProfRuns][3] stereo events after them
  int prof_runs = 0, prof_stereo_runs = 0;
};

// opt in to large dynamic shared memory etc.; defined in k_quadtree.cu
int configure_kernels(const PyrGeom& g);

}  // namespace orbfe
