// extractor.h — the object behind an orbfe_extractor handle.
#pragma once
#include <cuda_runtime.h>

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
  cudaEvent_t done = nullptr;   // recorded after the last queued work (cross-handle ordering for stereo)
  int32_t* h_flags = nullptr;   // pinned
  int64_t launches = 0;
  int last_batch = 0;
};

// opt in to large dynamic shared memory etc.; defined in k_quadtree.cu
int configure_kernels(const PyrGeom& g);

}  // namespace orbfe
