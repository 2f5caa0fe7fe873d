// common.cuh — error plumbing shared by all translation units of liborbfe.so.
#pragma once
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdint>
#include <cstdio>

#include "../../include/orbfe.h"

namespace orbfe {

// thread-local text behind orbfe_last_error()
void set_error(const char* fmt, ...);

#define ORBFE_CUDA(call)                                                                              \
  do {                                                                                                \
    cudaError_t err__ = (call);                                                                       \
    if (err__ != cudaSuccess) {                                                                       \
      orbfe::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(err__));      \
      return ORBFE_ERR_CUDA;                                                                          \
    }                                                                                                 \
  } while (0)

#define ORBFE_ARG(cond)                                                       \
  do {                                                                        \
    if (!(cond)) {                                                            \
      orbfe::set_error("%s:%d: bad argument: %s", __FILE__, __LINE__, #cond); \
      return ORBFE_ERR_ARG;                                                   \
    }                                                                         \
  } while (0)

static inline int div_up(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

// RAII device switch (handles live on a fixed device; callers may be on another)
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; }
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace orbfe
