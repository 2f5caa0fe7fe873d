// common.cuh — error plumbing shared by all translation units of liborbfe.so.
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <thread>

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

// A handle (stream + workspace) serves ONE call at a time: the reference's matcher objects are stack-local, one per
// calling thread.  Two host threads inside the same handle would overwrite each other's staged
// inputs; instead of wrong matches the second caller gets ORBFE_ERR_STATE.  Entry points that call other entry points
// of the same handle (same thread) pass through.
struct BusyState {
  std::atomic<int> busy{0};
  std::atomic<size_t> owner{0};
};
struct BusyGuard {
  BusyState* st;
  bool ok = false, outer = false;
  explicit BusyGuard(BusyState* s) : st(s) {
    const size_t me = std::hash<std::thread::id>()(std::this_thread::get_id()) | 1u;
    if (st->busy.load(std::memory_order_acquire) && st->owner.load(std::memory_order_relaxed) == me) { ok = true; return; }
    int expected = 0;
    ok = outer = st->busy.compare_exchange_strong(expected, 1, std::memory_order_acq_rel);
    if (outer) st->owner.store(me, std::memory_order_relaxed);
  }
  ~BusyGuard() {
    if (outer) {
      st->owner.store(0, std::memory_order_relaxed);
      st->busy.store(0, std::memory_order_release);
    }
  }
};
#define ORBFE_MATCHER_ENTER(m)                                                                                      \
  orbfe::BusyGuard busy_guard_(&(m)->busy);                                                                         \
  if (!busy_guard_.ok) {                                                                                            \
    orbfe::set_error("%s: this matcher handle is inside another call (one orbfe_matcher per host thread)", __func__); \
    return ORBFE_ERR_STATE;                                                                                         \
  }                                                                                                                 \
  orbfe::DeviceGuard guard((m)->device)

// Capture `issue()` (kernel launches on `stream`) into an executable graph.  Whatever fails between BeginCapture and
// Instantiate, the capture is always ended and the intermediate graph destroyed, so the stream never stays in capture
// mode; on failure *exec stays null and the caller runs the plain launches instead (and stops trying graphs).
template <class Issue>
static inline bool capture_graph(cudaStream_t stream, cudaGraphExec_t* exec, Issue&& issue) {
  *exec = nullptr;
  if (cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  issue();
  const cudaError_t launch_err = cudaGetLastError();
  cudaGraph_t graph = nullptr;
  const cudaError_t end_err = cudaStreamEndCapture(stream, &graph);  // also ends an invalidated capture
  bool ok = launch_err == cudaSuccess && end_err == cudaSuccess && graph != nullptr;
  if (ok && cudaGraphInstantiate(exec, graph, 0) != cudaSuccess) { *exec = nullptr; ok = false; }
  if (graph) cudaGraphDestroy(graph);
  if (!ok) cudaGetLastError();
  return ok;
}

}  // namespace orbfe
