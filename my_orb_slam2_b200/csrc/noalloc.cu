// noalloc.cu — orbfe_hamming_bruteforce_device_noalloc: the brute-force Hamming top-2 of orbfe_hamming_bruteforce_device
// for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// The existing entry point keeps the per-slice partial results (2 x slices x nq u64) in the matcher's growable candidate
// buffer (cudaMalloc).  Here they live in a file-scope static __device__ array of a fixed, modest size: every GPU that
// loads the library holds 4 MiB (kScratchEntries u64) until exit.  This file calls no cudaMalloc / cudaFree / allocating
// library routine; the kernels are the same ones (k_match.cu), launched on the caller's stream.
//
// Sharing: EVERY call on a device uses the same array.  Calls that overlap on the GPU (different streams, different
// host threads) race on it: serialise them (one stream, or events between streams).  The array keeps the last call's
// contents; nothing has to start at zero, because every entry the merge kernel reads (slices x nq of each half) is
// written by the brute-force kernel of the same call first.
#include <mutex>

#include "common.cuh"
#include "match_kernels.h"

namespace orbfe {

constexpr int kNoAllocMaxQueries = 32768;                      // stated limit on nq
constexpr size_t kScratchEntries = (size_t)512 * 1024;          // u64 entries: 4 MiB
static_assert(kScratchEntries * sizeof(unsigned long long) == (size_t)4 << 20, "the documented size: 4 MiB per GPU");
constexpr int kMaxDevices = 64;

// In HOST code this name is a host-side stand-in, not a device address: never pass it to a kernel; host code uses the
// pointer cudaGetSymbolAddress returns for the current device (below).
static __device__ __align__(256) unsigned long long g_bf_scratch[kScratchEntries];

static std::mutex g_scratch_mu;
static unsigned long long* g_scratch_ptr[kMaxDevices] = {};  // per device, resolved once with that device current

// the device address of g_bf_scratch on the CURRENT device (set-up: once per device)
static int scratch_for_current_device(unsigned long long** out) {
  int dev = -1;
  ORBFE_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("orbfe_hamming_bruteforce_device_noalloc: device index %d outside [0, %d)", dev, kMaxDevices);
    return ORBFE_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    ORBFE_CUDA(cudaGetSymbolAddress(&p, g_bf_scratch));
    if (!p) {
      set_error("orbfe_hamming_bruteforce_device_noalloc: cudaGetSymbolAddress returned a null address");
      return ORBFE_ERR_CUDA;
    }
    g_scratch_ptr[dev] = static_cast<unsigned long long*>(p);
  }
  *out = g_scratch_ptr[dev];
  return ORBFE_OK;
}

}  // namespace orbfe

using namespace orbfe;

extern "C" {

int orbfe_hamming_bruteforce_device_noalloc(void* stream, const uint8_t* d_query, int nq, const uint8_t* d_map, int nm,
                                            int64_t index_base, uint64_t* d_best, uint64_t* d_second) {
  ORBFE_ARG(d_query && d_best && d_second && nq >= 1 && nm >= 0);
  ORBFE_ARG(nm == 0 || d_map);
  ORBFE_ARG(index_base >= 0 && index_base + (int64_t)nm <= ((int64_t)1 << 32));
  if (nq > kNoAllocMaxQueries) {
    set_error("orbfe_hamming_bruteforce_device_noalloc: nq = %d exceeds the static scratch limit of %d queries "
              "(use orbfe_hamming_bruteforce_device, which allocates)", nq, kNoAllocMaxQueries);
    return ORBFE_ERR_CAPACITY;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (nm == 0) {  // no map rows: both keys are "none"
    ORBFE_CUDA(cudaMemsetAsync(d_best, 0xFF, (size_t)nq * 8, s));
    ORBFE_CUDA(cudaMemsetAsync(d_second, 0xFF, (size_t)nq * 8, s));
    return ORBFE_OK;
  }
  int slice_len = 0;
  // the slicing is capped so that the partial results fit the static array (fewer, longer slices for many queries)
  const int slices = bruteforce_slices(nq, nm, &slice_len, (int)(kScratchEntries / ((size_t)2 * nq)));
  const size_t need = (size_t)slices * nq * 2;
  if (need > kScratchEntries) {
    set_error("orbfe_hamming_bruteforce_device_noalloc: %d slices x %d queries need %zu scratch entries > %zu "
              "(use orbfe_hamming_bruteforce_device, which allocates)", slices, nq, need, (size_t)kScratchEntries);
    return ORBFE_ERR_CAPACITY;
  }
  unsigned long long* scratch = nullptr;
  const int rc = scratch_for_current_device(&scratch);
  if (rc != ORBFE_OK) return rc;
  unsigned long long* pb = scratch;
  unsigned long long* ps = scratch + (size_t)slices * nq;
  launch_hamming_bruteforce(d_query, nq, d_map, nm, slice_len, slices, index_base, pb, ps, s);
  launch_hamming_merge(pb, ps, slices, nq, reinterpret_cast<unsigned long long*>(d_best),
                       reinterpret_cast<unsigned long long*>(d_second), s);
  ORBFE_CUDA(cudaGetLastError());
  return ORBFE_OK;
}

}  // extern "C"
