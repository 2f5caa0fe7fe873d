// k_match.cu — Hamming kernels of the ORBmatcher path.
//
// ORBmatcher::DescriptorDistance (reference: src/ORBmatcher.cc:1715-1731) is a SWAR
// popcount of 8 x 32-bit XORs; on the device it is 8 x POPC.  The brute-force kernel
// (BASELINE.json config 5) is bound by the POPC pipe (16 results/clk/SM), so it is
// organised to keep every other cost off the critical pipe:
//   * each thread keeps QPT query descriptors in registers (8 words each);
//   * a CTA streams its slice of the map through shared memory in tiles; all lanes
//     read the same map descriptor (2 x LDS.128 broadcast) and reuse it for QPT pairs;
//   * best / second best are tracked as 32-bit keys dist<<20 | local index, updated
//     branch-free with 3 min/max (ties -> lowest index falls out of the packing);
//   * per (query, map slice) partial top-2 go to a small workspace and a second kernel
//     merges slices (and, multi-GPU, shards) as 64-bit keys dist<<32 | global index.
#include <climits>

#include "match_kernels.h"

namespace orbfe {

constexpr int BF_T = 128;        // threads per CTA
constexpr int BF_QPT = 4;        // queries per thread
constexpr int BF_TILE = 256;     // map descriptors per shared-memory tile (8 KB)

// popcount of 256 bits with 5 POPC instead of 8: three carry-save adders (2 LOP3 each: sum = a^b^c, carry = majority)
// compress the 8 words into two "ones" words and three "twos" words.  POPC issues at 16 lanes/clk/SM, LOP3 at 64, so
// trading 3 POPC for 6 LOP3 moves the brute-force loop from the POPC pipe towards a POPC/ALU balance.
__device__ __forceinline__ uint32_t hamming256_csa(uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, uint32_t x4,
                                                   uint32_t x5, uint32_t x6, uint32_t x7) {
  const uint32_t l1 = x0 ^ x1 ^ x2, h1 = (x0 & x1) | (x2 & (x0 | x1));
  const uint32_t l2 = x3 ^ x4 ^ x5, h2 = (x3 & x4) | (x5 & (x3 | x4));
  const uint32_t l3 = l1 ^ l2 ^ x6, h3 = (l1 & l2) | (x6 & (l1 | l2));
  return __popc(l3) + __popc(x7) + 2u * (__popc(h1) + __popc(h2) + __popc(h3));
}

// register-only POPC throughput probe (the roofline denominator of the Hamming kernels): every thread runs
// iters x 8 independent popc chains; out keeps the compiler honest
__global__ void __launch_bounds__(256) popc_peak_kernel(int iters, uint32_t seed, uint32_t* __restrict__ out) {
  uint32_t v[8], acc[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { v[i] = seed * (threadIdx.x + 1u) + i * 0x9E3779B9u; acc[i] = 0; }
#pragma unroll 1
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      acc[i] += __popc(v[i]);
      v[i] ^= acc[i] << 3;  // data dependence so the popc cannot be hoisted; LOP3/SHF go to the ALU pipe
    }
  }
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += acc[i];
  if (s == 0xFFFFFFFFu) out[0] = s;
}

double measure_popc_peak(cudaStream_t s) {
  uint32_t* d = nullptr;
  if (cudaMalloc(&d, 4) != cudaSuccess) return 0.0;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int sms = 148;
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int iters = 8192, blocks = sms * 8;
  popc_peak_kernel<<<blocks, 256, 0, s>>>(64, 1u, d);  // warm-up
  double best = 0.0;
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0, s);
    popc_peak_kernel<<<blocks, 256, 0, s>>>(iters, 12345u + rep, d);
    cudaEventRecord(e1, s);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double rate = (double)blocks * 256.0 * iters * 8.0 / (ms * 1e-3);
    if (rate > best) best = rate;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  return best;
}

__global__ void descriptor_distance_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int n,
                                           int32_t* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 a0 = __ldg(a + 2 * i), a1 = __ldg(a + 2 * i + 1), b0 = __ldg(b + 2 * i), b1 = __ldg(b + 2 * i + 1);
  out[i] = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// grid: (ceil(nq / (BF_T*BF_QPT)), n_slices).  Slice s covers map rows [s*slice_len, min(nm,(s+1)*slice_len)).
__global__ void __launch_bounds__(BF_T) hamming_bruteforce_kernel(const uint4* __restrict__ query, int nq,
                                                                  const uint4* __restrict__ map, int nm, int slice_len,
                                                                  long long index_base,
                                                                  unsigned long long* __restrict__ part_best,
                                                                  unsigned long long* __restrict__ part_second) {
  __shared__ uint4 tile[BF_TILE * 2];
  const int tid = threadIdx.x;
  const int q0 = (blockIdx.x * BF_T + tid) * BF_QPT;
  const int m_begin = blockIdx.y * slice_len;
  const int m_end = min(nm, m_begin + slice_len);
  uint32_t q[BF_QPT][8];
#pragma unroll
  for (int k = 0; k < BF_QPT; ++k) {
    const int qi = min(q0 + k, nq - 1);
    const uint4 lo = __ldg(query + 2 * qi), hi = __ldg(query + 2 * qi + 1);
    q[k][0] = lo.x; q[k][1] = lo.y; q[k][2] = lo.z; q[k][3] = lo.w;
    q[k][4] = hi.x; q[k][5] = hi.y; q[k][6] = hi.z; q[k][7] = hi.w;
  }
  uint32_t b1[BF_QPT], b2[BF_QPT];
#pragma unroll
  for (int k = 0; k < BF_QPT; ++k) { b1[k] = 0xFFFFFFFFu; b2[k] = 0xFFFFFFFFu; }

  for (int t0 = m_begin; t0 < m_end; t0 += BF_TILE) {
    const int tn = min(BF_TILE, m_end - t0);
    __syncthreads();
    for (int i = tid; i < tn * 2; i += BF_T) tile[i] = __ldg(map + 2 * (size_t)t0 + i);
    __syncthreads();
    const uint32_t jbase = (uint32_t)(t0 - m_begin);
#pragma unroll 4
    for (int j = 0; j < tn; ++j) {
      const uint4 lo = tile[2 * j], hi = tile[2 * j + 1];
#pragma unroll
      for (int k = 0; k < BF_QPT; ++k) {
        const uint32_t d = hamming256_csa(q[k][0] ^ lo.x, q[k][1] ^ lo.y, q[k][2] ^ lo.z, q[k][3] ^ lo.w,
                                          q[k][4] ^ hi.x, q[k][5] ^ hi.y, q[k][6] ^ hi.z, q[k][7] ^ hi.w);
        const uint32_t key = (d << 20) + (jbase + (uint32_t)j);
        b2[k] = min(b2[k], max(b1[k], key));
        b1[k] = min(b1[k], key);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < BF_QPT; ++k) {
    const int qi = q0 + k;
    if (qi < nq) {
      const size_t o = (size_t)blockIdx.y * nq + qi;
      const unsigned long long base = (unsigned long long)(index_base + m_begin);
      part_best[o] = b1[k] == 0xFFFFFFFFu ? ~0ull
                                          : ((unsigned long long)(b1[k] >> 20) << 32) | (base + (b1[k] & 0xFFFFFu));
      part_second[o] = b2[k] == 0xFFFFFFFFu ? ~0ull
                                            : ((unsigned long long)(b2[k] >> 20) << 32) | (base + (b2[k] & 0xFFFFFu));
    }
  }
}

// top-2 of the union of `parts` (best, second) pairs per query; arrays laid out [part][nq]
__global__ void hamming_merge_kernel(const unsigned long long* __restrict__ best_all,
                                     const unsigned long long* __restrict__ second_all, int parts, int nq,
                                     unsigned long long* __restrict__ best, unsigned long long* __restrict__ second) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  unsigned long long b1 = ~0ull, b2 = ~0ull;
  for (int p = 0; p < parts; ++p) {
    const unsigned long long x = best_all[(size_t)p * nq + qi], y = second_all[(size_t)p * nq + qi];
    if (x < b1) { b2 = b1; b1 = x; } else if (x < b2) b2 = x;
    if (y < b2) b2 = y;  // y >= x, so it can only be a second best
  }
  best[qi] = b1;
  second[qi] = b2;
}

void launch_descriptor_distance(const uint8_t* a, const uint8_t* b, int n, int32_t* out, cudaStream_t s) {
  descriptor_distance_kernel<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint4*>(a),
                                                             reinterpret_cast<const uint4*>(b), n, out);
}

int bruteforce_slices(int nq, int nm, int* slice_len) {
  // enough CTAs for ~3 waves of 148 SMs, slices a multiple of the tile, at most 2^20 rows per slice
  const int qblocks = (nq + BF_T * BF_QPT - 1) / (BF_T * BF_QPT);
  int want = (148 * 3 + qblocks - 1) / qblocks;
  int len = (nm + want - 1) / want;
  len = (len + BF_TILE - 1) / BF_TILE * BF_TILE;
  if (len < BF_TILE) len = BF_TILE;
  if (len > (1 << 20)) len = 1 << 20;
  *slice_len = len;
  return (nm + len - 1) / len;
}

void launch_hamming_bruteforce(const uint8_t* query, int nq, const uint8_t* map, int nm, int slice_len, int slices,
                               long long index_base, unsigned long long* part_best, unsigned long long* part_second,
                               cudaStream_t s) {
  dim3 grid((nq + BF_T * BF_QPT - 1) / (BF_T * BF_QPT), slices);
  hamming_bruteforce_kernel<<<grid, BF_T, 0, s>>>(reinterpret_cast<const uint4*>(query), nq,
                                                  reinterpret_cast<const uint4*>(map), nm, slice_len, index_base,
                                                  part_best, part_second);
}

void launch_hamming_merge(const unsigned long long* best_all, const unsigned long long* second_all, int parts, int nq,
                          unsigned long long* best, unsigned long long* second, cudaStream_t s) {
  hamming_merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(best_all, second_all, parts, nq, best, second);
}


// ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:252-318), batched over map points:
// one warp per point; for every observation i the median of its Hamming distances to all observations of the point
// (self included, element floor(0.5*(N-1)) of the sorted row, :303-305) is found by a 9-step bisection on the value
// range [0, 256] with warp-wide counts; the first observation with the least median wins (:307-311).
__device__ __forceinline__ int hamming_regs(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__global__ void __launch_bounds__(256) distinctive_kernel(const uint8_t* __restrict__ desc,
                                                          const int32_t* __restrict__ offsets, int n_points,
                                                          int32_t* __restrict__ best_idx,
                                                          uint8_t* __restrict__ best_desc) {
  const int lane = threadIdx.x & 31;
  const int p = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (p >= n_points) return;
  const int beg = offsets[p], N = offsets[p + 1] - beg;
  if (N <= 0) {
    if (lane == 0) best_idx[p] = -1;  // "if(vDescriptors.empty()) return;": the descriptor stays as it was
    return;
  }
  const uint4* D = reinterpret_cast<const uint4*>(desc + (size_t)beg * 32);
  const int k = (int)(0.5 * (double)(N - 1));  // vDists[0.5*(N-1)]
  int best_median = INT_MAX, best_i = 0;
  if (N <= 32) {  // one observation per lane: the row of distances lives in registers
    uint4 m0 = make_uint4(0, 0, 0, 0), m1 = m0;
    if (lane < N) { m0 = __ldg(D + 2 * lane); m1 = __ldg(D + 2 * lane + 1); }
    for (int i = 0; i < N; ++i) {
      uint4 a0, a1;
      a0.x = __shfl_sync(0xffffffffu, m0.x, i); a0.y = __shfl_sync(0xffffffffu, m0.y, i);
      a0.z = __shfl_sync(0xffffffffu, m0.z, i); a0.w = __shfl_sync(0xffffffffu, m0.w, i);
      a1.x = __shfl_sync(0xffffffffu, m1.x, i); a1.y = __shfl_sync(0xffffffffu, m1.y, i);
      a1.z = __shfl_sync(0xffffffffu, m1.z, i); a1.w = __shfl_sync(0xffffffffu, m1.w, i);
      const int d = lane < N ? hamming_regs(a0, a1, m0, m1) : 1 << 20;
      int lo = 0, hi = 256;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const int c = __popc(__ballot_sync(0xffffffffu, d <= mid));
        if (c >= k + 1) hi = mid; else lo = mid + 1;
      }
      if (lo < best_median) { best_median = lo; best_i = i; }
    }
  } else {  // long observation lists: a 257-bin histogram of the row in shared memory, k-th element by a warp scan
    __shared__ int s_hist[8][288];
    int* hist = s_hist[threadIdx.x >> 5];
    for (int i = 0; i < N; ++i) {
#pragma unroll
      for (int b = 0; b < 9; ++b) hist[lane * 9 + b] = 0;
      __syncwarp();
      const uint4 a0 = __ldg(D + 2 * i), a1 = __ldg(D + 2 * i + 1);
      for (int j = lane; j < N; j += 32) atomicAdd(&hist[hamming_regs(a0, a1, __ldg(D + 2 * j), __ldg(D + 2 * j + 1))], 1);
      __syncwarp();
      int mine = 0;
#pragma unroll
      for (int b = 0; b < 9; ++b) mine += hist[lane * 9 + b];
      int incl = mine;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const int excl = incl - mine;
      int median = -1;
      if (excl <= k && k < incl) {  // exactly one lane owns the k-th smallest
        int run = excl;
#pragma unroll
        for (int b = 0; b < 9; ++b) {
          run += hist[lane * 9 + b];
          if (median < 0 && k < run) median = lane * 9 + b;
        }
      }
      const unsigned owner = __ballot_sync(0xffffffffu, median >= 0);
      median = __shfl_sync(0xffffffffu, median, __ffs(owner) - 1);
      if (median < best_median) { best_median = median; best_i = i; }
      __syncwarp();
    }
  }
  if (lane == 0) best_idx[p] = best_i;
  if (best_desc && lane < 8)
    reinterpret_cast<uint32_t*>(best_desc + (size_t)p * 32)[lane] =
        reinterpret_cast<const uint32_t*>(desc + (size_t)(beg + best_i) * 32)[lane];
}

void launch_distinctive(const uint8_t* desc, const int32_t* offsets, int n_points, int32_t* best_idx,
                        uint8_t* best_desc, cudaStream_t s) {
  if (n_points > 0) distinctive_kernel<<<(n_points + 7) / 8, 256, 0, s>>>(desc, offsets, n_points, best_idx, best_desc);
}

}  // namespace orbfe
