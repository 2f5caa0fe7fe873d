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
// That is the round-1 kernel (hamming_bruteforce_kernel, kept selectable with ORBFE_BF_KERNEL=1 for A/B runs).  The
// default since round 2 is hamming_bruteforce2_kernel further down: 16 LOP3 + 4 POPC per pair instead of 27 + 5, a
// threshold-gated top-2 update and bulk-TMA map tiles (17.7 -> 11.1 ms for 8192 x 1 M).  The tensor-core variant of the
// same search lives in k_match_mma.cu.
#include <climits>
#include <cstdlib>

#include "match_kernels.h"

namespace orbfe {

#ifndef ORBFE_BF_T
#define ORBFE_BF_T 128
#endif
#ifndef ORBFE_BF_GROUP
#define ORBFE_BF_GROUP 4
#endif
constexpr int BF_T = ORBFE_BF_T;  // threads per CTA
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

// ---- v2 of the brute-force kernel (round 2): fewer ALU operations per pair, map tiles by bulk TMA ----
// The round-1 kernel above spends 27 LOP3 per pair (8 XOR + carry-save tree) and is bound by the ALU pipe (85 % busy),
// not by POPC.  Here the XOR is split off the LINEAR parts of the adder tree: with x_i = q_i ^ m_i,
//   l1 = x0^x1^x2 = (q0^q1^q2) ^ (m0^m1^m2),  l2 likewise for 3..5,  l3 = l1^l2^x6 = Q3 ^ M3,
// where the q-terms are per-query constants and the m-terms are computed once per map row and shared by the thread's
// BF_QPT queries; only the three majorities need the x_i themselves.  A fourth carry-save adder folds the three
// weight-2 words into one weight-2 and one weight-4 word: 16 LOP3 + 4 POPC per pair (was 27 + 5), and
//   d = popc(l3) + popc(x7) + 2 popc(hs) + 4 popc(hc).
// The running top-2 is only touched when a distance beats the query's current second best (one compare per pair and
// ONE rarely-taken branch per map row for the thread's queries) instead of 3 min/max + a key per pair: rows are scanned
// in increasing index, so a later row with an equal distance never wins and ties still go to the lowest index.
// Map tiles arrive by cp.async.bulk (one 8 KB bulk copy per tile, issued by one thread, completing on an mbarrier),
// double-buffered: the LSU only serves the broadcast reads of the tile.
// LOP3 through inline asm: opaque to the compiler's algebra, which otherwise re-associates the XORs into the majorities
// (5 LOP3 per majority instead of 3 + 1) and re-derives the shared map-row terms once per query
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ uint32_t xor2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("xor.b32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}
__device__ __forceinline__ unsigned bf_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bf_bulk_load(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  const unsigned b = bf_smem_u32(bar);
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(b), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::
                   "r"(bf_smem_u32(dst)), "l"(src), "r"(bytes), "r"(b)
               : "memory");
}
__device__ __forceinline__ void bf_mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "BF_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra BF_DONE;\n"
      "bra BF_WAIT;\n"
      "BF_DONE:\n"
      "}\n" ::"r"(bf_smem_u32(bar)),
      "r"(parity)
      : "memory");
}

constexpr int BF_GROUP = ORBFE_BF_GROUP;
__device__ __forceinline__ uint32_t bf_dist(const uint32_t (&q)[8], uint32_t qa, uint32_t qb, uint32_t qc,
                                            const uint4& lo, const uint4& hi, uint32_t ma, uint32_t mb, uint32_t mc) {
  const uint32_t h1 = maj3(xor2(q[0], lo.x), xor2(q[1], lo.y), xor2(q[2], lo.z));
  const uint32_t h2 = maj3(xor2(q[3], lo.w), xor2(q[4], hi.x), xor2(q[5], hi.y));
  const uint32_t h3 = maj3(xor2(qa, ma), xor2(qb, mb), xor2(q[6], hi.z));
  const uint32_t hs = xor3(h1, h2, h3), hc = maj3(h1, h2, h3);
  return __popc(xor2(qc, mc)) + __popc(xor2(q[7], hi.w)) + 2u * __popc(hs) + 4u * __popc(hc);
}

template <int QPT>
__global__ void __launch_bounds__(BF_T) hamming_bruteforce2_kernel(const uint4* __restrict__ query, int nq,
                                                                   const uint4* __restrict__ map, int nm, int slice_len,
                                                                   long long index_base,
                                                                   unsigned long long* __restrict__ part_best,
                                                                   unsigned long long* __restrict__ part_second) {
  __shared__ __align__(128) uint4 tile[2][BF_TILE * 2];
  __shared__ unsigned long long bar[2];
  const int tid = threadIdx.x;
  const int q0 = (blockIdx.x * BF_T + tid) * QPT;
  const int m_begin = blockIdx.y * slice_len;
  const int m_end = min(nm, m_begin + slice_len);
  const int ntiles = (m_end - m_begin + BF_TILE - 1) / BF_TILE;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(bf_smem_u32(&bar[0])) : "memory");
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(bf_smem_u32(&bar[1])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    for (int t = 0; t < 2 && t < ntiles; ++t) {
      const int t0 = m_begin + t * BF_TILE;
      bf_bulk_load(tile[t], map + 2 * (size_t)t0, (unsigned)min(BF_TILE, m_end - t0) * 32u, &bar[t]);
    }
  }
  uint32_t q[QPT][8], qa[QPT], qb[QPT], qc[QPT];
#pragma unroll
  for (int k = 0; k < QPT; ++k) {
    const int qi = min(q0 + k, nq - 1);
    const uint4 lo = __ldg(query + 2 * qi), hi = __ldg(query + 2 * qi + 1);
    q[k][0] = lo.x; q[k][1] = lo.y; q[k][2] = lo.z; q[k][3] = lo.w;
    q[k][4] = hi.x; q[k][5] = hi.y; q[k][6] = hi.z; q[k][7] = hi.w;
    qa[k] = lo.x ^ lo.y ^ lo.z;
    qb[k] = lo.w ^ hi.x ^ hi.y;
    qc[k] = qa[k] ^ qb[k] ^ hi.z;
  }
  uint32_t b1[QPT], b2[QPT], thr[QPT];  // keys dist << 20 | row within the slice; thr = distance of the second best
#pragma unroll
  for (int k = 0; k < QPT; ++k) { b1[k] = 0xFFFFFFFFu; b2[k] = 0xFFFFFFFFu; thr[k] = 0xFFFu; }
  __syncthreads();  // the barriers are initialised before anyone waits on them

  for (int t = 0; t < ntiles; ++t) {
    const int buf = t & 1;
    const int t0 = m_begin + t * BF_TILE;
    const int tn = min(BF_TILE, m_end - t0);
    bf_mbar_wait(&bar[buf], (unsigned)(t >> 1) & 1u);
    const uint4* tl = tile[buf];
    const uint32_t jbase = (uint32_t)(t0 - m_begin);
    // BF_GROUP rows at a time: the distances of a group feed a running minimum per query (no branch between rows, so
    // the rows' instruction streams overlap) and stay in registers; only when one of them beats a second best as it
    // stood BEFORE the group (a superset of the rows that update anything) are the group's distances walked again, in
    // row order, for the exact update.  A warp takes that path for ~10 % of the rows of a slice (128 queries x 2 / row).
    for (int j0 = 0; j0 < tn; j0 += BF_GROUP) {
      uint32_t md[QPT], d[BF_GROUP][QPT];
#pragma unroll
      for (int k = 0; k < QPT; ++k) md[k] = 0xFFFu;
#pragma unroll
      for (int g = 0; g < BF_GROUP; ++g) {
        const int j = min(j0 + g, tn - 1);  // a short last group repeats its last row (skipped in the update below)
        const uint4 lo = tl[2 * j], hi = tl[2 * j + 1];
        const uint32_t ma = xor3(lo.x, lo.y, lo.z), mb = xor3(lo.w, hi.x, hi.y), mc = xor3(ma, mb, hi.z);
#pragma unroll
        for (int k = 0; k < QPT; ++k) {
          d[g][k] = bf_dist(q[k], qa[k], qb[k], qc[k], lo, hi, ma, mb, mc);
          md[k] = min(md[k], d[g][k]);
        }
      }
      bool any = false;
#pragma unroll
      for (int k = 0; k < QPT; ++k) any |= md[k] < thr[k];
      if (any) {
#pragma unroll
        for (int g = 0; g < BF_GROUP; ++g) {
          if (j0 + g < tn) {
            const uint32_t jj = jbase + (uint32_t)(j0 + g);
#pragma unroll
            for (int k = 0; k < QPT; ++k) {
              if (d[g][k] < thr[k]) {
                const uint32_t key = (d[g][k] << 20) + jj;
                b2[k] = min(b2[k], max(b1[k], key));
                b1[k] = min(b1[k], key);
                thr[k] = b2[k] >> 20;
              }
            }
          }
        }
      }
    }
    __syncthreads();  // every thread has read this buffer
    if (tid == 0 && t + 2 < ntiles) {
      const int n0 = m_begin + (t + 2) * BF_TILE;
      asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
      bf_bulk_load(tile[buf], map + 2 * (size_t)n0, (unsigned)min(BF_TILE, m_end - n0) * 32u, &bar[buf]);
    }
  }
#pragma unroll
  for (int k = 0; k < QPT; ++k) {
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

int bruteforce_slices(int nq, int nm, int* slice_len, int max_slices) {
  // 12 CTAs of 128 threads per SM's worth of slices (5 are resident at 96 registers; short slices keep the tail
  // small: 3 / 6 / 12 per SM ran the kernel at 514 / 744 / 772 G pairs/s), slices a multiple of the tile, at most
  // 2^20 rows per slice;
  // max_slices > 0 caps the number of slices (callers with a fixed partial-result buffer)
  const int qblocks = (nq + BF_T * BF_QPT - 1) / (BF_T * BF_QPT);
  static const int waves = [] { const char* e = getenv("ORBFE_BF_WAVES"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 12; }();
  int want = (148 * waves + qblocks - 1) / qblocks;
  if (max_slices > 0 && want > max_slices) want = max_slices;
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
  static const int version = [] { const char* e = getenv("ORBFE_BF_KERNEL"); return e && e[0] == '1' ? 1 : 2; }();
  if (version == 1)
    hamming_bruteforce_kernel<<<grid, BF_T, 0, s>>>(reinterpret_cast<const uint4*>(query), nq,
                                                    reinterpret_cast<const uint4*>(map), nm, slice_len, index_base,
                                                    part_best, part_second);
  else
    hamming_bruteforce2_kernel<BF_QPT><<<grid, BF_T, 0, s>>>(reinterpret_cast<const uint4*>(query), nq,
                                                             reinterpret_cast<const uint4*>(map), nm, slice_len,
                                                             index_base, part_best, part_second);
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
