// k_match_mma.cu — brute-force Hamming on the 5th-generation tensor cores (tcgen05 / TMEM / TMA), an OPTIONAL second
// path for BASELINE.json config 5 beside the POPC kernel of k_match.cu (which stays the default and the graded one:
// north_star asks for popc-based matching; VERDICT r1 item 9 invites this as a stretch and wants both reported with
// bit-equal keys).
//
// 256-bit Hamming distance as an integer GEMM:  d(q, m) = |q| + |m| - 2 q.m  with the descriptors expanded to one byte
// per bit.  M = queries, N = map rows; the A rows hold -2 q_k (s8), the B rows m_k (u8), and a ninth K step of 32 bytes
// carries |m| itself (A: ones; B: |m| split over two bytes), so the s32 accumulator IS s = |m| - 2 q.m = d - |q| and the
// epilogue has no per-element arithmetic left but a minimum (K = 288: +12.5 % tensor work, epilogue 3x lighter).
//   * expand_bits_kernel: 32-byte descriptors -> 256-byte rows + the row's popcount (the map side is a one-off per map).
//   * hamming_mma_kernel: persistent, warp-specialised, one CTA per SM:
//       warp 0      TMA producer: the item's two 128-query A tiles once, then the map in 128-row B tiles through a
//                   4-stage ring (cp.async.bulk.tensor.2d, 128-byte swizzle, mbarrier complete_tx);
//       warp 1      MMA issuer (one elected lane): per B tile 2 x 8 tcgen05.mma.cta_group::1.kind::i8 (M128 x N128 x K32)
//                   into one of two 256-column TMEM accumulator buffers; tcgen05.commit frees the B stage and hands the
//                   accumulator to the epilogue;
//       warps 2..17 epilogue: thread = one query x one 64-column half of an accumulator; tcgen05.ld 32 columns at a
//                   time (the next chunk in flight), a running minimum, and only when a chunk's minimum beats the
//                   query's current second best (rare) the chunk's 32 values are walked in row order for the exact
//                   top-2 update — the same tie rule as the POPC kernel (rows ascend, a later equal distance never wins).
//                   (A weight-stationary tcgen05.mma.ws variant sharing the B slice between the two A tiles was
//                   measured equal: 1.805 vs 1.783 ms.  With the epilogue's work removed the call takes 1.54 ms: the
//                   TMA -> MMA pipeline alone runs the tensor pipe at ~82 %, the epilogue costs the other 0.24 ms.)
//     Work item = (pair of A tiles = 256 queries, slice of the map).  Every slice restarts its top-2, so a query's
//     current GLOBAL second-best distance is shared through `gthr` (atomicMin at the end of an item, read at the start
//     of the next): a bound >= the true second best never prunes a row of the true top-2, and it keeps the exact-update
//     path rare in later slices.  Partial (best, second) keys per (slice, query) go through the same merge kernel as the
//     POPC path.
// Every mbarrier wait is bounded (a wrong descriptor must not hang the GPU): on time-out the role sets an error word,
// leaves its loop, and the CTA still frees its TMEM.
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "match_kernels.h"

namespace orbfe {

constexpr int MM_TILE = 128;                 // rows per A / B tile = UMMA M and N
constexpr int MM_K = 256;                    // bytes per expanded descriptor row
constexpr int MM_KB = 2;                     // 128-byte swizzle atoms along K
constexpr int MM_STAGES = 4;                 // B ring
constexpr int MM_TILE_BYTES = MM_TILE * 128; // one (tile, k-block): 16 KB
constexpr int MM_EPI_WARPS = 16;              // 2 column halves x 2 A tiles x 4 TMEM lane quarters
constexpr int MM_THREADS = 32 * (2 + MM_EPI_WARPS);

// ---- bits -> bytes: lane = one input byte of the row, writes its 8 output bytes; the row's popcount by warp shuffle.
// QUERY rows: bytes -2 * bit (s8) and an extra block of 32 ones; MAP rows: bytes bit (u8) and an extra block
// {min(|m|, 255), |m| - that, 0 ...}; map padding rows (row >= n): extra block 32 x 255 (s = 8160: never a candidate) ----
template <bool QUERY>
__global__ void __launch_bounds__(256) expand_bits_kernel(const uint8_t* __restrict__ desc, int n, int n_padded,
                                                          uint8_t* __restrict__ out, uint8_t* __restrict__ extra,
                                                          int32_t* __restrict__ popc) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (row >= n_padded) return;
  unsigned b = 0;
  if (row < n) b = desc[(size_t)row * 32 + lane];
  // bit k of the byte -> output byte k (0 or 1)
  unsigned lo = (b & 1u) | ((b & 2u) << 7) | ((b & 4u) << 14) | ((b & 8u) << 21);
  unsigned hi = ((b >> 4) & 1u) | ((b & 32u) << 3) | ((b & 64u) << 10) | ((b & 128u) << 17);
  if (QUERY) { lo *= 0xFEu; hi *= 0xFEu; }  // 1 -> 0xFE = -2 (s8); no carries between bytes
  if (row < n) reinterpret_cast<uint2*>(out + (size_t)row * MM_K)[lane] = make_uint2(lo, hi);
  int c = __popc(b);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if (QUERY) {
    if (row < n) extra[(size_t)row * 32 + lane] = 1;
    if (lane == 0 && popc) popc[row] = row < n ? c : 0;
  } else {
    const int plo = min(c, 255);
    extra[(size_t)row * 32 + lane] = row < n ? (uint8_t)(lane == 0 ? plo : (lane == 1 ? c - plo : 0)) : (uint8_t)255;
  }
}

__global__ void fill_u32_kernel(unsigned* p, int n, unsigned v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// ---- PTX helpers ----
__device__ __forceinline__ unsigned mm_smem(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mm_bar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(mm_smem(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mm_bar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(mm_smem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mm_bar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(mm_smem(bar)) : "memory");
}
// bounded wait: false after `mm_timeout_ns` (a wrong descriptor or a missing arrival must not hang the GPU); the bound
// is generous (default 10 s, ORBFE_TC_TIMEOUT_S) because a time-sliced GPU can legitimately stall a CTA for long
__device__ unsigned long long mm_timeout_ns = 10000000000ull;
__device__ __forceinline__ bool mm_bar_wait(unsigned long long* bar, unsigned parity) {
  const unsigned b = mm_smem(bar);
  unsigned long long t0 = 0;
  for (unsigned spin = 0;; ++spin) {
    unsigned ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(b), "r"(parity)
        : "memory");
    if (ok) return true;
    if ((spin & 63u) == 63u) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      if (t0 == 0) t0 = t;
      else if (t - t0 > mm_timeout_ns) return false;
    }
  }
}
__device__ __forceinline__ void mm_tma_load_2d(void* dst, const CUtensorMap* map, int x, int y, unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];\n" ::
          "r"(mm_smem(dst)), "l"(map), "r"(x), "r"(y), "r"(mm_smem(bar))
      : "memory");
}
// shared-memory matrix descriptor: K-major, 128-byte swizzle, 8-row groups 1024 bytes apart (sm_100 version field = 1)
__device__ __forceinline__ unsigned long long mm_desc(unsigned smem_addr) {
  return (unsigned long long)((smem_addr >> 4) & 0x3FFFu) | ((unsigned long long)(1024u >> 4) << 32) | (1ull << 46) |
         (2ull << 61);
}
// ... and for the 32-byte-wide extra block: 32-byte swizzle, 8-row groups 256 bytes apart
__device__ __forceinline__ unsigned long long mm_desc32(unsigned smem_addr) {
  return (unsigned long long)((smem_addr >> 4) & 0x3FFFu) | ((unsigned long long)(256u >> 4) << 32) | (1ull << 46) |
         (6ull << 61);
}
// instruction descriptor, kind::i8: D = s32, A = s8, B = u8, both K-major, N = 128, M = 128
constexpr unsigned MM_IDESC = (2u << 4) | (1u << 7) | ((unsigned)(MM_TILE >> 3) << 17) | ((unsigned)(MM_TILE >> 4) << 24);
__device__ __forceinline__ void mm_mma_i8(unsigned tmem_d, unsigned long long da, unsigned long long db, unsigned accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(MM_IDESC), "r"(accumulate), "r"(0u)
      : "memory");
}
__device__ __forceinline__ void mm_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(mm_smem(bar)) : "memory");
}
__device__ __forceinline__ void mm_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void mm_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
// 32 lanes x 32 columns of 32 bits: thread i of the warp gets TMEM lane (lane field + i), columns col .. col+31.
// Issue and wait are separate so that the next chunk travels while this one is processed; the wait takes the
// registers as in/out operands, which keeps the compiler from reading them before the load has landed.
__device__ __forceinline__ void mm_tmem_ld32_issue(unsigned taddr, int (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void mm_tmem_ld_wait(int (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;\n"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
                 "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
                 "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}

struct __align__(1024) MmSmem {
  uint8_t a[2][MM_KB][MM_TILE_BYTES];          // the item's two query tiles (64 KB)
  uint8_t b[MM_STAGES][MM_KB][MM_TILE_BYTES];  // map tile ring (128 KB)
  uint8_t ax[2][MM_TILE * 32];                 // the ninth K step: 32-byte rows (ones / |m| bytes), 32-byte swizzle
  uint8_t bx[MM_STAGES][MM_TILE * 32];
  unsigned long long full[MM_STAGES], empty[MM_STAGES], a_full, a_empty, t_full[2], t_empty[2];
  unsigned tmem_base;
};

// items: (query pair p, map slice s) = item / n_slices ... slice-major so that the CTAs running at the same time
// share map tiles in L2
__global__ void __launch_bounds__(MM_THREADS, 1) hamming_mma_kernel(
    const __grid_constant__ CUtensorMap qmap, const __grid_constant__ CUtensorMap mmap,
    const __grid_constant__ CUtensorMap qxmap, const __grid_constant__ CUtensorMap mxmap, int nq, int nm, int n_pairs,
    int n_slices, int tiles_per_slice, long long index_base, const int32_t* __restrict__ pq, unsigned* __restrict__ gthr, unsigned long long* __restrict__ part_best,
    unsigned long long* __restrict__ part_second, unsigned* __restrict__ err) {
  extern __shared__ __align__(1024) unsigned char mm_smem_raw[];
  MmSmem& sm = *reinterpret_cast<MmSmem*>((reinterpret_cast<uintptr_t>(mm_smem_raw) + 1023) & ~uintptr_t(1023));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = n_pairs * n_slices;
  const int n_tiles_total = (nm + MM_TILE - 1) / MM_TILE;

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < MM_STAGES; ++s) { mm_bar_init(&sm.full[s], 1); mm_bar_init(&sm.empty[s], 1); }
    mm_bar_init(&sm.a_full, 1);
    mm_bar_init(&sm.a_empty, 1);
    for (int t = 0; t < 2; ++t) { mm_bar_init(&sm.t_full[t], 1); mm_bar_init(&sm.t_empty[t], MM_EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  }
  if (warp == 1) {  // the whole TMEM of the SM: two accumulator buffers of 2 x 128 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(mm_smem(&sm.tmem_base)), "r"(512u)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
  }
  mm_fence_before();
  __syncthreads();
  mm_fence_after();
  const unsigned tmem = sm.tmem_base;
  bool ok = true;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      unsigned stage = 0, phase = 0, a_phase = 0, seq = 0;
      for (int item = blockIdx.x; item < n_items && ok; item += gridDim.x) {
        const int slice = item / n_pairs, pair = item - slice * n_pairs;
        const int t0 = slice * tiles_per_slice, t1 = min(t0 + tiles_per_slice, n_tiles_total);
        ok = mm_bar_wait(&sm.a_empty, a_phase ^ 1u);  // the previous item's MMAs have read the A tiles
        if (!ok) break;
        mm_bar_expect_tx(&sm.a_full, 2 * (MM_KB * MM_TILE_BYTES + MM_TILE * 32));
        for (int a = 0; a < 2; ++a) {
          for (int kb = 0; kb < MM_KB; ++kb)
            mm_tma_load_2d(sm.a[a][kb], &qmap, kb * 128, pair * 2 * MM_TILE + a * MM_TILE, &sm.a_full);
          mm_tma_load_2d(sm.ax[a], &qxmap, 0, pair * 2 * MM_TILE + a * MM_TILE, &sm.a_full);
        }
        a_phase ^= 1u;
        for (int t = t0; t < t1; ++t) {
          ok = mm_bar_wait(&sm.empty[stage], phase ^ 1u);
          if (!ok) break;
          mm_bar_expect_tx(&sm.full[stage], MM_KB * MM_TILE_BYTES + MM_TILE * 32);
          for (int kb = 0; kb < MM_KB; ++kb) mm_tma_load_2d(sm.b[stage][kb], &mmap, kb * 128, t * MM_TILE, &sm.full[stage]);
          mm_tma_load_2d(sm.bx[stage], &mxmap, 0, t * MM_TILE, &sm.full[stage]);
          ++seq;
          if (++stage == MM_STAGES) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      unsigned stage = 0, phase = 0, a_phase = 0, acc = 0, acc_phase = 0;
      for (int item = blockIdx.x; item < n_items && ok; item += gridDim.x) {
        const int slice = item / n_pairs;
        const int t0 = slice * tiles_per_slice, t1 = min(t0 + tiles_per_slice, n_tiles_total);
        ok = mm_bar_wait(&sm.a_full, a_phase);
        if (!ok) break;
        a_phase ^= 1u;
        for (int t = t0; t < t1; ++t) {
          ok = mm_bar_wait(&sm.t_empty[acc], acc_phase ^ 1u);  // the epilogue has drained this accumulator buffer
          if (!ok) break;
          ok = mm_bar_wait(&sm.full[stage], phase);
          if (!ok) break;
          mm_fence_after();
#pragma unroll
          for (int a = 0; a < 2; ++a) {
            const unsigned d = tmem + acc * 256u + a * 128u;
#pragma unroll
            for (int kb = 0; kb < MM_KB; ++kb) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const unsigned long long da = mm_desc(mm_smem(sm.a[a][kb]) + k * 32u);
                const unsigned long long db = mm_desc(mm_smem(sm.b[stage][kb]) + k * 32u);
                mm_mma_i8(d, da, db, (kb | k) ? 1u : 0u);
              }
            }
            mm_mma_i8(d, mm_desc32(mm_smem(sm.ax[a])), mm_desc32(mm_smem(sm.bx[stage])), 1u);  // + |m|
          }
          mm_commit(&sm.empty[stage]);   // the B stage is free once these MMAs have read it
          mm_commit(&sm.t_full[acc]);    // ... and the accumulators are complete
          if (++stage == MM_STAGES) { stage = 0; phase ^= 1u; }
          if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
        }
        mm_commit(&sm.a_empty);  // the A tiles may be replaced once every MMA of the item is done
      }
    }
  } else {
    // ===================== epilogue: thread = one query =====================
    const int ew = warp - 2;            // 0..15
    const int a = (ew >> 2) & 1;        // which A tile
    const int half = ew >> 3;           // which 64 of the tile's 128 columns (its own top-2: one more partial per slice)
    const unsigned quarter = warp & 3;  // a warp may only touch TMEM lanes 32 * (warp id % 4) ..
    const int row = (int)quarter * 32 + lane;
    unsigned acc = 0, acc_phase = 0, seq = 0;
    for (int item = blockIdx.x; item < n_items && ok; item += gridDim.x) {
      const int slice = item / n_pairs, pair = item - slice * n_pairs;
      const int t0 = slice * tiles_per_slice, t1 = min(t0 + tiles_per_slice, n_tiles_total);
      const int qi = pair * 2 * MM_TILE + a * MM_TILE + row;
      const bool qvalid = qi < nq;
      const int my_pq = qvalid ? __ldg(pq + qi) : 0;
      // s = d - |q|; the bound is the query's global second-best distance so far (inclusive: an equal
      // distance at a lower row index can still be the second best), in s units
      int thr = qvalid ? (int)min(__ldcg(gthr + qi), 0xFFFu) + 1 - my_pq : -0x4000;
      unsigned b1 = 0xFFFFFFFFu, b2 = 0xFFFFFFFFu;  // (d << 20) | row within the slice
      for (int t = t0; t < t1; ++t, ++seq) {
        ok = mm_bar_wait(&sm.t_full[acc], acc_phase);
        if (!ok) break;
        mm_fence_after();
        const int j0 = t * MM_TILE;
        const unsigned tbase = tmem + ((quarter * 32u) << 16) + acc * 256u + a * 128u + half * 64u;
        // the chunk after the one being processed is already on its way out of TMEM
        int r0[32], r1[32];
        mm_tmem_ld32_issue(tbase, r0);
        auto process = [&](int (&r)[32], int c) {
          int mn = 0x7FFFFFFF;  // the accumulators ARE s = |m| - 2 q.m
#pragma unroll
          for (int i = 0; i < 32; i += 2) mn = min(mn, min(r[i], r[i + 1]));
          if (mn < thr) {  // rare: exact update in row order
            const unsigned jl = (unsigned)(j0 - t0 * MM_TILE + c * 32);
#pragma unroll
            for (int i = 0; i < 32; ++i) {
              if (r[i] < thr) {
                const unsigned key = ((unsigned)(r[i] + my_pq) << 20) + (jl + (unsigned)i);
                b2 = min(b2, max(b1, key));
                b1 = min(b1, key);
                if (b2 != 0xFFFFFFFFu) thr = min(thr, (int)(b2 >> 20) - my_pq);
              }
            }
          }
        };
        mm_tmem_ld_wait(r0);
        mm_tmem_ld32_issue(tbase + 32u, r1);
        process(r0, 2 * half);
        mm_tmem_ld_wait(r1);
        // the accumulator buffer is free as soon as its last chunk is in registers
        mm_fence_before();
        __syncwarp();
        if (lane == 0) mm_bar_arrive(&sm.t_empty[acc]);
        process(r1, 2 * half + 1);
        if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
      }
      if (!ok) break;
      if (qvalid) {
        const size_t o = (size_t)(slice * 2 + half) * nq + qi;
        const unsigned long long base = (unsigned long long)(index_base + (long long)t0 * MM_TILE);
        part_best[o] = b1 == 0xFFFFFFFFu ? ~0ull : ((unsigned long long)(b1 >> 20) << 32) | (base + (b1 & 0xFFFFFu));
        part_second[o] = b2 == 0xFFFFFFFFu ? ~0ull : ((unsigned long long)(b2 >> 20) << 32) | (base + (b2 & 0xFFFFFu));
        if (b2 != 0xFFFFFFFFu) atomicMin(gthr + qi, b2 >> 20);
      }
    }
  }
  if (!ok && lane == 0) atomicExch(err, 1u);
  // ---- teardown: every path reaches it, the TMEM is always released ----
  mm_fence_before();
  __syncthreads();
  if (warp == 1) {
    mm_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(512u) : "memory");
  }
}

// ---- host side ----
static int mm_encode_map(CUtensorMap* out, const void* base, int rows, int row_bytes, int box_bytes,
                         CUtensorMapSwizzle swizzle) {
  typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  static EncodeTiled fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<EncodeTiled>(f);
  }();
  if (!fn) return -1;
  const cuuint64_t dims[2] = {(cuuint64_t)row_bytes, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)row_bytes};
  const cuuint32_t box[2] = {(cuuint32_t)box_bytes, (cuuint32_t)MM_TILE};
  const cuuint32_t estr[2] = {1u, 1u};
  return fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS ? 0 : -1;
}

int hamming_mma_slices(int nq, int nm, int sms, int* tiles_per_slice) {
  const int n_pairs = (nq + 2 * MM_TILE - 1) / (2 * MM_TILE);
  const int n_tiles = (nm + MM_TILE - 1) / MM_TILE;
  // about 8 items per SM (load balance of the persistent CTAs), slices of at least 16 tiles, at most 2^20 rows
  static const int items = [] { const char* e = getenv("ORBFE_TC_ITEMS"); const int v = e ? atoi(e) : 0; return v > 0 ? v : 8; }();
  int slices = (items * sms + n_pairs - 1) / n_pairs;
  slices = std::max(1, std::min(slices, (n_tiles + 15) / 16));
  int tps = (n_tiles + slices - 1) / slices;
  tps = std::min(tps, (1 << 20) / MM_TILE);
  *tiles_per_slice = tps;
  return (n_tiles + tps - 1) / tps;
}

size_t hamming_mma_scratch_bytes(int nq, int nm, int slices) {
  const size_t nm_pad = ((size_t)nm + MM_TILE - 1) / MM_TILE * MM_TILE, nq_pad = ((size_t)nq + 255) / 256 * 256;
  return (size_t)nq * MM_K + (size_t)nm * MM_K + nq_pad * 32 + nm_pad * 32 + nq_pad * 4 + nq_pad * 4 +
         (size_t)4 * slices * nq * 8 + 16 * 256;
}

// scratch layout (all 256-byte aligned): Qb | Mb | Qx | Mx | pq | gthr | part_best | part_second | err
int launch_hamming_mma(const uint8_t* query, int nq, const uint8_t* map, int nm, long long index_base, uint8_t* scratch,
                       int slices, int tiles_per_slice, int sms, unsigned long long* best, unsigned long long* second,
                       unsigned** err_out, cudaStream_t s) {
  const size_t nm_pad = ((size_t)nm + MM_TILE - 1) / MM_TILE * MM_TILE, nq_pad = ((size_t)nq + 255) / 256 * 256;
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  uint8_t* p = scratch;
  uint8_t* qb = p; p += up((size_t)nq * MM_K);
  uint8_t* mb = p; p += up((size_t)nm * MM_K);
  uint8_t* qx = p; p += up(nq_pad * 32);
  uint8_t* mx = p; p += up(nm_pad * 32);
  int32_t* pq = reinterpret_cast<int32_t*>(p); p += up(nq_pad * 4);
  unsigned* gthr = reinterpret_cast<unsigned*>(p); p += up(nq_pad * 4);
  unsigned long long* pb = reinterpret_cast<unsigned long long*>(p); p += up((size_t)2 * slices * nq * 8);
  unsigned long long* ps = reinterpret_cast<unsigned long long*>(p); p += up((size_t)2 * slices * nq * 8);
  unsigned* err = reinterpret_cast<unsigned*>(p);
  *err_out = err;
  CUtensorMap qmap, mmap, qxmap, mxmap;
  if (mm_encode_map(&qmap, qb, nq, MM_K, 128, CU_TENSOR_MAP_SWIZZLE_128B) != 0 ||
      mm_encode_map(&mmap, mb, nm, MM_K, 128, CU_TENSOR_MAP_SWIZZLE_128B) != 0 ||
      mm_encode_map(&qxmap, qx, nq, 32, 32, CU_TENSOR_MAP_SWIZZLE_32B) != 0 ||
      mm_encode_map(&mxmap, mx, (int)nm_pad, 32, 32, CU_TENSOR_MAP_SWIZZLE_32B) != 0)
    return -1;
  const int smem = (int)sizeof(MmSmem) + 1024;
  // a per-device attribute (several GPUs may be driven from one process): set on every call, it costs microseconds
  if (cudaFuncSetAttribute(hamming_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -2;
  if (const char* e = getenv("ORBFE_TC_TIMEOUT_S")) {
    const double sec = atof(e);
    const unsigned long long ns = (unsigned long long)((sec > 0.001 ? sec : 0.001) * 1e9);
    cudaMemcpyToSymbolAsync(mm_timeout_ns, &ns, sizeof(ns), 0, cudaMemcpyHostToDevice, s);
  }
  cudaMemsetAsync(err, 0, 4, s);
  expand_bits_kernel<true><<<(int)((nq_pad + 7) / 8), 256, 0, s>>>(query, nq, (int)nq_pad, qb, qx, pq);
  expand_bits_kernel<false><<<(int)((nm_pad + 7) / 8), 256, 0, s>>>(map, nm, (int)nm_pad, mb, mx, nullptr);
  fill_u32_kernel<<<(int)((nq_pad + 255) / 256), 256, 0, s>>>(gthr, (int)nq_pad, 0xFFFu);
  const int n_pairs = (nq + 2 * MM_TILE - 1) / (2 * MM_TILE);
  const int grid = std::min(sms, n_pairs * slices);
  hamming_mma_kernel<<<grid, MM_THREADS, smem, s>>>(qmap, mmap, qxmap, mxmap, nq, nm, n_pairs, slices, tiles_per_slice,
                                                    index_base, pq, gthr, pb, ps, err);
  launch_hamming_merge(pb, ps, 2 * slices, nq, best, second, s);  // two column halves per slice
  return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // namespace orbfe
