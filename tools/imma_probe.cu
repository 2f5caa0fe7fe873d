// Throughput probe: legacy mma.sync.m16n8k32 u8 x u8 -> s32 on sm_100a (register operands only).
// Measured on B200 (round 2): 138.9 G mma/s = 1.14 Pop/s, one m16n8k32 per 8.4 clk per SM sub-partition — a quarter of the
// tcgen05 kind::i8 rate, which is why the tensor-core brute force (csrc/k_match_mma.cu) is written with tcgen05 / TMEM.
// (mma.sync m16n8k256 b1 and.popc still compiles for sm_100a but is emulated: 8 IMMA.16832 + ~140 unpack instructions.)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o imma_probe imma_probe.cu && ./imma_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) k(int iters, int* out) {
  uint32_t a[4] = {threadIdx.x, threadIdx.x * 3u, threadIdx.x * 5u, threadIdx.x * 7u};
  uint32_t b[2] = {threadIdx.x * 11u, threadIdx.x * 13u};
  int c[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                   : "+r"(c[i][0]), "+r"(c[i][1]), "+r"(c[i][2]), "+r"(c[i][3])
                   : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0] + i), "r"(b[1]));
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
  if (s == 0x7fffffff) out[0] = s;
}
int main() {
  int* d; cudaMalloc(&d, 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 4096, blocks = 148 * 8;
  k<<<blocks, 256>>>(16, d);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(e0); k<<<blocks, 256>>>(iters, d); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double mmas = (double)blocks * 8 /*warps*/ * iters * 8;
    printf("m16n8k32 u8: %.3f ms, %.2f G mma/s, %.1f TOPS (2*16*8*32 per mma), %.2f clk per mma per SM sub-partition @1.965 GHz\n", ms,
           mmas / ms / 1e6, mmas * 2 * 16 * 8 * 32 / ms / 1e9, 592.0 * 1.965e9 / (mmas / (ms * 1e-3)));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
