#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <algorithm>
#include <vector>
constexpr int RS=68;
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
int ref(const uint8_t* p, int stride) {
  const int v = p[0]; int d[25];
  for (int k = 0; k < 16; ++k) d[k] = v - p[kRingDy[k] * stride + kRingDx[k]];
  for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
  int best = 0;
  for (int s = 0; s < 16; ++s) { int mn = d[s], mx = d[s]; for (int k = 1; k < 9; ++k) { mn = std::min(mn, d[s + k]); mx = std::max(mx, d[s + k]); } best = std::max(best, std::max(mn, -mx)); }
  return best;
}
__device__ __forceinline__ int corner_strength(const uint8_t* p) {
  const int v = p[0];
  int d[16];
  d[0] = v - p[3 * RS];       d[1] = v - p[3 * RS + 1];   d[2] = v - p[2 * RS + 2];   d[3] = v - p[RS + 3];
  d[4] = v - p[3];            d[5] = v - p[-RS + 3];      d[6] = v - p[-2 * RS + 2];  d[7] = v - p[-3 * RS + 1];
  d[8] = v - p[-3 * RS];      d[9] = v - p[-3 * RS - 1];  d[10] = v - p[-2 * RS - 2]; d[11] = v - p[-RS - 3];
  d[12] = v - p[-3];          d[13] = v - p[RS - 3];      d[14] = v - p[2 * RS - 2];  d[15] = v - p[3 * RS - 1];
  // min over every 9-arc of d (darker ring) and of e = -d (brighter ring).  NB: written with two min chains on
  // purpose: ptxas 12.9 drops the negation when max(a, -b) is fused into VIMNMX3 on sm_100a (measured).
  int e[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) e[i] = -d[i];
  int a1[16], b1[16], a2[16], b2[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { a1[i] = min(d[i], d[(i + 1) & 15]); b1[i] = min(e[i], e[(i + 1) & 15]); }
#pragma unroll
  for (int i = 0; i < 16; ++i) { a2[i] = min(a1[i], a1[(i + 2) & 15]); b2[i] = min(b1[i], b1[(i + 2) & 15]); }
#pragma unroll
  for (int i = 0; i < 16; ++i) { a1[i] = min(a2[i], a2[(i + 4) & 15]); b1[i] = min(b2[i], b2[(i + 4) & 15]); }
  int best = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    best = max(best, min(a1[i], d[(i + 8) & 15]));
    best = max(best, min(b1[i], e[(i + 8) & 15]));
  }
  return best;
}
__device__ int naive(const uint8_t* p) {
  const int dx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
  const int dy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};
  const int v = p[0]; int d[25];
  for (int k = 0; k < 16; ++k) d[k] = v - p[dy[k] * RS + dx[k]];
  for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
  int best = 0;
  for (int s = 0; s < 16; ++s) { int mn = d[s], mx = d[s]; for (int k = 1; k < 9; ++k) { mn = min(mn, d[s + k]); mx = max(mx, d[s + k]); } best = max(best, max(mn, -mx)); if (blockIdx.x==0 && threadIdx.x==0) printf("s %d d %d mn %d mx %d best %d\n", s, d[s], mn, mx, best); }
  return best;
}
__global__ void k(const uint8_t* img, int n, int* out, int* out2) {
  __shared__ uint8_t roi[RS*7];
  int t = blockIdx.x;
  for (int i = threadIdx.x; i < RS*7; i += blockDim.x) roi[i] = img[(size_t)t*RS*7+i];
  __syncthreads();
  if (t == 0 && threadIdx.x == 0) { printf("dev center %d ring0 %d ring4 %d row0:", roi[3*RS+3], roi[6*RS+3], roi[3*RS+6]); for (int i=0;i<8;i++) printf(" %d", roi[i]); printf("\n"); }
  if (threadIdx.x < 60) { out[t*60+threadIdx.x] = corner_strength(roi + 3*RS + 3 + threadIdx.x); out2[t*60+threadIdx.x] = naive(roi + 3*RS + 3 + threadIdx.x); }
}
int main(){ int n=2000; std::vector<uint8_t> h((size_t)n*RS*7); for(int t=0;t<n;t++){int base=rand()%256, amp=rand()%100+1; for(int i=0;i<RS*7;i++) h[(size_t)t*RS*7+i]=(uint8_t)std::min(255,std::max(0,base+rand()%amp-amp/2));}
  printf("host center %d ring0 %d ring4 %d row0:", h[3*RS+3], h[6*RS+3], h[3*RS+6]); for (int i=0;i<8;i++) printf(" %d", h[i]); printf("\n");
  uint8_t* d; int* o; cudaMalloc(&d,h.size()); cudaMalloc(&o,n*60*4); cudaMemcpy(d,h.data(),h.size(),cudaMemcpyHostToDevice);
  int* o2; cudaMalloc(&o2,n*60*4); k<<<n,64>>>(d,n,o,o2); std::vector<int> r2(n*60); cudaMemcpy(r2.data(),o2,n*60*4,cudaMemcpyDeviceToHost); int bad2=0;  std::vector<int> r(n*60); cudaMemcpy(r.data(),o,n*60*4,cudaMemcpyDeviceToHost);
  int bad=0; for(int t=0;t<n;t++) for(int x=0;x<60;x++){ int a=ref(h.data()+(size_t)t*RS*7+3*RS+3+x,RS); if(a!=r2[t*60+x]) bad2++; if(a!=r[t*60+x]){ if(bad<5) printf("t %d x %d ref %d gpu %d\n",t,x,a,r[t*60+x]); bad++; } }
  printf("naive-bad %d ", bad2); printf("bad %d of %d err %s\n",bad,n*60,cudaGetErrorString(cudaGetLastError())); }
