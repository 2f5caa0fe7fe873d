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
int main(){ int n=2000; std::vector<uint8_t> h((size_t)n*RS*7); for(int t=0;t<n;t++){int base=rand()%256, amp=rand()%100+1; for(int i=0;i<RS*7;i++) h[(size_t)t*RS*7+i]=(uint8_t)std::min(255,std::max(0,base+rand()%amp-amp/2));}
 const uint8_t* p = h.data()+3*RS+3; printf("center %d\n", p[0]); for(int k=0;k<16;k++) printf("%d ", p[0]-p[kRingDy[k]*RS+kRingDx[k]]); printf("\nref %d\n", ref(p,RS)); }
