// orb_internal.h — internal declarations of the CPU ORACLE (test infrastructure only).
#ifndef ORB_INTERNAL_H_
#define ORB_INTERNAL_H_

#include <cmath>
#include <cstdint>
#include <vector>

#include "../include/orbfe_types.h"

namespace orc {

// cvRound(): round half to even (SSE cvtss2si / lrint in the default rounding mode).
static inline int round_half_even(float v) { return (int)std::nearbyintf(v); }
static inline int round_half_even(double v) { return (int)std::nearbyint(v); }

struct FastPoint {
  int x, y, score;
};

void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
void cvt_gray_u8(const uint8_t* src, int w, int h, int sstride, int channels, bool rgb, uint8_t* dst, int dstride);
void remap_linear_u8(const uint8_t* src, int sw, int sh, int sstride, const float* mapx, const float* mapy, int dw, int dh,
                     uint8_t* dst, int dstride);
int fast_corner_strength(const uint8_t* p, int stride);
int fast9_16(const uint8_t* roi, int cols, int rows, int stride, int threshold, bool nms, std::vector<FastPoint>& out);

// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1715-1731): sum over 8 x u32 of popcount(a^b).
// The reference uses a SWAR bit-hack; popcount of the xor is the same integer.
static inline int descriptor_distance(const uint8_t* a, const uint8_t* b) {
  const uint64_t* pa = reinterpret_cast<const uint64_t*>(a);
  const uint64_t* pb = reinterpret_cast<const uint64_t*>(b);
  uint64_t w0, w1, w2, w3, v0, v1, v2, v3;
  __builtin_memcpy(&w0, pa + 0, 8); __builtin_memcpy(&v0, pb + 0, 8);
  __builtin_memcpy(&w1, pa + 1, 8); __builtin_memcpy(&v1, pb + 1, 8);
  __builtin_memcpy(&w2, pa + 2, 8); __builtin_memcpy(&v2, pb + 2, 8);
  __builtin_memcpy(&w3, pa + 3, 8); __builtin_memcpy(&v3, pb + 3, 8);
  return __builtin_popcountll(w0 ^ v0) + __builtin_popcountll(w1 ^ v1) + __builtin_popcountll(w2 ^ v2) +
         __builtin_popcountll(w3 ^ v3);
}

// the literal SWAR form of the reference, used by the KAT test to show both agree
int descriptor_distance_swar(const uint8_t* a, const uint8_t* b);

int distribute_octtree(const int* xs, const int* ys, const int* scores, int n, int minX, int maxX, int minY, int maxY,
                       int N, int flags, std::vector<int>& kept);
float ic_angle(const uint8_t* center, int step, const int* umax);
void sincos_deg(float angle_deg, bool libm, float* c, float* s);
void orb_descriptor(const uint8_t* center, int step, float angle_deg, bool libm, const int* pattern, uint8_t* desc);
const int* pattern_table();

struct Image {
  int w = 0, h = 0;
  std::vector<uint8_t> px;  // tightly packed, stride == w
  const uint8_t* row(int y) const { return px.data() + (size_t)y * w; }
  uint8_t* row(int y) { return px.data() + (size_t)y * w; }
};

}  // namespace orc

// opaque extractor object (orb_extractor.cc)
struct orc_extractor {
  orbfe_extractor_params params;
  int flags = 0;
  orbfe_extractor_info info;
  std::vector<int> pattern;  // 1024 ints
  std::vector<orc::Image> pyramid;
  std::vector<orc::Image> blurred;
  std::vector<std::vector<orc::FastPoint>> candidates;     // per level, relative to minBorder
  std::vector<std::vector<orbfe_keypoint>> level_keypoints;  // per level, level coords
};

#endif
