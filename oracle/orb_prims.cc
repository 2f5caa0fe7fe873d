// orb_prims.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h).
//
// Integer restatements of the OpenCV primitives the reference's front end
// calls.  The reference does not contain this arithmetic (OpenCV is an
// un-vendored, un-pinned system dependency: CMakeLists.txt:32-38); semantics are
// pinned to OpenCV 4.13.0 and verified bit-for-bit against python cv2 4.13.0 in
// tests/test_oracle_cv2.py.  Call sites in the reference:
//   resize        src/ORBextractor.cc:1142
//   GaussianBlur  src/ORBextractor.cc:1108
//   FAST          src/ORBextractor.cc:830,835
//   fastAtan2     src/ORBextractor.cc:103
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../include/orb_math_shared.h"
#include "orb_internal.h"
#include "orb_oracle.h"

namespace orc {

// ---------------------------------------------------------------------------
// cv::resize(..., INTER_LINEAR) for CV_8UC1
// ---------------------------------------------------------------------------
void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
  const double scale_x = (double)sw / dw;
  const double scale_y = (double)sh / dh;
  std::vector<int> xofs(dw);
  std::vector<short> xa0(dw), xa1(dw);
  for (int x = 0; x < dw; ++x) {
    float fx = (float)((x + 0.5) * scale_x - 0.5);
    int sx = (int)std::floor(fx);
    fx -= sx;
    if (sx < 0) { sx = 0; fx = 0.f; }
    if (sx >= sw - 1) { sx = sw - 1; fx = 0.f; }
    xofs[x] = sx;
    xa0[x] = (short)round_half_even((1.f - fx) * 2048.f);
    xa1[x] = (short)round_half_even(fx * 2048.f);
  }
  std::vector<int> row0(dw), row1(dw);
  int cached0 = -1, cached1 = -1;
  auto hpass = [&](int sy, std::vector<int>& out) {
    const uint8_t* S = src + (size_t)sy * sstride;
    for (int x = 0; x < dw; ++x) {
      const int sx = xofs[x];
      const int sx1 = std::min(sx + 1, sw - 1);
      out[x] = S[sx] * xa0[x] + S[sx1] * xa1[x];
    }
  };
  for (int y = 0; y < dh; ++y) {
    float fy = (float)((y + 0.5) * scale_y - 0.5);
    int sy = (int)std::floor(fy);
    fy -= sy;
    const int b0 = (short)round_half_even((1.f - fy) * 2048.f);
    const int b1 = (short)round_half_even(fy * 2048.f);
    const int y0 = std::min(std::max(sy, 0), sh - 1);
    const int y1 = std::min(std::max(sy + 1, 0), sh - 1);
    // reuse horizontally filtered rows where possible (rows advance monotonically)
    if (cached1 == y0) { std::swap(row0, row1); std::swap(cached0, cached1); }
    if (cached0 != y0) { hpass(y0, row0); cached0 = y0; }
    if (cached1 != y1) {
      if (y1 == y0) { row1 = row0; } else { hpass(y1, row1); }
      cached1 = y1;
    }
    uint8_t* D = dst + (size_t)y * dstride;
    for (int x = 0; x < dw; ++x) {
      const int v = (((b0 * (row0[x] >> 4)) >> 16) + ((b1 * (row1[x] >> 4)) >> 16) + 2) >> 2;
      D[x] = (uint8_t)v;
    }
  }
}

// ---------------------------------------------------------------------------
// cv::GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) for CV_8UC1
// 8.8 fixed point kernel, 16.16 accumulate.
// ---------------------------------------------------------------------------
static inline int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) {
    if (p < 0) p = -p;
    else p = 2 * (len - 1) - p;
  }
  return p;
}

void gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
  static const int K[7] = {18, 34, 48, 56, 48, 34, 18};
  std::vector<uint16_t> hbuf((size_t)w * h);
  for (int y = 0; y < h; ++y) {
    const uint8_t* S = src + (size_t)y * sstride;
    uint16_t* H = hbuf.data() + (size_t)y * w;
    for (int x = 0; x < w; ++x) {
      int acc = 0;
      if (x >= 3 && x < w - 3) {
        for (int k = 0; k < 7; ++k) acc += K[k] * S[x + k - 3];
      } else {
        for (int k = 0; k < 7; ++k) acc += K[k] * S[reflect101(x + k - 3, w)];
      }
      H[x] = (uint16_t)acc;
    }
  }
  for (int y = 0; y < h; ++y) {
    const uint16_t* R[7];
    for (int k = 0; k < 7; ++k) R[k] = hbuf.data() + (size_t)reflect101(y + k - 3, h) * w;
    uint8_t* D = dst + (size_t)y * dstride;
    for (int x = 0; x < w; ++x) {
      uint32_t acc = 0;
      for (int k = 0; k < 7; ++k) acc += (uint32_t)K[k] * R[k][x];
      D[x] = (uint8_t)((acc + 32768u) >> 16);
    }
  }
}

// ---------------------------------------------------------------------------
// cv::FAST(roi, kps, threshold, nonmaxSuppression) TYPE_9_16
// ---------------------------------------------------------------------------
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// Largest t such that the pixel is still a FAST-9 corner at threshold t, plus one
// ( = max over 9-arcs of the min |difference| with a common sign ); 0 if none.
int fast_corner_strength(const uint8_t* p, int stride) {
  const int v = p[0];
  int d[25];
  for (int k = 0; k < 16; ++k) d[k] = v - p[kRingDy[k] * stride + kRingDx[k]];
  for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
  int best = 0;
  for (int s = 0; s < 16; ++s) {
    int mn = d[s], mx = d[s];
    for (int k = 1; k < 9; ++k) {
      mn = std::min(mn, d[s + k]);
      mx = std::max(mx, d[s + k]);
    }
    best = std::max(best, std::max(mn, -mx));
  }
  return best;
}

static inline bool has_run9(uint32_t m16) {
  uint32_t x = m16 | (m16 << 16);
  uint32_t r = x & (x >> 1);
  r &= r >> 2;
  r &= r >> 4;
  r &= x >> 8;
  return r != 0;
}

int fast9_16(const uint8_t* roi, int cols, int rows, int stride, int threshold, bool nms, std::vector<FastPoint>& out) {
  out.clear();
  if (cols < 7 || rows < 7) return 0;
  threshold = std::min(std::max(threshold, 0), 255);
  // score map (0 where not a corner / outside the detection area)
  std::vector<uint8_t> score((size_t)cols * rows, 0);
  int off[16];
  for (int k = 0; k < 16; ++k) off[k] = kRingDy[k] * stride + kRingDx[k];
  for (int y = 3; y < rows - 3; ++y) {
    const uint8_t* row = roi + (size_t)y * stride;
    uint8_t* srow = score.data() + (size_t)y * cols;
    for (int x = 3; x < cols - 3; ++x) {
      const uint8_t* p = row + x;
      const int v = p[0];
      const int hi = v + threshold, lo = v - threshold;
      // quick reject on the vertical pair (ring 0 and 8), as any 9-arc contains one of them
      const int a = p[off[0]], b = p[off[8]];
      if (!((a > hi) || (a < lo) || (b > hi) || (b < lo))) continue;
      uint32_t bright = 0, dark = 0;
      for (int k = 0; k < 16; ++k) {
        const int q = p[off[k]];
        bright |= (uint32_t)(q > hi) << k;
        dark |= (uint32_t)(q < lo) << k;
      }
      if (has_run9(bright) || has_run9(dark)) {
        const int m = fast_corner_strength(p, stride);
        srow[x] = (uint8_t)(m - 1);
        if (!nms) out.push_back({x, y, m - 1});
      }
    }
  }
  if (!nms) return (int)out.size();
  for (int y = 3; y < rows - 3; ++y) {
    const uint8_t* s = score.data() + (size_t)y * cols;
    for (int x = 3; x < cols - 3; ++x) {
      const int c = s[x];
      if (c == 0) {
        // a corner always has score >= threshold >= 0; score 0 only happens for threshold 0
        if (threshold > 0) continue;
      }
      // strict maximum over the 8 neighbours (scores outside the detection area are 0)
      // NB: a corner at threshold 0 with score 0 can never be a strict maximum.
      if (c > s[x - 1] && c > s[x + 1] && c > s[x - cols - 1] && c > s[x - cols] && c > s[x - cols + 1] &&
          c > s[x + cols - 1] && c > s[x + cols] && c > s[x + cols + 1]) {
        // need to know it is a corner: score>0 implies corner (score = m-1 with m>threshold>=0)
        out.push_back({x, y, c});
      }
    }
  }
  return (int)out.size();
}


// ---- input side of the pipeline ----
// cv::cvtColor(src, dst, CV_{RGB,BGR,RGBA,BGRA}2GRAY) on 8-bit pixels (src/Tracking.cc:189-214): OpenCV 4.x fixed
// point, 15 fractional bits: (R*9798 + G*19235 + B*3735 + 2^14) >> 15 (verified against cv2 4.13 on all (R,G) and
// every third B).
void cvt_gray_u8(const uint8_t* src, int w, int h, int sstride, int channels, bool rgb, uint8_t* dst, int dstride) {
  for (int y = 0; y < h; ++y) {
    const uint8_t* s = src + (size_t)y * sstride;
    uint8_t* d = dst + (size_t)y * dstride;
    for (int x = 0; x < w; ++x, s += channels) {
      const int r = rgb ? s[0] : s[2], g = s[1], b = rgb ? s[2] : s[0];
      d[x] = (uint8_t)((r * 9798 + g * 19235 + b * 3735 + (1 << 14)) >> 15);
    }
  }
}

// cv::remap(src, dst, map1, map2, INTER_LINEAR) with CV_32FC1 maps, BORDER_CONSTANT 0, 8-bit single channel
// (Examples/Stereo/stereo_euroc.cc:136-137): coordinates go to 1/32 pixel with cvRound(map * 32) (fp32 product, round
// half to even), the four taps are weighted with the 15-bit integer table (32-ay)(32-ax)*32 ... and the sum is rounded
// with + 2^14 >> 15; taps outside the source read 0.
void remap_linear_u8(const uint8_t* src, int sw, int sh, int sstride, const float* mapx, const float* mapy, int dw, int dh,
                     uint8_t* dst, int dstride) {
  for (int y = 0; y < dh; ++y)
    for (int x = 0; x < dw; ++x) {
      const int sx = round_half_even(mapx[(size_t)y * dw + x] * 32.0f);
      const int sy = round_half_even(mapy[(size_t)y * dw + x] * 32.0f);
      int ix = sx >> 5, iy = sy >> 5;
      ix = ix < -32768 ? -32768 : (ix > 32767 ? 32767 : ix);  // saturate_cast<short>
      iy = iy < -32768 ? -32768 : (iy > 32767 ? 32767 : iy);
      const int ax = sx & 31, ay = sy & 31;
      auto px = [&](int yy, int xx) -> int {
        return (xx >= 0 && xx < sw && yy >= 0 && yy < sh) ? src[(size_t)yy * sstride + xx] : 0;
      };
      const int v = px(iy, ix) * ((32 - ay) * (32 - ax) * 32) + px(iy, ix + 1) * ((32 - ay) * ax * 32) +
                    px(iy + 1, ix) * (ay * (32 - ax) * 32) + px(iy + 1, ix + 1) * (ay * ax * 32);
      dst[(size_t)y * dstride + x] = (uint8_t)((v + (1 << 14)) >> 15);
    }
}

}  // namespace orc

// ---------------------------------------------------------------------------
// C API
// ---------------------------------------------------------------------------
extern "C" {

void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
  orc::resize_linear_u8(src, sw, sh, sstride, dst, dw, dh, dstride);
}

void orc_gaussian_blur7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
  orc::gaussian_blur7_u8(src, w, h, sstride, dst, dstride);
}

int orc_fast9_16(const uint8_t* roi, int cols, int rows, int stride, int threshold, int nms, int cap, int32_t* xs,
                 int32_t* ys, int32_t* scores) {
  std::vector<orc::FastPoint> pts;
  orc::fast9_16(roi, cols, rows, stride, threshold, nms != 0, pts);
  const int n = (int)pts.size();
  for (int i = 0; i < n && i < cap; ++i) {
    xs[i] = pts[i].x;
    ys[i] = pts[i].y;
    scores[i] = pts[i].score;
  }
  return n;
}

float orc_fast_atan2(float y, float x) { return orb_fast_atan2_deg(y, x); }

int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) { return orc::descriptor_distance(a, b); }

void orc_cvt_gray_u8(const uint8_t* src, int w, int h, int sstride, int channels, int rgb, uint8_t* dst, int dstride) {
  orc::cvt_gray_u8(src, w, h, sstride, channels, rgb != 0, dst, dstride);
}

void orc_remap_linear_u8(const uint8_t* src, int sw, int sh, int sstride, const float* mapx, const float* mapy, int dw,
                         int dh, uint8_t* dst, int dstride) {
  orc::remap_linear_u8(src, sw, sh, sstride, mapx, mapy, dw, dh, dst, dstride);
}

}  // extern "C"
