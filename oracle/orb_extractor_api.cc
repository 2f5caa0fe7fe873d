// orb_extractor_api.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h).
// ORBextractor::operator() (reference: src/ORBextractor.cc:1065-1127) and the C API around it.
#include <algorithm>
#include <cstring>

#include "../include/orb_math_shared.h"
#include "orb_internal.h"
#include "orb_oracle.h"

namespace orc {
void build_info_public(const orbfe_extractor_params& p, orbfe_extractor_info* info);
void compute_pyramid_public(orc_extractor* e, const uint8_t* image, int width, int height, int stride);
void detect_cells_public(const Image& im, int iniTh, int minTh, std::vector<FastPoint>& out);
}  // namespace orc

extern "C" {

orc_extractor* orc_extractor_create(const orbfe_extractor_params* p, int flags) {
  if (!p || p->nlevels < 1 || p->nlevels > ORBFE_MAX_LEVELS || p->nfeatures < 1) return nullptr;
  orc_extractor* e = new orc_extractor();
  e->params = *p;
  e->flags = flags;
  orc::build_info_public(*p, &e->info);
  e->pattern.assign(orc::pattern_table(), orc::pattern_table() + 1024);  // :448-450
  return e;
}

void orc_extractor_destroy(orc_extractor* e) { delete e; }

void orc_extractor_info(const orc_extractor* e, orbfe_extractor_info* info) { *info = e->info; }

int orc_extract(orc_extractor* e, const uint8_t* image, int width, int height, int stride, orbfe_keypoint* kps,
                uint8_t* desc, int capacity) {
  if (!image || width <= 0 || height <= 0) return 0;  // empty image: silent return (:1068)
  const int L = e->info.nlevels;
  const bool libm = (e->flags & ORC_FLAG_LIBM_SINCOS) != 0;
  orc::compute_pyramid_public(e, image, width, height, stride);  // :1075
  e->candidates.assign(L, {});
  e->level_keypoints.assign(L, {});
  e->blurred.assign(L, orc::Image());
  // ComputeKeyPointsOctTree (:776-875)
  std::vector<int> xs, ys, sc, kept;
  for (int lv = 0; lv < L; ++lv) {
    const orc::Image& im = e->pyramid[lv];
    std::vector<orc::FastPoint>& cand = e->candidates[lv];
    orc::detect_cells_public(im, e->params.ini_th_fast, e->params.min_th_fast, cand);
    const int minBX = 16, minBY = 16, maxBX = im.w - 16, maxBY = im.h - 16;
    xs.resize(cand.size()); ys.resize(cand.size()); sc.resize(cand.size());
    for (size_t i = 0; i < cand.size(); ++i) { xs[i] = cand[i].x; ys[i] = cand[i].y; sc[i] = cand[i].score; }
    int r = 0;
    if (!cand.empty())
      r = orc::distribute_octtree(xs.data(), ys.data(), sc.data(), (int)cand.size(), minBX, maxBX, minBY, maxBY,
                                  e->info.features_per_level[lv], e->flags, kept);
    else kept.clear();
    if (r < 0) return -2;
    const int scaledPatch = (int)(31 * e->info.scale_factors[lv]);  // :857 (int = PATCH_SIZE*float)
    std::vector<orbfe_keypoint>& out = e->level_keypoints[lv];
    out.resize(kept.size());
    for (size_t i = 0; i < kept.size(); ++i) {
      const orc::FastPoint& c = cand[kept[i]];
      orbfe_keypoint k;
      k.x = (float)c.x + (float)minBX;  // :865-866
      k.y = (float)c.y + (float)minBY;
      k.size = (float)scaledPatch;
      k.angle = -1.f;
      k.response = (float)c.score;
      k.octave = lv;
      k.class_id = -1;
      out[i] = k;
    }
  }
  // computeOrientation on the un-blurred levels (:873-874, :472-479)
  for (int lv = 0; lv < L; ++lv) {
    const orc::Image& im = e->pyramid[lv];
    for (orbfe_keypoint& k : e->level_keypoints[lv]) {
      const int cx = orc::round_half_even(k.x), cy = orc::round_half_even(k.y);
      k.angle = orc::ic_angle(im.row(cy) + cx, im.w, e->info.umax);
    }
  }
  // per level: blur, describe, rescale, append (:1096-1126)
  int n = 0;
  for (int lv = 0; lv < L; ++lv) {
    std::vector<orbfe_keypoint>& lk = e->level_keypoints[lv];
    if (lk.empty()) continue;
    const orc::Image& im = e->pyramid[lv];
    orc::Image& bl = e->blurred[lv];
    bl.w = im.w; bl.h = im.h; bl.px.resize(im.px.size());
    orc::gaussian_blur7_u8(im.px.data(), im.w, im.h, im.w, bl.px.data(), bl.w);  // :1107-1108
    const float scale = e->info.scale_factors[lv];
    for (const orbfe_keypoint& k0 : lk) {
      if (n < capacity) {
        const int cx = orc::round_half_even(k0.x), cy = orc::round_half_even(k0.y);
        orc::orb_descriptor(bl.row(cy) + cx, bl.w, k0.angle, libm, e->pattern.data(), desc + (size_t)n * 32);
        orbfe_keypoint k = k0;
        if (lv != 0) { k.x *= scale; k.y *= scale; }  // :1117-1123
        kps[n] = k;
      }
      ++n;
    }
  }
  return n;
}

int orc_extractor_level_size(const orc_extractor* e, int level, int* w, int* h) {
  if (level < 0 || level >= (int)e->pyramid.size()) return -1;
  *w = e->pyramid[level].w;
  *h = e->pyramid[level].h;
  return 0;
}

const uint8_t* orc_extractor_level_ptr(const orc_extractor* e, int level, int* stride) {
  if (level < 0 || level >= (int)e->pyramid.size()) return nullptr;
  *stride = e->pyramid[level].w;
  return e->pyramid[level].px.data();
}

const uint8_t* orc_extractor_blurred_ptr(const orc_extractor* e, int level, int* stride) {
  if (level < 0 || level >= (int)e->blurred.size() || e->blurred[level].px.empty()) return nullptr;
  *stride = e->blurred[level].w;
  return e->blurred[level].px.data();
}

int orc_extractor_candidates(const orc_extractor* e, int level, int cap, int32_t* xs, int32_t* ys, int32_t* scores) {
  if (level < 0 || level >= (int)e->candidates.size()) return -1;
  const auto& c = e->candidates[level];
  for (size_t i = 0; i < c.size() && (int)i < cap; ++i) { xs[i] = c[i].x; ys[i] = c[i].y; scores[i] = c[i].score; }
  return (int)c.size();
}

int orc_extractor_level_keypoints(const orc_extractor* e, int level, int cap, orbfe_keypoint* kps) {
  if (level < 0 || level >= (int)e->level_keypoints.size()) return -1;
  const auto& c = e->level_keypoints[level];
  for (size_t i = 0; i < c.size() && (int)i < cap; ++i) kps[i] = c[i];
  return (int)c.size();
}

int orc_distribute_octtree(const int32_t* xs, const int32_t* ys, const int32_t* scores, int n, int minX, int maxX,
                           int minY, int maxY, int N, int flags, int cap, int32_t* kept_index) {
  std::vector<int> kept;
  const int r = orc::distribute_octtree(xs, ys, scores, n, minX, maxX, minY, maxY, N, flags, kept);
  if (r < 0) return r;
  for (int i = 0; i < r && i < cap; ++i) kept_index[i] = kept[i];
  return r;
}

void orc_sincos_deg(float angle_deg, int use_libm, float* c, float* s) { orc::sincos_deg(angle_deg, use_libm != 0, c, s); }

float orc_ic_angle(const uint8_t* img, int stride, int x, int y) {
  orbfe_extractor_params p = {1000, 1.2f, 8, 20, 7};
  orbfe_extractor_info info;
  orc::build_info_public(p, &info);
  return orc::ic_angle(img + (size_t)y * stride + x, stride, info.umax);
}

void orc_orb_descriptor(const uint8_t* blurred, int stride, int x, int y, float angle_deg, int use_libm,
                        uint8_t* desc32) {
  orc::orb_descriptor(blurred + (size_t)y * stride + x, stride, angle_deg, use_libm != 0, orc::pattern_table(), desc32);
}

}  // extern "C"
