// orb_extractor.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h).
//
// Restatement of ORBextractor (reference: src/ORBextractor.cc) on plain arrays.
// Each function cites the reference lines it follows.  Structure is kept
// faithful where it costs time in the reference (per-cell FAST calls, a
// std::list based quadtree, a full-level blur) because this code is also the
// timed CPU baseline.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <list>
#include <utility>

#include "../include/orb_math_shared.h"
#include "orb_internal.h"
#include "orb_oracle.h"

namespace orc {

static const int kPatch = 31;      // PATCH_SIZE        (src/ORBextractor.cc:72)
static const int kHalfPatch = 15;  // HALF_PATCH_SIZE   (:73)
static const int kEdge = 19;       // EDGE_THRESHOLD    (:74)

static const int kPattern[1024] = {
#include "../include/orb_pattern_31.inc"
};

// ---- constructor tables (src/ORBextractor.cc:410-470) ----------------------
static void build_info(const orbfe_extractor_params& p, orbfe_extractor_info* info) {
  std::memset(info, 0, sizeof(*info));
  const int L = p.nlevels;
  const double sf = (double)p.scale_factor;  // member `double scaleFactor` (include/ORBextractor.h:95)
  info->nlevels = L;
  info->scale_factor = p.scale_factor;
  info->scale_factors[0] = 1.0f;
  info->level_sigma2[0] = 1.0f;
  for (int i = 1; i < L; ++i) {
    info->scale_factors[i] = (float)((double)info->scale_factors[i - 1] * sf);  // :421
    info->level_sigma2[i] = info->scale_factors[i] * info->scale_factors[i];    // :422
  }
  for (int i = 0; i < L; ++i) {
    info->inv_scale_factors[i] = 1.0f / info->scale_factors[i];  // :429
    info->inv_level_sigma2[i] = 1.0f / info->level_sigma2[i];    // :430
  }
  // per-level quota (:435-446)
  const float factor = (float)(1.0 / sf);
  float desired = (float)p.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)L));
  int sum = 0;
  for (int lv = 0; lv < L - 1; ++lv) {
    info->features_per_level[lv] = round_half_even(desired);
    sum += info->features_per_level[lv];
    desired *= factor;
  }
  info->features_per_level[L - 1] = std::max(p.nfeatures - sum, 0);
  // umax: end of each row of the radius-15 disc (:454-469)
  int* umax = info->umax;
  const int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
  const int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
  const double hp2 = (double)kHalfPatch * kHalfPatch;
  for (int v = 0; v <= vmax; ++v) umax[v] = round_half_even(std::sqrt(hp2 - (double)v * v));
  for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
    while (umax[v0] == umax[v0 + 1]) ++v0;
    umax[v] = v0;
    ++v0;
  }
  int cap = 0;
  for (int lv = 0; lv < L; ++lv) cap += info->features_per_level[lv] + 3;
  info->capacity = cap + 64;
}

// ---- ComputePyramid (src/ORBextractor.cc:1129-1154) ------------------------
// The 19 px reflect-101 frame produced by copyMakeBorder is not materialised:
// nothing on the path reads it.
static void compute_pyramid(orc_extractor* e, const uint8_t* image, int width, int height, int stride) {
  const int L = e->info.nlevels;
  e->pyramid.resize(L);
  for (int lv = 0; lv < L; ++lv) {
    const float s = e->info.inv_scale_factors[lv];
    Image& im = e->pyramid[lv];
    im.w = round_half_even((float)width * s);   // :1134
    im.h = round_half_even((float)height * s);
    im.px.resize((size_t)im.w * im.h);
    if (lv == 0) {
      for (int y = 0; y < height; ++y) std::memcpy(im.row(y), image + (size_t)y * stride, width);
    } else {
      const Image& prev = e->pyramid[lv - 1];
      resize_linear_u8(prev.px.data(), prev.w, prev.h, prev.w, im.px.data(), im.w, im.h, im.w);  // :1142
    }
  }
}

// ---- cell loop of ComputeKeyPointsOctTree (src/ORBextractor.cc:776-850) ----
static void detect_cells(const Image& im, int iniTh, int minTh, std::vector<FastPoint>& out) {
  out.clear();
  const float W = 30;
  const int minBX = kEdge - 3, minBY = minBX;          // :785-786
  const int maxBX = im.w - kEdge + 3, maxBY = im.h - kEdge + 3;
  const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
  const int nCols = (int)(width / W), nRows = (int)(height / W);  // :796-797
  if (nCols <= 0 || nRows <= 0) return;
  const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
  std::vector<FastPoint> cell;
  for (int i = 0; i < nRows; ++i) {
    const float iniY = (float)(minBY + i * hCell);
    float maxY = iniY + hCell + 6;
    if (iniY >= maxBY - 3) continue;  // :806
    if (maxY > maxBY) maxY = (float)maxBY;
    for (int j = 0; j < nCols; ++j) {
      const float iniX = (float)(minBX + j * wCell);
      float maxX = iniX + wCell + 6;
      if (iniX >= maxBX - 6) continue;  // :815
      if (maxX > maxBX) maxX = (float)maxBX;
      const int x0 = (int)iniX, y0 = (int)iniY, cw = (int)maxX - x0, ch = (int)maxY - y0;
      const uint8_t* roi = im.row(y0) + x0;
      fast9_16(roi, cw, ch, im.w, iniTh, true, cell);                    // :830
      if (cell.empty()) fast9_16(roi, cw, ch, im.w, minTh, true, cell);  // :833-837
      for (const FastPoint& p : cell) out.push_back({p.x + j * wCell, p.y + i * hCell, p.score});  // :841-846
    }
  }
}

// ---- DistributeOctTree + DivideNode (src/ORBextractor.cc:481-765) ----------
struct QNode {
  int ulx, uly, urx, ury, blx, bly, brx, bry;  // UL, UR, BL, BR
  std::vector<int> keys;                        // indices into the candidate list, in vKeys order
  bool no_more = false;
  unsigned seq = 0;                             // creation order (tie rule)
  std::list<QNode>::iterator self;
};

// DivideNode (:481-537)
static void divide(const QNode& n, const int* xs, const int* ys, QNode out[4]) {
  const int halfX = (int)std::ceil((float)(n.urx - n.ulx) / 2);
  const int halfY = (int)std::ceil((float)(n.bry - n.uly) / 2);
  QNode &n1 = out[0], &n2 = out[1], &n3 = out[2], &n4 = out[3];
  n1.ulx = n.ulx; n1.uly = n.uly; n1.urx = n.ulx + halfX; n1.ury = n.uly;
  n1.blx = n.ulx; n1.bly = n.uly + halfY; n1.brx = n.ulx + halfX; n1.bry = n.uly + halfY;
  n2.ulx = n1.urx; n2.uly = n1.ury; n2.urx = n.urx; n2.ury = n.ury;
  n2.blx = n1.brx; n2.bly = n1.bry; n2.brx = n.urx; n2.bry = n.uly + halfY;
  n3.ulx = n1.blx; n3.uly = n1.bly; n3.urx = n1.brx; n3.ury = n1.bry;
  n3.blx = n.blx; n3.bly = n.bly; n3.brx = n1.brx; n3.bry = n.bly;
  n4.ulx = n3.urx; n4.uly = n3.ury; n4.urx = n2.brx; n4.ury = n2.bry;
  n4.blx = n3.brx; n4.bly = n3.bry; n4.brx = n.brx; n4.bry = n.bry;
  for (int q = 0; q < 4; ++q) { out[q].keys.clear(); out[q].keys.reserve(n.keys.size()); out[q].no_more = false; }
  for (int k : n.keys) {
    const float x = (float)xs[k], y = (float)ys[k];
    if (x < (float)n1.urx) {
      if (y < (float)n1.bry) n1.keys.push_back(k); else n3.keys.push_back(k);
    } else if (y < (float)n1.bry) n2.keys.push_back(k);
    else n4.keys.push_back(k);
  }
  for (int q = 0; q < 4; ++q) if (out[q].keys.size() == 1) out[q].no_more = true;
}

typedef std::pair<int, std::pair<uintptr_t, QNode*>> SizeKeyNode;

int distribute_octtree(const int* xs, const int* ys, const int* scores, int n, int minX, int maxX, int minY, int maxY,
                       int N, int flags, std::vector<int>& kept) {
  kept.clear();
  if (n == 0) return 0;
  const bool by_address = (flags & ORC_FLAG_TIE_BY_ADDRESS) != 0;
  const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));  // :543
  if (nIni < 1) return -1;  // the reference divides by zero here (portrait level)
  const float hX = (float)(maxX - minX) / nIni;                            // :545
  std::list<QNode> nodes;
  std::vector<QNode*> roots(nIni);
  unsigned seq = 0;
  for (int i = 0; i < nIni; ++i) {
    QNode r;
    r.ulx = (int)(hX * (float)i); r.uly = 0;
    r.urx = (int)(hX * (float)(i + 1)); r.ury = 0;
    r.blx = r.ulx; r.bly = maxY - minY;
    r.brx = r.urx; r.bry = maxY - minY;
    r.seq = seq++;
    nodes.push_back(r);
    roots[i] = &nodes.back();
  }
  for (int k = 0; k < n; ++k) {
    int ri = (int)((float)xs[k] / hX);  // :569
    if (ri >= nIni) ri = nIni - 1;      // cannot happen for in-range x; keeps the oracle memory safe
    roots[ri]->keys.push_back(k);
  }
  for (auto it = nodes.begin(); it != nodes.end();) {  // :572-583
    if (it->keys.size() == 1) { it->no_more = true; ++it; }
    else if (it->keys.empty()) it = nodes.erase(it);
    else ++it;
  }
  auto tie_key = [&](QNode* p) -> uintptr_t { return by_address ? (uintptr_t)p : (uintptr_t)p->seq; };
  // pushes the non-empty children to the front (:623-662); returns how many were pushed
  auto push_children = [&](QNode ch[4], std::vector<SizeKeyNode>& expandable, int* nToExpand) {
    for (int q = 0; q < 4; ++q) {
      if (ch[q].keys.empty()) continue;
      ch[q].seq = seq++;
      nodes.push_front(ch[q]);
      if (nodes.front().keys.size() > 1) {
        if (nToExpand) ++*nToExpand;
        QNode* p = &nodes.front();
        expandable.push_back(std::make_pair((int)p->keys.size(), std::make_pair(tie_key(p), p)));
        p->self = nodes.begin();
      }
    }
  };
  bool finish = false;
  std::vector<SizeKeyNode> expandable;
  QNode ch[4];
  while (!finish) {
    const int prevSize = (int)nodes.size();
    int nToExpand = 0;
    expandable.clear();
    for (auto it = nodes.begin(); it != nodes.end();) {  // :603-664 (new children sit before begin)
      if (it->no_more) { ++it; continue; }
      divide(*it, xs, ys, ch);
      push_children(ch, expandable, &nToExpand);
      it = nodes.erase(it);
    }
    if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
      finish = true;  // :669-672
    } else if ((int)nodes.size() + nToExpand * 3 > N) {  // :675
      while (!finish) {
        const int prev2 = (int)nodes.size();
        std::vector<SizeKeyNode> prevExp = expandable;
        expandable.clear();
        std::sort(prevExp.begin(), prevExp.end());  // :686 (size, then address | creation order)
        for (int j = (int)prevExp.size() - 1; j >= 0; --j) {
          QNode* p = prevExp[j].second.second;
          divide(*p, xs, ys, ch);
          push_children(ch, expandable, nullptr);
          nodes.erase(p->self);  // :730
          if ((int)nodes.size() >= N) break;
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prev2) finish = true;
      }
    }
  }
  // best response per node, first wins (:746-762)
  for (const QNode& nd : nodes) {
    int best = nd.keys[0];
    for (size_t k = 1; k < nd.keys.size(); ++k)
      if ((float)scores[nd.keys[k]] > (float)scores[best]) best = nd.keys[k];
    kept.push_back(best);
  }
  return (int)kept.size();
}

// ---- IC_Angle (src/ORBextractor.cc:77-104) ---------------------------------
float ic_angle(const uint8_t* center, int step, const int* umax) {
  int m01 = 0, m10 = 0;
  for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * center[u];
  for (int v = 1; v <= kHalfPatch; ++v) {
    int vsum = 0;
    const int d = umax[v];
    for (int u = -d; u <= d; ++u) {
      const int plus = center[u + v * step], minus = center[u - v * step];
      vsum += plus - minus;
      m10 += u * (plus + minus);
    }
    m01 += v * vsum;
  }
  return orb_fast_atan2_deg((float)m01, (float)m10);
}

void sincos_deg(float angle_deg, bool libm, float* c, float* s) {
  if (libm) {
    const float factorPI = (float)(3.14159265358979323846 / 180.f);  // :107
    const float a = angle_deg * factorPI;
    *c = cosf(a);
    *s = sinf(a);
  } else {
    orb_sincos_deg(angle_deg, c, s);
  }
}

// ---- computeOrbDescriptor (src/ORBextractor.cc:108-147) --------------------
void orb_descriptor(const uint8_t* center, int step, float angle_deg, bool libm, const int* pattern, uint8_t* desc) {
  float a, b;
  sincos_deg(angle_deg, libm, &a, &b);
  for (int i = 0; i < 32; ++i) {
    int val = 0;
    for (int j = 0; j < 8; ++j) {
      const int* pt = pattern + (i * 16 + 2 * j) * 2;
      const float x0 = (float)pt[0], y0 = (float)pt[1], x1 = (float)pt[2], y1 = (float)pt[3];
      const int r0 = round_half_even(x0 * b + y0 * a), c0 = round_half_even(x0 * a - y0 * b);
      const int r1 = round_half_even(x1 * b + y1 * a), c1 = round_half_even(x1 * a - y1 * b);
      const int t0 = center[r0 * step + c0], t1 = center[r1 * step + c1];
      val |= (t0 < t1) << j;
    }
    desc[i] = (uint8_t)val;
  }
}

const int* pattern_table() { return kPattern; }

}  // namespace orc

namespace orc {
void build_info_public(const orbfe_extractor_params& p, orbfe_extractor_info* info) { build_info(p, info); }
void compute_pyramid_public(orc_extractor* e, const uint8_t* image, int width, int height, int stride) {
  compute_pyramid(e, image, width, height, stride);
}
void detect_cells_public(const Image& im, int iniTh, int minTh, std::vector<FastPoint>& out) {
  detect_cells(im, iniTh, minTh, out);
}
int descriptor_distance_swar(const uint8_t* a, const uint8_t* b) {
  // literal form of src/ORBmatcher.cc:1715-1731
  int dist = 0;
  for (int i = 0; i < 8; ++i) {
    uint32_t pa, pb;
    std::memcpy(&pa, a + 4 * i, 4);
    std::memcpy(&pb, b + 4 * i, 4);
    uint32_t v = pa ^ pb;
    v = v - ((v >> 1) & 0x55555555u);
    v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
    dist += (int)((((v + (v >> 4)) & 0xF0F0F0Fu) * 0x1010101u) >> 24);
  }
  return dist;
}
}  // namespace orc
