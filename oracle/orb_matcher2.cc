// orb_matcher2.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h).
//
// Restatement of the seven remaining ORBmatcher members (reference: src/ORBmatcher.cc) over the
// struct-of-array views of include/orbfe_types.h:
//   SearchByBoW(KeyFrame*, Frame&, ...)                        :182-319
//   SearchByProjection(KeyFrame*, Scw, points, matched, th)    :321-434
//   SearchForInitialization                                    :446-561
//   SearchByBoW(KeyFrame*, KeyFrame*, ...)                     :563-696
//   Fuse(KeyFrame*, points, th) / Fuse(KeyFrame*, Scw, ...)    :879-1029 / :1033-1156
//   SearchBySim3                                               :1158-1382
//   SearchByProjection(Frame&, KeyFrame*, found, th, ORBdist)  :1540-1667
//
// cv::Mat arithmetic is restated with OpenCV's small-matrix conventions (the same ones orb_matcher.cc uses for the
// last-frame overload): a 3x3 * 3x1 product (+ 3x1) is an in-order fp32 dot product followed by an fp32 add;
// a product with a transposed operand, Mat::dot and cv::norm accumulate in double; `M / s` and `s * M` scale by
// (float)(1.0 / s) resp. (float)s element-wise.  No FMA contraction (-ffp-contract=off).
// Documented deviation 4: `log` in MapPoint::PredictScale is orb_log_f32 (include/orb_math_shared.h), the float
// rounding of a double-precision evaluation, instead of the machine's libm logf.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <vector>

#include "../include/orb_math_shared.h"
#include "orb_internal.h"
#include "orb_matcher_common.h"
#include "orb_oracle.h"

namespace orc {

struct Pose {
  float R[9], t[3], Ow[3];
};

// Rcw, tcw, Ow of the target: a plain pose, or the decomposition of a Sim3 (src/ORBmatcher.cc:329-334, :1041-1046)
static Pose decompose(const orbfe_projection* pr) {
  Pose P;
  const float* T = pr->tcw;
  if (pr->is_sim3) {
    const double d = ((double)T[0] * (double)T[0] + (double)T[1] * (double)T[1]) + (double)T[2] * (double)T[2];
    const float scw = (float)std::sqrt(d);           // const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)))
    const float inv = (float)(1.0 / (double)scw);    // sRcw/scw, .../scw
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 3; ++c) P.R[r * 3 + c] = T[r * 4 + c] * inv;
      P.t[r] = T[r * 4 + 3] * inv;
    }
  } else {
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 3; ++c) P.R[r * 3 + c] = T[r * 4 + c];
      P.t[r] = T[r * 4 + 3];
    }
  }
  if (pr->has_ow) {
    for (int i = 0; i < 3; ++i) P.Ow[i] = pr->ow[i];
  } else {
    for (int i = 0; i < 3; ++i) {  // -Rcw.t()*tcw: transposed operand -> double accumulation
      double s = 0;
      for (int k = 0; k < 3; ++k) s += (double)P.R[k * 3 + i] * (double)P.t[k];
      P.Ow[i] = (float)(-1.0 * s);
    }
  }
  return P;
}

static inline void rigid(const float* R, const float* t, const float* X, float* o) {
  o[0] = ((R[0] * X[0] + R[1] * X[1]) + R[2] * X[2]) + t[0];
  o[1] = ((R[3] * X[0] + R[4] * X[1]) + R[5] * X[2]) + t[1];
  o[2] = ((R[6] * X[0] + R[7] * X[1]) + R[8] * X[2]) + t[2];
}

static inline float norm3(const float* v) {  // cv::norm of a 3x1 float Mat
  const double s = ((double)v[0] * (double)v[0] + (double)v[1] * (double)v[1]) + (double)v[2] * (double)v[2];
  return (float)std::sqrt(s);
}

static inline double dot3(const float* a, const float* b) {  // Mat::dot
  return ((double)a[0] * (double)b[0] + (double)a[1] * (double)b[1]) + (double)a[2] * (double)b[2];
}

// KeyFrame::IsInImage (src/KeyFrame.cc:624-627)
static inline bool in_image(const orbfe_frame_view* F, float x, float y) {
  return x >= F->min_x && x < F->max_x && y >= F->min_y && y < F->max_y;
}

}  // namespace orc

extern "C" {

// ---- SearchByBoW, both overloads.  kf_kf = 0: (KeyFrame*, Frame&) :182-319, matches[f] (size k2->n) = index of the
// keyframe feature whose map point goes into vpMapPointMatches[f]; kf_kf = 1: (KeyFrame*, KeyFrame*) :563-696,
// matches[idx1] (size k1->n) = idx2 with vpMatches12[idx1] = vpMapPoints2[idx2].  has_mp = "map point present and
// not bad".  -1 = NULL.
int orc_search_by_bow(const orbfe_keyframe_view* k1, const orbfe_keyframe_view* k2, int kf_kf, float nnratio,
                      int check_ori, int32_t* matches) {
  const int n_out = kf_kf ? k1->n : k2->n;
  for (int i = 0; i < n_out; ++i) matches[i] = -1;
  std::vector<char> matched2(k2->n, 0);  // vpMapPointMatches[realIdxF]!=NULL (:240) / vbMatched2 (:576)
  std::vector<int> rotHist[ORBFE_HISTO_LENGTH];
  int nmatches = 0;
  int a = 0, b = 0;
  while (a < k1->n_nodes && b < k2->n_nodes) {
    if (k1->node_ids[a] == k2->node_ids[b]) {
      for (int p1 = k1->node_offsets[a]; p1 < k1->node_offsets[a + 1]; ++p1) {
        const int idx1 = k1->node_features[p1];
        if (!k1->has_mp[idx1]) continue;  // :224-228, :600-603
        const uint8_t* d1 = k1->desc + (size_t)idx1 * 32;
        int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
        for (int p2 = k2->node_offsets[b]; p2 < k2->node_offsets[b + 1]; ++p2) {
          const int idx2 = k2->node_features[p2];
          if (matched2[idx2]) continue;
          if (kf_kf && !k2->has_mp[idx2]) continue;  // :617-621
          const int dist = orc::descriptor_distance(d1, k2->desc + (size_t)idx2 * 32);
          if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
          else if (dist < bestDist2) bestDist2 = dist;
        }
        const bool pass = kf_kf ? bestDist1 < ORBFE_TH_LOW : bestDist1 <= ORBFE_TH_LOW;  // :639 vs :259
        if (pass && (float)bestDist1 < nnratio * (float)bestDist2) {
          matched2[bestIdx2] = 1;
          if (kf_kf) matches[idx1] = bestIdx2; else matches[bestIdx2] = idx1;
          if (check_ori) {
            const int bin = orc::rot_bin(k1->keys_un[idx1].angle, k2->keys_un[bestIdx2].angle);
            rotHist[bin].push_back(kf_kf ? idx1 : bestIdx2);
          }
          ++nmatches;
        }
      }
      ++a; ++b;
    } else if (k1->node_ids[a] < k2->node_ids[b]) {
      ++a;  // lower_bound on the sorted ids == advance
    } else {
      ++b;
    }
  }
  if (check_ori) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    orc::three_maxima(rotHist, ORBFE_HISTO_LENGTH, ind1, ind2, ind3);
    for (int h = 0; h < ORBFE_HISTO_LENGTH; ++h) {
      if (h == ind1 || h == ind2 || h == ind3) continue;
      for (int j : rotHist[h]) { matches[j] = -1; --nmatches; }
    }
  }
  return nmatches;
}

// ---- SearchForInitialization (:446-561).  prev_matched: n1 x 2 floats (vbPrevMatched), updated in place (:556-558).
int orc_search_for_initialization(const orbfe_frame_view* F1, const orbfe_frame_view* F2, float* prev_matched,
                                  int window, float nnratio, int check_ori, int32_t* matches12) {
  orc::Grid* grid = new orc::Grid();
  orc::build_grid(F2, *grid);
  int nmatches = 0;
  for (int i = 0; i < F1->n; ++i) matches12[i] = -1;
  std::vector<int> rotHist[ORBFE_HISTO_LENGTH];
  std::vector<int> vMatchedDistance(F2->n, INT_MAX), vnMatches21(F2->n, -1);
  std::vector<int> idxs;
  for (int i1 = 0; i1 < F1->n; ++i1) {
    const int level1 = F1->keys_un[i1].octave;
    if (level1 > 0) continue;
    orc::features_in_area(F2, *grid, prev_matched[2 * i1], prev_matched[2 * i1 + 1], (float)window, level1, level1, idxs);
    if (idxs.empty()) continue;
    const uint8_t* d1 = F1->desc + (size_t)i1 * 32;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : idxs) {
      const int dist = orc::descriptor_distance(d1, F2->desc + (size_t)i2 * 32);
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) bestDist2 = dist;
    }
    if (bestDist <= ORBFE_TH_LOW) {
      if ((float)bestDist < (float)bestDist2 * nnratio) {
        if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; --nmatches; }
        matches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        ++nmatches;
        if (check_ori) rotHist[orc::rot_bin(F1->keys_un[i1].angle, F2->keys_un[bestIdx2].angle)].push_back(i1);
      }
    }
  }
  if (check_ori) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    orc::three_maxima(rotHist, ORBFE_HISTO_LENGTH, ind1, ind2, ind3);
    for (int h = 0; h < ORBFE_HISTO_LENGTH; ++h) {
      if (h == ind1 || h == ind2 || h == ind3) continue;
      for (int idx1 : rotHist[h])
        if (matches12[idx1] >= 0) { matches12[idx1] = -1; --nmatches; }
    }
  }
  for (int i1 = 0; i1 < F1->n; ++i1)
    if (matches12[i1] >= 0) {
      prev_matched[2 * i1] = F2->keys_un[matches12[i1]].x;
      prev_matched[2 * i1 + 1] = F2->keys_un[matches12[i1]].y;
    }
  delete grid;
  return nmatches;
}

// ---- SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (:1540-1667).  cur->mp_obs[i] >= 0 means
// CurrentFrame.mvpMapPoints[i] != NULL.  assigned[f] = row written into mvpMapPoints[f], -1 untouched, -2 = NULLed by
// the rotation check.
int orc_search_by_projection_reloc(const orbfe_frame_view* cur, const orbfe_projection* pr,
                                   const orbfe_points3d_view* pts, float th, int orb_dist, int check_ori,
                                   int32_t* assigned) {
  orc::Grid* grid = new orc::Grid();
  orc::build_grid(cur, *grid);
  for (int i = 0; i < cur->n; ++i) assigned[i] = -1;
  std::vector<char> occ(cur->n);
  for (int i = 0; i < cur->n; ++i) occ[i] = cur->mp_obs[i] >= 0;
  const orc::Pose P = orc::decompose(pr);
  std::vector<int> rotHist[ORBFE_HISTO_LENGTH];
  std::vector<int> idxs;
  int nmatches = 0;
  for (int i = 0; i < pts->m; ++i) {
    if (!pts->valid[i]) continue;  // :1560-1562
    const float* X = pts->world_pos + 3 * i;
    float xc[3];
    orc::rigid(P.R, P.t, X, xc);
    const float invzc = (float)(1.0 / (double)xc[2]);  // :1570
    const float u = pr->fx * xc[0] * invzc + pr->cx;
    const float v = pr->fy * xc[1] * invzc + pr->cy;
    if (u < cur->min_x || u > cur->max_x) continue;
    if (v < cur->min_y || v > cur->max_y) continue;
    const float PO[3] = {X[0] - P.Ow[0], X[1] - P.Ow[1], X[2] - P.Ow[2]};
    const float dist3D = orc::norm3(PO);
    const float maxDistance = pts->max_dist_inv[i];  // GetMaxDistanceInvariance(), MapPoint.cc:394-404
    const float minDistance = pts->min_dist_inv[i];
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    const int lvl = orb_predict_scale(pts->max_distance[i], dist3D, pr->log_scale_factor, cur->nlevels);
    const float radius = th * cur->scale_factors[lvl];
    orc::features_in_area(cur, *grid, u, v, radius, lvl - 1, lvl + 1, idxs);
    if (idxs.empty()) continue;
    const uint8_t* dMP = pts->desc + (size_t)i * 32;
    int bestDist = 256, bestIdx2 = -1;
    for (int i2 : idxs) {
      if (occ[i2]) continue;  // :1609
      const int dist = orc::descriptor_distance(dMP, cur->desc + (size_t)i2 * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= orb_dist) {
      assigned[bestIdx2] = i;
      occ[bestIdx2] = 1;
      ++nmatches;
      if (check_ori) rotHist[orc::rot_bin(pts->angle[i], cur->keys_un[bestIdx2].angle)].push_back(bestIdx2);
    }
  }
  if (check_ori) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    orc::three_maxima(rotHist, ORBFE_HISTO_LENGTH, ind1, ind2, ind3);
    for (int h = 0; h < ORBFE_HISTO_LENGTH; ++h) {
      if (h == ind1 || h == ind2 || h == ind3) continue;
      for (int f : rotHist[h]) { assigned[f] = -2; --nmatches; }
    }
  }
  delete grid;
  return nmatches;
}

// ---- SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (:321-434).  kf->mp_obs[i] >= 0 means
// vpMatched[i] != NULL on entry; pts->valid = !isBad() && !spAlreadyFound.count(pMP).
int orc_search_by_projection_sim3(const orbfe_frame_view* kf, const orbfe_projection* pr,
                                  const orbfe_points3d_view* pts, int th, int32_t* assigned) {
  orc::Grid* grid = new orc::Grid();
  orc::build_grid(kf, *grid);
  for (int i = 0; i < kf->n; ++i) assigned[i] = -1;
  std::vector<char> occ(kf->n);
  for (int i = 0; i < kf->n; ++i) occ[i] = kf->mp_obs[i] >= 0;
  const orc::Pose P = orc::decompose(pr);
  std::vector<int> idxs;
  int nmatches = 0;
  for (int iMP = 0; iMP < pts->m; ++iMP) {
    if (!pts->valid[iMP]) continue;
    const float* X = pts->world_pos + 3 * iMP;
    float pc[3];
    orc::rigid(P.R, P.t, X, pc);
    if (pc[2] < 0.0) continue;
    const float invz = 1 / pc[2];
    const float x = pc[0] * invz, y = pc[1] * invz;
    const float u = pr->fx * x + pr->cx, v = pr->fy * y + pr->cy;
    if (!orc::in_image(kf, u, v)) continue;
    const float maxDistance = pts->max_dist_inv[iMP], minDistance = pts->min_dist_inv[iMP];
    const float PO[3] = {X[0] - P.Ow[0], X[1] - P.Ow[1], X[2] - P.Ow[2]};
    const float dist = orc::norm3(PO);
    if (dist < minDistance || dist > maxDistance) continue;
    if (orc::dot3(PO, pts->normal + 3 * iMP) < 0.5 * dist) continue;
    const int lvl = orb_predict_scale(pts->max_distance[iMP], dist, pr->log_scale_factor, kf->nlevels);
    const float radius = th * kf->scale_factors[lvl];
    orc::features_in_area(kf, *grid, u, v, radius, -1, -1, idxs);  // KeyFrame::GetFeaturesInArea has no level gate
    if (idxs.empty()) continue;
    const uint8_t* dMP = pts->desc + (size_t)iMP * 32;
    int bestDist = 256, bestIdx = -1;
    for (int idx : idxs) {
      if (occ[idx]) continue;
      const int kpLevel = kf->keys_un[idx].octave;
      if (kpLevel < lvl - 1 || kpLevel > lvl) continue;
      const int d = orc::descriptor_distance(dMP, kf->desc + (size_t)idx * 32);
      if (d < bestDist) { bestDist = d; bestIdx = idx; }
    }
    if (bestDist <= ORBFE_TH_LOW) {
      assigned[bestIdx] = iMP;
      occ[bestIdx] = 1;
      ++nmatches;
    }
  }
  delete grid;
  return nmatches;
}

// ---- Fuse, both overloads (:879-1029 with pr->is_sim3 == 0, :1033-1156 with is_sim3 == 1): the SEARCH part.
// best_idx[i] = bestIdx of row i when bestDist <= TH_LOW, else -1.  The map surgery that follows (Replace /
// AddObservation / AddMapPoint, :1008-1023) stays with the caller, which walks best_idx in row order.
int orc_fuse(const orbfe_frame_view* kf, const orbfe_projection* pr, const orbfe_points3d_view* pts, float th,
             const float* inv_level_sigma2, int32_t* best_idx) {
  orc::Grid* grid = new orc::Grid();
  orc::build_grid(kf, *grid);
  const orc::Pose P = orc::decompose(pr);
  std::vector<int> idxs;
  int nFused = 0;
  for (int i = 0; i < pts->m; ++i) {
    best_idx[i] = -1;
    if (!pts->valid[i]) continue;
    const float* X = pts->world_pos + 3 * i;
    float pc[3];
    orc::rigid(P.R, P.t, X, pc);
    if (pc[2] < 0.0f) continue;
    const float invz = pr->is_sim3 ? (float)(1.0 / (double)pc[2]) : 1 / pc[2];  // :1075 vs :913
    const float x = pc[0] * invz, y = pc[1] * invz;
    const float u = pr->fx * x + pr->cx, v = pr->fy * y + pr->cy;
    if (!orc::in_image(kf, u, v)) continue;
    const float ur = u - pr->bf * invz;
    const float maxDistance = pts->max_dist_inv[i], minDistance = pts->min_dist_inv[i];
    const float PO[3] = {X[0] - P.Ow[0], X[1] - P.Ow[1], X[2] - P.Ow[2]};
    const float dist3D = orc::norm3(PO);
    if (dist3D < minDistance || dist3D > maxDistance) continue;
    if (orc::dot3(PO, pts->normal + 3 * i) < 0.5 * dist3D) continue;
    const int lvl = orb_predict_scale(pts->max_distance[i], dist3D, pr->log_scale_factor, kf->nlevels);
    const float radius = th * kf->scale_factors[lvl];
    orc::features_in_area(kf, *grid, u, v, radius, -1, -1, idxs);
    if (idxs.empty()) continue;
    const uint8_t* dMP = pts->desc + (size_t)i * 32;
    int bestDist = 256, bestIdx = -1;
    for (int idx : idxs) {
      const orbfe_keypoint& kp = kf->keys_un[idx];
      const int kpLevel = kp.octave;
      if (kpLevel < lvl - 1 || kpLevel > lvl) continue;
      if (!pr->is_sim3) {  // :968-992
        const float ex = u - kp.x, ey = v - kp.y;
        if (kf->u_right[idx] >= 0) {
          const float er = ur - kf->u_right[idx];
          const float e2 = ex * ex + ey * ey + er * er;
          if (e2 * inv_level_sigma2[kpLevel] > 7.8) continue;
        } else {
          const float e2 = ex * ex + ey * ey;
          if (e2 * inv_level_sigma2[kpLevel] > 5.99) continue;
        }
      }
      const int d = orc::descriptor_distance(dMP, kf->desc + (size_t)idx * 32);
      if (d < bestDist) { bestDist = d; bestIdx = idx; }
    }
    if (bestDist <= ORBFE_TH_LOW) { best_idx[i] = bestIdx; ++nFused; }
  }
  delete grid;
  return nFused;
}

// ---- SearchBySim3 (:1158-1382).  p1/p2 = GetMapPointMatches() of the two keyframes (m == n), valid = present,
// not bad and not vbAlreadyMatched.  match12[i1] = idx2 for the agreeing pairs (vpMatches12[i1] =
// vpMapPoints2[idx2]), -1 otherwise.
int orc_search_by_sim3(const orbfe_frame_view* kf1, const orbfe_frame_view* kf2, const orbfe_points3d_view* p1,
                       const orbfe_points3d_view* p2, const orbfe_sim3_pair* s, float th, int32_t* match12) {
  float R1w[9], t1w[3], R2w[9], t2w[3];
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) { R1w[r * 3 + c] = s->t1w[r * 4 + c]; R2w[r * 3 + c] = s->t2w[r * 4 + c]; }
    t1w[r] = s->t1w[r * 4 + 3];
    t2w[r] = s->t2w[r * 4 + 3];
  }
  float sR12[9], sR21[9], t21[3];
  const float inv_s = (float)(1.0 / (double)s->s12);
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) {
      sR12[r * 3 + c] = s->r12[r * 3 + c] * s->s12;  // s12*R12
      sR21[r * 3 + c] = s->r12[c * 3 + r] * inv_s;   // (1.0/s12)*R12.t()
    }
  for (int r = 0; r < 3; ++r)  // t21 = -sR21*t12: gemm with alpha = -1 on the fp32 dot product
    t21[r] = -((sR21[r * 3] * s->t12[0] + sR21[r * 3 + 1] * s->t12[1]) + sR21[r * 3 + 2] * s->t12[2]);
  auto search = [&](const orbfe_points3d_view* pts, const float* Ra, const float* ta, const float* Rb, const float* tb,
                    const orbfe_frame_view* target, float log_sf, std::vector<int>& out) {
    orc::Grid* grid = new orc::Grid();
    orc::build_grid(target, *grid);
    out.assign(pts->m, -1);
    std::vector<int> idxs;
    for (int i = 0; i < pts->m; ++i) {
      if (!pts->valid[i]) continue;
      float pa[3], pb[3];
      orc::rigid(Ra, ta, pts->world_pos + 3 * i, pa);
      orc::rigid(Rb, tb, pa, pb);
      if (pb[2] < 0.0) continue;
      const float invz = (float)(1.0 / (double)pb[2]);
      const float x = pb[0] * invz, y = pb[1] * invz;
      const float u = s->fx * x + s->cx, v = s->fy * y + s->cy;
      if (!orc::in_image(target, u, v)) continue;
      const float maxDistance = pts->max_dist_inv[i], minDistance = pts->min_dist_inv[i];
      const float dist3D = orc::norm3(pb);
      if (dist3D < minDistance || dist3D > maxDistance) continue;
      const int lvl = orb_predict_scale(pts->max_distance[i], dist3D, log_sf, target->nlevels);
      const float radius = th * target->scale_factors[lvl];
      orc::features_in_area(target, *grid, u, v, radius, -1, -1, idxs);
      if (idxs.empty()) continue;
      const uint8_t* dMP = pts->desc + (size_t)i * 32;
      int bestDist = INT_MAX, bestIdx = -1;
      for (int idx : idxs) {
        const int oct = target->keys_un[idx].octave;
        if (oct < lvl - 1 || oct > lvl) continue;
        const int d = orc::descriptor_distance(dMP, target->desc + (size_t)idx * 32);
        if (d < bestDist) { bestDist = d; bestIdx = idx; }
      }
      if (bestDist <= ORBFE_TH_HIGH) out[i] = bestIdx;
    }
    delete grid;
  };
  std::vector<int> vnMatch1, vnMatch2;
  search(p1, R1w, t1w, sR21, t21, kf2, s->log_scale_factor2, vnMatch1);   // :1204-1281
  search(p2, R2w, t2w, sR12, s->t12, kf1, s->log_scale_factor1, vnMatch2);  // :1284-1361
  int nFound = 0;
  for (int i1 = 0; i1 < p1->m; ++i1) {
    match12[i1] = -1;
    const int idx2 = vnMatch1[i1];
    if (idx2 >= 0 && idx2 < (int)vnMatch2.size() && vnMatch2[idx2] == i1) { match12[i1] = idx2; ++nFound; }
  }
  return nFound;
}

// ---- Tracking::SearchLocalPoints (src/Tracking.cc:1258-1284): Frame::isInFrustum (src/Frame.cc:285-349) per local
// map point, then SearchByProjection(Frame&, vpMapPoints, th) on the fields it filled.
int orc_search_local_points(const orbfe_frame_view* F, const orbfe_projection* pr, const orbfe_points3d_view* pts,
                            const int32_t* obs, float viewingCosLimit, float th, float nnratio, int32_t* assigned,
                            uint8_t* in_view, float* proj_x, float* proj_y, float* proj_xr, int32_t* level,
                            float* view_cos) {
  const orc::Pose P = orc::decompose(pr);
  for (int i = 0; i < pts->m; ++i) {
    in_view[i] = 0;  // pMP->mbTrackInView = false
    proj_x[i] = proj_y[i] = proj_xr[i] = view_cos[i] = 0.f;
    level[i] = 0;
    if (!pts->valid[i]) continue;
    const float* X = pts->world_pos + 3 * i;
    float Pc[3];
    orc::rigid(P.R, P.t, X, Pc);
    if (Pc[2] < 0.0f) continue;
    const float invz = 1.0f / Pc[2];
    const float u = pr->fx * Pc[0] * invz + pr->cx;
    const float v = pr->fy * Pc[1] * invz + pr->cy;
    if (u < F->min_x || u > F->max_x) continue;
    if (v < F->min_y || v > F->max_y) continue;
    const float PO[3] = {X[0] - P.Ow[0], X[1] - P.Ow[1], X[2] - P.Ow[2]};
    const float dist = orc::norm3(PO);
    if (dist < pts->min_dist_inv[i] || dist > pts->max_dist_inv[i]) continue;
    const float viewCos = (float)(orc::dot3(PO, pts->normal + 3 * i) / (double)dist);
    if (viewCos < viewingCosLimit) continue;
    proj_x[i] = u; proj_y[i] = v; view_cos[i] = viewCos;
    level[i] = orb_predict_scale(pts->max_distance[i], dist, pr->log_scale_factor, F->nlevels);
    proj_xr[i] = u - pr->bf * invz;
    in_view[i] = 1;
  }
  orbfe_mappoints_view mv;
  mv.m = pts->m; mv.proj_x = proj_x; mv.proj_y = proj_y; mv.proj_xr = proj_xr; mv.level = level;
  mv.view_cos = view_cos; mv.in_view = in_view; mv.obs = obs; mv.desc = pts->desc;
  return orc_search_by_projection_map(F, &mv, th, nnratio, assigned);
}

// ---- MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:252-318) for a batch of points (CSR of descriptors)
void orc_compute_distinctive_descriptors(const uint8_t* desc, const int32_t* offsets, int n_points, int32_t* best_idx,
                                         uint8_t* best_desc) {
  for (int p = 0; p < n_points; ++p) {
    const int beg = offsets[p], N = offsets[p + 1] - beg;
    if (best_desc) std::memset(best_desc + (size_t)p * 32, 0, 32);
    if (N <= 0) { best_idx[p] = -1; continue; }
    std::vector<std::vector<float>> Distances(N, std::vector<float>(N));
    for (int i = 0; i < N; ++i) {
      Distances[i][i] = 0;
      for (int j = i + 1; j < N; ++j) {
        const int distij = orc::descriptor_distance(desc + (size_t)(beg + i) * 32, desc + (size_t)(beg + j) * 32);
        Distances[i][j] = distij;
        Distances[j][i] = distij;
      }
    }
    int BestMedian = INT_MAX, BestIdx = 0;
    for (int i = 0; i < N; ++i) {
      std::vector<int> vDists(Distances[i].begin(), Distances[i].end());
      std::sort(vDists.begin(), vDists.end());
      const int median = vDists[(size_t)(0.5 * (N - 1))];
      if (median < BestMedian) { BestMedian = median; BestIdx = i; }
    }
    best_idx[p] = BestIdx;
    if (best_desc) std::memcpy(best_desc + (size_t)p * 32, desc + (size_t)(beg + BestIdx) * 32, 32);
  }
}

float orc_log_f32(float x) { return orb_log_f32(x); }
int orc_predict_scale(float max_distance, float dist, float log_scale_factor, int nlevels) {
  return orb_predict_scale(max_distance, dist, log_scale_factor, nlevels);
}

}  // extern "C"
