// orb_matcher.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h).
//
// Restatement of the ORBmatcher methods on the hot path (reference:
// src/ORBmatcher.cc) and of the Frame grid they query (src/Frame.cc:243-258,
// :351-417) over struct-of-array views (include/orbfe_types.h).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "orb_internal.h"
#include "orb_matcher_common.h"
#include "orb_oracle.h"



extern "C" {

// ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:46-132)
// assigned[f] = index of the map point written into F.mvpMapPoints[f] by this call, -1 = untouched.
int orc_search_by_projection_map(const orbfe_frame_view* F, const orbfe_mappoints_view* mps, float th, float nnratio,
                                 int32_t* assigned) {
  orc::Grid* grid = new orc::Grid();
  orc::build_grid(F, *grid);
  for (int i = 0; i < F->n; ++i) assigned[i] = -1;
  // observation count of the map point currently sitting in F.mvpMapPoints[f] (-1 = NULL)
  std::vector<int> cur_obs(F->mp_obs, F->mp_obs + F->n);
  int nmatches = 0;
  const bool bFactor = th != 1.0;
  std::vector<int> idxs;
  for (int iMP = 0; iMP < mps->m; ++iMP) {
    if (!mps->in_view[iMP]) continue;  // :53-57
    const int lvl = mps->level[iMP];
    float r = mps->view_cos[iMP] > 0.998 ? 2.5f : 4.0f;  // RadiusByViewingCos (:134-140)
    if (bFactor) r *= th;
    orc::features_in_area(F, *grid, mps->proj_x[iMP], mps->proj_y[iMP], r * F->scale_factors[lvl], lvl - 1, lvl, idxs);
    if (idxs.empty()) continue;
    const uint8_t* dMP = mps->desc + (size_t)iMP * 32;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int idx : idxs) {
      if (cur_obs[idx] > 0) continue;  // :90-92
      if (F->u_right[idx] > 0) {
        const float er = std::fabs(mps->proj_xr[iMP] - F->u_right[idx]);
        if (er > r * F->scale_factors[lvl]) continue;
      }
      const int dist = orc::descriptor_distance(dMP, F->desc + (size_t)idx * 32);
      if (dist < bestDist) {
        bestDist2 = bestDist; bestDist = dist;
        bestLevel2 = bestLevel; bestLevel = F->keys_un[idx].octave;
        bestIdx = idx;
      } else if (dist < bestDist2) {
        bestLevel2 = F->keys_un[idx].octave;
        bestDist2 = dist;
      }
    }
    if (bestDist <= ORBFE_TH_HIGH) {
      if (bestLevel == bestLevel2 && (float)bestDist > nnratio * (float)bestDist2) continue;
      assigned[bestIdx] = iMP;  // :126
      cur_obs[bestIdx] = mps->obs[iMP];
      ++nmatches;
    }
  }
  delete grid;
  return nmatches;
}

// ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) (src/ORBmatcher.cc:1392-1538)
int orc_search_by_projection_last(const orbfe_frame_view* cur, const orbfe_camera_pose* cam,
                                  const orbfe_lastframe_view* last, float th, int mono, int check_ori,
                                  int32_t* assigned) {
  orc::Grid* grid = new orc::Grid();
  orc::build_grid(cur, *grid);
  for (int i = 0; i < cur->n; ++i) assigned[i] = -1;
  std::vector<int> cur_obs(cur->mp_obs, cur->mp_obs + cur->n);
  int nmatches = 0;
  std::vector<int> rotHist[ORBFE_HISTO_LENGTH];
  const float* T = cam->tcw;
  const float* Tl = last->tcw;
  // twc = -Rcw^T * tcw: cv::gemm general path with the transpose flag accumulates in double (:1405)
  float twc[3];
  for (int i = 0; i < 3; ++i) {
    double s = 0;
    for (int k = 0; k < 3; ++k) s += (double)T[k * 4 + i] * (double)T[k * 4 + 3];
    twc[i] = (float)(-1.0 * s);
  }
  // tlc = Rlw*twc + tlw: small-matrix path, fp32 dot product (:1410)
  const float tlcz = ((Tl[8] * twc[0] + Tl[9] * twc[1]) + Tl[10] * twc[2]) + Tl[11];
  const bool bForward = tlcz > cam->mb && !mono;
  const bool bBackward = -tlcz > cam->mb && !mono;
  std::vector<int> idxs;
  for (int i = 0; i < last->n; ++i) {
    if (!last->valid[i]) continue;
    const float* X = last->world_pos + 3 * i;
    const float xc = ((T[0] * X[0] + T[1] * X[1]) + T[2] * X[2]) + T[3];
    const float yc = ((T[4] * X[0] + T[5] * X[1]) + T[6] * X[2]) + T[7];
    const float zc = ((T[8] * X[0] + T[9] * X[1]) + T[10] * X[2]) + T[11];
    const float invzc = (float)(1.0 / zc);
    if (invzc < 0) continue;
    const float u = cam->fx * xc * invzc + cam->cx;
    const float v = cam->fy * yc * invzc + cam->cy;
    if (u < cur->min_x || u > cur->max_x) continue;
    if (v < cur->min_y || v > cur->max_y) continue;
    const int nLastOctave = last->octave[i];
    const float radius = th * cur->scale_factors[nLastOctave];
    if (bForward) orc::features_in_area(cur, *grid, u, v, radius, nLastOctave, -1, idxs);
    else if (bBackward) orc::features_in_area(cur, *grid, u, v, radius, 0, nLastOctave, idxs);
    else orc::features_in_area(cur, *grid, u, v, radius, nLastOctave - 1, nLastOctave + 1, idxs);
    if (idxs.empty()) continue;
    const uint8_t* dMP = last->desc + (size_t)i * 32;
    int bestDist = 256, bestIdx2 = -1;
    for (int i2 : idxs) {
      if (cur_obs[i2] > 0) continue;
      if (cur->u_right[i2] > 0) {
        const float ur = u - cam->mbf * invzc;
        const float er = std::fabs(ur - cur->u_right[i2]);
        if (er > radius) continue;
      }
      const int dist = orc::descriptor_distance(dMP, cur->desc + (size_t)i2 * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= ORBFE_TH_HIGH) {
      assigned[bestIdx2] = i;
      cur_obs[bestIdx2] = last->obs[i];
      ++nmatches;
      if (check_ori) rotHist[orc::rot_bin(last->angle[i], cur->keys_un[bestIdx2].angle)].push_back(bestIdx2);
    }
  }
  if (check_ori) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    orc::three_maxima(rotHist, ORBFE_HISTO_LENGTH, ind1, ind2, ind3);
    for (int b = 0; b < ORBFE_HISTO_LENGTH; ++b) {
      if (b == ind1 || b == ind2 || b == ind3) continue;
      for (int f : rotHist[b]) { assigned[f] = -2; --nmatches; }  // :1527 mvpMapPoints[f] = NULL (not 'untouched')
    }
  }
  delete grid;
  return nmatches;
}

// ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:702-872) + CheckDistEpipolarLine (:147-167)
int orc_search_for_triangulation(const orbfe_keyframe_view* k1, const orbfe_keyframe_view* k2, const float* F12,
                                 float ex, float ey, int only_stereo, int check_ori, int32_t* pairs, int cap) {
  int nmatches = 0;
  std::vector<int> matches12(k1->n, -1);
  std::vector<int> rotHist[ORBFE_HISTO_LENGTH];
  int a = 0, b = 0;
  while (a < k1->n_nodes && b < k2->n_nodes) {
    if (k1->node_ids[a] == k2->node_ids[b]) {
      for (int p1 = k1->node_offsets[a]; p1 < k1->node_offsets[a + 1]; ++p1) {
        const int idx1 = k1->node_features[p1];
        if (k1->has_mp[idx1]) continue;
        const bool bStereo1 = k1->u_right[idx1] >= 0;
        if (only_stereo && !bStereo1) continue;
        const orbfe_keypoint& kp1 = k1->keys_un[idx1];
        const uint8_t* d1 = k1->desc + (size_t)idx1 * 32;
        int bestDist = ORBFE_TH_LOW, bestIdx2 = -1;
        for (int p2 = k2->node_offsets[b]; p2 < k2->node_offsets[b + 1]; ++p2) {
          const int idx2 = k2->node_features[p2];
          if (k2->has_mp[idx2]) continue;  // vbMatched2 is never set in the reference (:722,773)
          const bool bStereo2 = k2->u_right[idx2] >= 0;
          if (only_stereo && !bStereo2) continue;
          const int dist = orc::descriptor_distance(d1, k2->desc + (size_t)idx2 * 32);
          if (dist > ORBFE_TH_LOW || dist > bestDist) continue;
          const orbfe_keypoint& kp2 = k2->keys_un[idx2];
          if (!bStereo1 && !bStereo2) {
            const float dex = ex - kp2.x, dey = ey - kp2.y;
            if (dex * dex + dey * dey < 100 * k2->scale_factors[kp2.octave]) continue;
          }
          // CheckDistEpipolarLine
          const float la = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
          const float lb = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
          const float lc = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
          const float num = la * kp2.x + lb * kp2.y + lc;
          const float den = la * la + lb * lb;
          if (den == 0) continue;
          const float dsqr = num * num / den;
          if ((double)dsqr < 3.84 * (double)k2->level_sigma2[kp2.octave]) { bestIdx2 = idx2; bestDist = dist; }
        }
        if (bestIdx2 >= 0) {
          matches12[idx1] = bestIdx2;
          ++nmatches;
          if (check_ori) rotHist[orc::rot_bin(kp1.angle, k2->keys_un[bestIdx2].angle)].push_back(idx1);
        }
      }
      ++a; ++b;
    } else if (k1->node_ids[a] < k2->node_ids[b]) {
      ++a;  // lower_bound on a sorted id list == advance
    } else {
      ++b;
    }
  }
  if (check_ori) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    orc::three_maxima(rotHist, ORBFE_HISTO_LENGTH, ind1, ind2, ind3);
    for (int h = 0; h < ORBFE_HISTO_LENGTH; ++h) {
      if (h == ind1 || h == ind2 || h == ind3) continue;
      for (int f : rotHist[h]) { matches12[f] = -1; --nmatches; }
    }
  }
  int n = 0;
  for (int i = 0; i < k1->n; ++i) {
    if (matches12[i] < 0) continue;
    if (n < cap) { pairs[2 * n] = i; pairs[2 * n + 1] = matches12[i]; }
    ++n;
  }
  return n;
}

// Brute-force Hamming (BASELINE.json config 5): per query the two smallest (dist<<32 | index).
void orc_hamming_bruteforce(const uint8_t* query, int nq, const uint8_t* map, int nm, int64_t index_base, uint64_t* best,
                            uint64_t* second) {
  for (int q = 0; q < nq; ++q) {
    uint64_t b1 = ~0ull, b2 = ~0ull;
    const uint8_t* dq = query + (size_t)q * 32;
    for (int j = 0; j < nm; ++j) {
      const uint64_t key = ((uint64_t)orc::descriptor_distance(dq, map + (size_t)j * 32) << 32) | (uint64_t)(index_base + j);
      if (key < b1) { b2 = b1; b1 = key; }
      else if (key < b2) b2 = key;
    }
    best[q] = b1;
    second[q] = b2;
  }
}

}  // extern "C"
