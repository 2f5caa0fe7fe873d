// orb_matcher_common.h — helpers shared by the matcher restatements of the CPU ORACLE (test infrastructure only,
// see orb_oracle.h): the Frame grid (src/Frame.cc:243-258, :351-417), ComputeThreeMaxima and the rotation bin.
#ifndef ORB_MATCHER_COMMON_H_
#define ORB_MATCHER_COMMON_H_
#include <algorithm>
#include <cmath>
#include <vector>

#include "orb_internal.h"

namespace orc {

// Frame::AssignFeaturesToGrid + PosInGrid (src/Frame.cc:243-258, :407-417)
struct Grid {
  std::vector<int> cell[ORBFE_GRID_COLS][ORBFE_GRID_ROWS];
};

inline void build_grid(const orbfe_frame_view* F, Grid& g) {
  for (int i = 0; i < F->n; ++i) {
    const orbfe_keypoint& kp = F->keys_un[i];
    const int px = (int)std::round((kp.x - F->min_x) * F->grid_inv_w);
    const int py = (int)std::round((kp.y - F->min_y) * F->grid_inv_h);
    if (px < 0 || px >= ORBFE_GRID_COLS || py < 0 || py >= ORBFE_GRID_ROWS) continue;
    g.cell[px][py].push_back(i);
  }
}

// Frame::GetFeaturesInArea (src/Frame.cc:351-405)
inline void features_in_area(const orbfe_frame_view* F, const Grid& g, float x, float y, float r, int minLevel,
                             int maxLevel, std::vector<int>& out) {
  out.clear();
  const int nMinCellX = std::max(0, (int)std::floor((x - F->min_x - r) * F->grid_inv_w));
  if (nMinCellX >= ORBFE_GRID_COLS) return;
  const int nMaxCellX = std::min(ORBFE_GRID_COLS - 1, (int)std::ceil((x - F->min_x + r) * F->grid_inv_w));
  if (nMaxCellX < 0) return;
  const int nMinCellY = std::max(0, (int)std::floor((y - F->min_y - r) * F->grid_inv_h));
  if (nMinCellY >= ORBFE_GRID_ROWS) return;
  const int nMaxCellY = std::min(ORBFE_GRID_ROWS - 1, (int)std::ceil((y - F->min_y + r) * F->grid_inv_h));
  if (nMaxCellY < 0) return;
  const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
  for (int ix = nMinCellX; ix <= nMaxCellX; ++ix) {
    for (int iy = nMinCellY; iy <= nMaxCellY; ++iy) {
      const std::vector<int>& cell = g.cell[ix][iy];
      for (int idx : cell) {
        const orbfe_keypoint& kp = F->keys_un[idx];
        if (bCheckLevels) {
          if (kp.octave < minLevel) continue;
          if (maxLevel >= 0 && kp.octave > maxLevel) continue;
        }
        const float dx = kp.x - x, dy = kp.y - y;
        if (std::fabs(dx) < r && std::fabs(dy) < r) out.push_back(idx);
      }
    }
  }
}

// ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1669-1710)
inline void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; ++i) {
    const int s = (int)histo[i].size();
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

static inline int rot_bin(float a1, float a2) {
  const float factor = 1.0f / ORBFE_HISTO_LENGTH;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * factor);
  if (bin == ORBFE_HISTO_LENGTH) bin = 0;
  return bin;
}

}  // namespace orc

#endif
