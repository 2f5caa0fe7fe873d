// orb_stereo.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h).
//
// Restatement of Frame::ComputeStereoMatches (reference: src/Frame.cc:496-686).
// Deviation 2 of SURVEY.md 8(c): `mb` is an explicit argument (the reference
// reads it before it is initialised, Frame.cc:534 vs :127), and an empty
// vDistIdx returns early instead of indexing element 0 (:673).
#include <algorithm>
#include <climits>
#include <cmath>
#include <utility>
#include <vector>

#include "orb_internal.h"
#include "orb_oracle.h"

extern "C" int orc_stereo_matches(const orc_extractor* left, const orc_extractor* right, const orbfe_keypoint* kpL,
                                  const uint8_t* descL, int nL, const orbfe_keypoint* kpR, const uint8_t* descR,
                                  int nR, float mbf, float mb, float* u_right, float* depth) {
  for (int i = 0; i < nL; ++i) { u_right[i] = -1.0f; depth[i] = -1.0f; }  // :498-499
  const int thOrbDist = (ORBFE_TH_HIGH + ORBFE_TH_LOW) / 2;                // :501
  const int nRows = left->pyramid[0].h;                                    // :503
  const float* scaleF = left->info.scale_factors;
  const float* invScaleF = left->info.inv_scale_factors;

  // row table (:511-531)
  std::vector<std::vector<size_t>> rowIdx(nRows);
  for (int i = 0; i < nRows; ++i) rowIdx[i].reserve(200);
  for (int iR = 0; iR < nR; ++iR) {
    const float kpY = kpR[iR].y;
    const float r = 2.0f * scaleF[kpR[iR].octave];
    const int maxr = (int)std::ceil(kpY + r);
    const int minr = (int)std::floor(kpY - r);
    for (int yi = minr; yi <= maxr; ++yi)
      if (yi >= 0 && yi < nRows) rowIdx[yi].push_back(iR);  // the reference does not bounds-check (:529-530)
  }

  const float minZ = mb;
  const float minD = 0;
  const float maxD = mbf / minZ;  // :536

  std::vector<std::pair<int, int>> distIdx;
  distIdx.reserve(nL);

  for (int iL = 0; iL < nL; ++iL) {
    const orbfe_keypoint& kL = kpL[iL];
    const int levelL = kL.octave;
    const float vL = kL.y, uL = kL.x;
    const int rowi = (int)vL;  // vRowIndices[vL]: float -> size_t truncation (:549)
    if (rowi < 0 || rowi >= nRows) continue;
    const std::vector<size_t>& cands = rowIdx[rowi];
    if (cands.empty()) continue;
    const float minU = uL - maxD, maxU = uL - minD;
    if (maxU < 0) continue;

    int bestDist = ORBFE_TH_HIGH;
    size_t bestIdxR = 0;
    const uint8_t* dL = descL + (size_t)iL * 32;
    for (size_t iC = 0; iC < cands.size(); ++iC) {
      const size_t iR = cands[iC];
      const orbfe_keypoint& kR = kpR[iR];
      if (kR.octave < levelL - 1 || kR.octave > levelL + 1) continue;
      const float uR = kR.x;
      if (uR >= minU && uR <= maxU) {
        const int dist = orc::descriptor_distance(dL, descR + iR * 32);
        if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
      }
    }

    if (bestDist < thOrbDist) {  // :591
      const float uR0 = kpR[bestIdxR].x;
      const float sf = invScaleF[kL.octave];
      const float scaleduL = std::round(kL.x * sf);
      const float scaledvL = std::round(kL.y * sf);
      const float scaleduR0 = std::round(uR0 * sf);
      const int w = 5;
      const orc::Image& imL = left->pyramid[kL.octave];
      const orc::Image& imR = right->pyramid[kL.octave];
      const int cuL = (int)scaleduL, cvL = (int)scaledvL, cuR = (int)scaleduR0;
      // memory-safety guard (never hit on this path)
      if (cvL - w < 0 || cvL + w >= imL.h || cuL - w < 0 || cuL + w >= imL.w) continue;
      const int centerL = imL.row(cvL)[cuL];

      int bestSad = INT_MAX, bestInc = 0;
      const int L = 5;
      float dists[2 * 5 + 1];
      const float iniu = scaleduR0 + L - w;
      const float endu = scaleduR0 + L + w + 1;
      if (iniu < 0 || endu >= imR.w) continue;  // :612-615
      if (cuR - L - w < 0) continue;            // memory-safety guard (never hit)
      for (int inc = -L; inc <= L; ++inc) {
        const int centerR = imR.row(cvL)[cuR + inc];
        int sad = 0;
        for (int dy = -w; dy <= w; ++dy) {
          const uint8_t* rl = imL.row(cvL + dy) + cuL;
          const uint8_t* rr = imR.row(cvL + dy) + cuR + inc;
          for (int dx = -w; dx <= w; ++dx) sad += std::abs((rl[dx] - centerL) - (rr[dx] - centerR));
        }
        const float dist = (float)sad;  // cv::norm(IL, IR, NORM_L1): exact integer in fp
        if (dist < (float)bestSad) { bestSad = (int)dist; bestInc = inc; }
        dists[L + inc] = dist;
      }
      if (bestInc == -L || bestInc == L) continue;  // :633

      const float d1 = dists[L + bestInc - 1], d2 = dists[L + bestInc], d3 = dists[L + bestInc + 1];
      const float deltaR = (d1 - d3) / (2.0f * (d1 + d3 - 2.0f * d2));  // :644
      if (deltaR < -1 || deltaR > 1) continue;

      float bestuR = scaleF[kL.octave] * ((float)scaleduR0 + (float)bestInc + deltaR);  // :653
      float disparity = uL - bestuR;
      if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) {
          disparity = 0.01;           // double literal -> float (:661)
          bestuR = uL - 0.01;         // float - double -> float (:662)
        }
        depth[iL] = mbf / disparity;
        u_right[iL] = bestuR;
        distIdx.push_back(std::pair<int, int>(bestSad, iL));
      }
    }
  }

  if (distIdx.empty()) return 0;
  std::sort(distIdx.begin(), distIdx.end());  // :672
  const float median = (float)distIdx[distIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;  // :674
  int kept = (int)distIdx.size();
  for (int i = (int)distIdx.size() - 1; i >= 0; --i) {
    if ((float)distIdx[i].first < thDist) break;
    u_right[distIdx[i].second] = -1;
    depth[distIdx[i].second] = -1;
    --kept;
  }
  return kept;
}
