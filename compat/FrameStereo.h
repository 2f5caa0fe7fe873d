// FrameStereo.h — body of the drop-in Frame::ComputeStereoMatches (reference: src/Frame.cc:496-686).
// In the reference's tree, src/Frame.cc keeps everything but this method, whose body becomes:
//
//     void Frame::ComputeStereoMatches() {
//         ORB_SLAM2::compat::ComputeStereoMatches(*mpORBextractorLeft, *mpORBextractorRight, N, mbf, mbf / fx,
//                                                 mvuRight, mvDepth);
//     }
//
// It runs on the device-resident pyramids / keypoints / descriptors that the two extractor handles produced for
// this frame (src/Frame.cc:89-92), so no image or descriptor crosses PCIe again; only mvuRight / mvDepth come back.
#pragma once
#include <cstdio>
#include <vector>

#include "../include/orbfe.h"
#include "ORBextractor.h"

namespace ORB_SLAM2 {
namespace compat {

inline int ComputeStereoMatches(ORBextractor& left, ORBextractor& right, int N, float mbf, float mb,
                                std::vector<float>& mvuRight, std::vector<float>& mvDepth) {
  mvuRight.assign(N, -1.0f);  // src/Frame.cc:498-499
  mvDepth.assign(N, -1.0f);
  if (N == 0 || !left.handle() || !right.handle()) return 0;
  int32_t matched = 0;
  const int rc = orbfe_stereo_match(left.handle(), right.handle(), 1, mbf, mb, mvuRight.data(), mvDepth.data(), N, &matched);
  if (rc != ORBFE_OK) {
    std::fprintf(stderr, "ComputeStereoMatches: %s\n", orbfe_last_error());
    mvuRight.assign(N, -1.0f);
    mvDepth.assign(N, -1.0f);
    return 0;
  }
  return matched;
}

}  // namespace compat
}  // namespace ORB_SLAM2
