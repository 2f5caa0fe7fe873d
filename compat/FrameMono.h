// FrameMono.h — bodies of the drop-in Frame::UndistortKeyPoints, Frame::ComputeImageBounds and
// Frame::ComputeStereoFromRGBD (reference: src/Frame.cc:429-494, :688-712), the steps the mono and RGB-D Frame
// constructors (:118-228) run after ExtractORB.  In the reference's tree, src/Frame.cc keeps everything else:
//
//     void Frame::UndistortKeyPoints() {
//         ORB_SLAM2::compat::UndistortKeyPoints(*mpORBextractorLeft, mvKeys, Camera(mK, mDistCoef), mvKeysUn);
//     }
//     void Frame::ComputeImageBounds(const cv::Mat& imLeft) {
//         float b[4];
//         orbfe_image_bounds(&cam, imLeft.cols, imLeft.rows, b);  mnMinX = b[0]; mnMaxX = b[1]; mnMinY = b[2]; mnMaxY = b[3];
//     }
//     void Frame::ComputeStereoFromRGBD(const cv::Mat& imDepth) {
//         ORB_SLAM2::compat::ComputeStereoFromRGBD(*mpORBextractorLeft, N, imDepth.ptr<float>(), imDepth.step / 4, mbf,
//                                                  mvuRight, mvDepth);
//     }
//
// They run on the keypoints the extractor handle produced for this frame (still resident on the device); mvKeysUn,
// mvuRight and mvDepth come back.
#pragma once
#include <cstdio>
#include <vector>

#include "../include/orbfe.h"
#include "ORBextractor.h"

namespace ORB_SLAM2 {
namespace compat {

// mK (3x3 CV_32F: fx, fy, cx, cy) and mDistCoef (4x1 or 5x1 CV_32F) as Tracking builds them (src/Tracking.cc:58-100)
inline orbfe_camera Camera(const cv::Mat& K, const cv::Mat& distCoef) {
  orbfe_camera c;
  c.fx = K.at<float>(0, 0); c.fy = K.at<float>(1, 1); c.cx = K.at<float>(0, 2); c.cy = K.at<float>(1, 2);
  for (int i = 0; i < 5; ++i) c.k[i] = i < distCoef.rows ? distCoef.at<float>(i) : 0.0f;
  return c;
}

inline bool UndistortKeyPoints(ORBextractor& ex, const std::vector<cv::KeyPoint>& mvKeys, const orbfe_camera& cam,
                               std::vector<cv::KeyPoint>& mvKeysUn) {
  static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint is the 28-byte record of the C ABI");
  const int N = (int)mvKeys.size();
  mvKeysUn.resize(N);
  if (N == 0) return true;
  if (!ex.handle() ||
      orbfe_undistort_keypoints(ex.handle(), 1, &cam, reinterpret_cast<orbfe_keypoint*>(mvKeysUn.data()), N) != ORBFE_OK) {
    std::fprintf(stderr, "UndistortKeyPoints: %s\n", orbfe_last_error());
    mvKeysUn = mvKeys;  // the reference has no error path; keep the frame usable
    return false;
  }
  return true;
}

inline bool ComputeStereoFromRGBD(ORBextractor& ex, int N, const float* depth, int row_stride_floats, float mbf,
                                  std::vector<float>& mvuRight, std::vector<float>& mvDepth) {
  mvuRight.assign(N, -1.0f);  // src/Frame.cc:690-691
  mvDepth.assign(N, -1.0f);
  if (N == 0) return true;
  if (!ex.handle() || orbfe_stereo_from_rgbd(ex.handle(), 1, depth, 0, row_stride_floats, mbf, mvuRight.data(),
                                             mvDepth.data(), N) != ORBFE_OK) {
    std::fprintf(stderr, "ComputeStereoFromRGBD: %s\n", orbfe_last_error());
    mvuRight.assign(N, -1.0f);
    mvDepth.assign(N, -1.0f);
    return false;
  }
  return true;
}

}  // namespace compat
}  // namespace ORB_SLAM2
