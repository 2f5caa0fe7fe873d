// orb_frame.cc — CPU ORACLE (test infrastructure only, see orb_oracle.h) of what the mono / RGB-D Frame constructors do
// after ExtractORB: Frame::UndistortKeyPoints (src/Frame.cc:429-460), Frame::ComputeImageBounds (:462-494) and
// Frame::ComputeStereoFromRGBD (:688-712), restated line by line on plain arrays.  cv::undistortPoints is the
// third-party part (OpenCV, un-vendored): orb_undistort_point of include/orb_math_shared.h restates OpenCV 4.13's
// cvUndistortPointsInternal and is pinned bit-for-bit to python-cv2 4.13 (tests/golden/undistort.npz).
#include <algorithm>

#include "../include/orb_math_shared.h"
#include "orb_oracle.h"

extern "C" {

// :429-460
void orc_undistort_keypoints(const orbfe_keypoint* keys, int n, const orbfe_camera* cam, orbfe_keypoint* keys_un) {
  if (cam->k[0] == 0.0f) {  // :431-435  mvKeysUn = mvKeys
    for (int i = 0; i < n; ++i) keys_un[i] = keys[i];
    return;
  }
  for (int i = 0; i < n; ++i) {  // :438-458: the N x 2 matrix of pt.x, pt.y through cv::undistortPoints(mat, mat, mK, mDistCoef, Mat(), mK)
    orbfe_keypoint kp = keys[i];
    float xu, yu;
    orb_undistort_point(kp.x, kp.y, cam->fx, cam->fy, cam->cx, cam->cy, cam->k, &xu, &yu);
    kp.x = xu;
    kp.y = yu;
    keys_un[i] = kp;
  }
}

// :462-494
void orc_image_bounds(const orbfe_camera* cam, int cols, int rows, float* bounds) {
  if (cam->k[0] != 0.0f) {
    const float mx[4] = {0.0f, (float)cols, 0.0f, (float)cols};  // :466-470
    const float my[4] = {0.0f, 0.0f, (float)rows, (float)rows};
    float x[4], y[4];
    for (int i = 0; i < 4; ++i) orb_undistort_point(mx[i], my[i], cam->fx, cam->fy, cam->cx, cam->cy, cam->k, &x[i], &y[i]);
    bounds[0] = std::min(x[0], x[2]);  // mnMinX (:477)
    bounds[1] = std::max(x[1], x[3]);  // mnMaxX
    bounds[2] = std::min(y[0], y[1]);  // mnMinY
    bounds[3] = std::max(y[2], y[3]);  // mnMaxY
  } else {  // :487-493
    bounds[0] = 0.0f;
    bounds[1] = (float)cols;
    bounds[2] = 0.0f;
    bounds[3] = (float)rows;
  }
}

// :688-712
void orc_stereo_from_rgbd(const orbfe_keypoint* keys, const orbfe_keypoint* keys_un, int n, const float* depth, int w,
                          int h, int stride, float mbf, float* u_right, float* depth_out) {
  for (int i = 0; i < n; ++i) {
    u_right[i] = -1.0f;  // :690-691
    depth_out[i] = -1.0f;
    const float v = keys[i].y, u = keys[i].x;          // :698-699 (the distorted position)
    const int iv = (int)v, iu = (int)u;                // imDepth.at<float>(v, u): float -> int conversion
    if (iv < 0 || iv >= h || iu < 0 || iu >= w) continue;  // memory guard (never taken: keypoints are >= 19 px inside)
    const float d = depth[(size_t)iv * stride + iu];   // :701
    if (d > 0) {                                        // :703
      depth_out[i] = d;                                 // :705
      u_right[i] = keys_un[i].x - mbf / d;              // :706
    }
  }
}

}  // extern "C"
