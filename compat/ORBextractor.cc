// ORBextractor.cc — drop-in ORB_SLAM2::ORBextractor on top of the C ABI (include/orbfe.h).
// Replaces the reference's src/ORBextractor.cc in libORB_SLAM2.so (CMakeLists.txt:48-68).
#include "ORBextractor.h"

#include <cstdio>
#include <cstring>

#include "../include/orbfe.h"

namespace ORB_SLAM2 {

int ORBextractor::sDevice = 0;

static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint must be 28 bytes (x,y,size,angle,response,octave,class_id)");

ORBextractor::ORBextractor(int _nfeatures, float _scaleFactor, int _nlevels, int _iniThFAST, int _minThFAST)
    : nfeatures(_nfeatures), scaleFactor(_scaleFactor), nlevels(_nlevels), iniThFAST(_iniThFAST), minThFAST(_minThFAST),
      mDevice(sDevice) {
  // the getter tables do not depend on the image size: take them from a probe handle (src/ORBextractor.cc:415-446)
  orbfe_extractor_params p = {nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST};
  orbfe_extractor* probe = nullptr;
  if (orbfe_extractor_create(&p, 256, 256, 1, mDevice, &probe) != ORBFE_OK) {
    std::fprintf(stderr, "ORBextractor: %s\n", orbfe_last_error());
    return;
  }
  orbfe_extractor_info info;
  orbfe_extractor_get_info(probe, &info);
  orbfe_extractor_destroy(probe);
  mvScaleFactor.assign(info.scale_factors, info.scale_factors + nlevels);
  mvInvScaleFactor.assign(info.inv_scale_factors, info.inv_scale_factors + nlevels);
  mvLevelSigma2.assign(info.level_sigma2, info.level_sigma2 + nlevels);
  mvInvLevelSigma2.assign(info.inv_level_sigma2, info.inv_level_sigma2 + nlevels);
  mnFeaturesPerLevel.assign(info.features_per_level, info.features_per_level + nlevels);
  umax.assign(info.umax, info.umax + 16);
  mvImagePyramid.resize(nlevels);
}

ORBextractor::~ORBextractor() { orbfe_extractor_destroy(mHandle); }

void ORBextractor::setDevice(int device) {
  if (device == mDevice) return;
  orbfe_extractor_destroy(mHandle);  // the arenas live on the old device; the next operator() re-creates them
  mHandle = nullptr;
  mWidth = mHeight = 0;
  mDevice = device;
}

void ORBextractor::ensureHandle(int width, int height) {
  if (mHandle && width == mWidth && height == mHeight) return;
  orbfe_extractor_destroy(mHandle);
  mHandle = nullptr;
  orbfe_extractor_params p = {nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST};
  if (orbfe_extractor_create(&p, width, height, 1, mDevice, &mHandle) != ORBFE_OK) {
    std::fprintf(stderr, "ORBextractor: %s\n", orbfe_last_error());
    mHandle = nullptr;
    return;
  }
  orbfe_extractor_info info;
  orbfe_extractor_get_info(mHandle, &info);
  mCapacity = info.capacity;
  mDescScratch.resize((size_t)mCapacity * 32);
  mWidth = width;
  mHeight = height;
}

void ORBextractor::operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                              cv::OutputArray _descriptors) {
#ifdef ORBFE_HAVE_OPENCV
  cv::Mat image = _image.getMat();
#else
  const cv::Mat& image = _image;
#endif
  if (image.empty()) return;  // src/ORBextractor.cc:1068
  ensureHandle(image.cols, image.rows);
  _keypoints.clear();
  mOk = mHandle != nullptr;
  if (!mHandle) {  // the reference has no error channel: empty outputs + a log line
    _descriptors.release();
    return;
  }
  _keypoints.resize(mCapacity);
  unsigned char* desc = mDescScratch.data();
  int32_t n = 0;
  const int rc = orbfe_extract(mHandle, image.ptr(0), (int)image.step, reinterpret_cast<orbfe_keypoint*>(_keypoints.data()),
                               desc, mCapacity, &n);
  if (rc != ORBFE_OK) {
    std::fprintf(stderr, "ORBextractor: %s\n", orbfe_last_error());
    mOk = false;
    n = 0;
  }
  _keypoints.resize(n);
  if (n == 0) {  // :1086-1087
    _descriptors.release();
  } else {
    _descriptors.create(n, 32, CV_8U);  // :1090
#ifdef ORBFE_HAVE_OPENCV
    cv::Mat out = _descriptors.getMat();
#else
    cv::Mat& out = _descriptors;
#endif
    std::memcpy(out.ptr(0), desc, (size_t)n * 32);
  }
  if (keepHostPyramid) {
    for (int lv = 0; lv < nlevels; ++lv) {
      int w = 0, h = 0, pitch = 0;
      orbfe_extractor_level(mHandle, 0, lv, nullptr, &pitch, &w, &h);
      mvImagePyramid[lv].create(h, w, CV_8UC1);
      orbfe_extractor_download_level(mHandle, 0, lv, mvImagePyramid[lv].ptr(0), (int)mvImagePyramid[lv].step);
    }
  }
}

}  // namespace ORB_SLAM2
