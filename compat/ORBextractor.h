// ORBextractor.h — drop-in replacement of the reference's include/ORBextractor.h:45-108.
// Same namespace, class name, constructor, operator() signature, getters and the PUBLIC member mvImagePyramid, so
// that Frame, Tracking, LocalMapping and LoopClosing compile and link unchanged.  All work is done by liborbfe.so
// (include/orbfe.h) on the GPU; there is no CPU implementation behind this class.
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <list>
#include <vector>

#include "cv_min.h"

struct orbfe_extractor;

namespace ORB_SLAM2 {

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };  // include/ORBextractor.h:49 (unused by the reference too)

  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
  ~ORBextractor();
  ORBextractor(const ORBextractor&) = delete;
  ORBextractor& operator=(const ORBextractor&) = delete;

  // Compute the ORB features and descriptors on an image.  Mask is ignored (as in the reference).
  void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints,
                  cv::OutputArray descriptors);

  int inline GetLevels() { return nlevels; }
  float inline GetScaleFactor() { return (float)scaleFactor; }
  std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
  std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
  std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
  std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

  // Public data member of the reference (read by Frame::ComputeStereoMatches).  The pyramid lives on the device;
  // this vector is filled (one D2H per level) by operator() only when keepHostPyramid is true.  The compat
  // ComputeStereoMatches (compat/FrameStereo.h) does not need it: it runs on the device-resident pyramids.
  std::vector<cv::Mat> mvImagePyramid;
  bool keepHostPyramid = false;

  // the device handle (for compat/FrameStereo.h)
  orbfe_extractor* handle() const { return mHandle; }
  // The GPU the NEXT constructed extractor lives on (process-wide default), and the device of THIS object: the two
  // extractors of a rig, or the rigs of several cameras, can sit on different GPUs.
  static void SetDevice(int device) { sDevice = device; }
  void setDevice(int device);
  int device() const { return mDevice; }
  // false when the last operator() failed on the device side (lost device, CUDA error): the reference has no error
  // channel, so the outputs are empty like a featureless frame; this tells the two apart.
  bool ok() const { return mOk; }

 protected:
  int nfeatures;
  double scaleFactor;
  int nlevels;
  int iniThFAST;
  int minThFAST;
  std::vector<int> mnFeaturesPerLevel;
  std::vector<int> umax;
  std::vector<float> mvScaleFactor;
  std::vector<float> mvInvScaleFactor;
  std::vector<float> mvLevelSigma2;
  std::vector<float> mvInvLevelSigma2;

 private:
  void ensureHandle(int width, int height);
  orbfe_extractor* mHandle = nullptr;
  int mWidth = 0, mHeight = 0, mCapacity = 0;
  int mDevice = 0;
  bool mOk = true;
  std::vector<unsigned char> mDescScratch;  // capacity x 32, kept across calls (no per-frame allocation)
  static int sDevice;
};

}  // namespace ORB_SLAM2

#endif
