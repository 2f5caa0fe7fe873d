// slam_min.h — the members of the reference's Frame / KeyFrame / MapPoint (and DBoW2::FeatureVector) that the
// GPU-backed ORBmatcher methods touch, with the reference's names and types, for machines WITHOUT OpenCV C++, Eigen
// and DBoW2 (this image).  compat/ORBmatcher.cc compiles against this header here and against the reference's own
// include/Frame.h, include/KeyFrame.h, include/MapPoint.h when ORBFE_HAVE_OPENCV is defined.  Nothing here computes
// matches: these are plain data holders for the compile check and the parity driver (compat/test_compat_matcher.cc).
#pragma once
#include <map>
#include <set>
#include <vector>

#include "cv_min.h"

namespace DBoW2 {
typedef unsigned int NodeId;
typedef unsigned int WordId;
typedef double WordValue;
typedef std::map<NodeId, std::vector<unsigned int> > FeatureVector;  // Thirdparty/DBoW2/DBoW2/FeatureVector.h
typedef std::map<WordId, WordValue> BowVector;                        // Thirdparty/DBoW2/DBoW2/BowVector.h
}  // namespace DBoW2

namespace ORB_SLAM2 {

class KeyFrame;

class MapPoint {  // include/MapPoint.h:92-97, src/MapPoint.cc:320-324,394-444
 public:
  float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
  bool mbTrackInView = false;
  int mnTrackScaleLevel = 0;
  float mTrackViewCos = 0;
  bool isBad() { return mbBad; }
  int Observations() { return nObs; }
  cv::Mat GetDescriptor() { return mDescriptor; }
  cv::Mat GetWorldPos() { return mWorldPos; }
  cv::Mat GetNormal() { return mNormalVector; }
  float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }  // src/MapPoint.cc:394-404
  float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }
  float GetMaxDistance() { return mfMaxDistance; }  // the one-line accessor of INTEGRATION.md 4 (ORBFE_MAPPOINT_RAW_DISTANCE)
  int GetIndexInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) ? (int)mObservations[pKF] : -1; }
  bool IsInKeyFrame(KeyFrame* pKF) { return mObservations.count(pKF) != 0; }
  void AddObservation(KeyFrame* pKF, size_t idx) { if (!mObservations.count(pKF)) { mObservations[pKF] = idx; ++nObs; } }
  void Replace(MapPoint* pMP) { if (pMP != this) { mbBad = true; mpReplaced = pMP; } }  // bookkeeping stand-in
  // test-side state
  bool mbBad = false;
  int nObs = 0;
  cv::Mat mDescriptor;    // 1 x 32 CV_8U
  cv::Mat mWorldPos;      // 3 x 1 CV_32F
  cv::Mat mNormalVector;  // 3 x 1 CV_32F
  float mfMinDistance = 0, mfMaxDistance = 0;
  std::map<KeyFrame*, size_t> mObservations;
  MapPoint* mpReplaced = NULL;
};

class Frame {  // include/Frame.h
 public:
  int N = 0;
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
  std::vector<float> mvuRight;
  cv::Mat mDescriptors;  // N x 32 CV_8U
  std::vector<MapPoint*> mvpMapPoints;
  std::vector<bool> mvbOutlier;
  std::vector<float> mvScaleFactors, mvLevelSigma2;
  DBoW2::FeatureVector mFeatVec;
  int mnScaleLevels = 0;
  float mfLogScaleFactor = 0;
  cv::Mat mTcw;  // 4 x 4 CV_32F
  float mbf = 0, mb = 0;
  static float fx, fy, cx, cy;
  static float mnMinX, mnMaxX, mnMinY, mnMaxY;
  static float mfGridElementWidthInv, mfGridElementHeightInv;
};

class KeyFrame {  // include/KeyFrame.h
 public:
  int N = 0;
  std::vector<cv::KeyPoint> mvKeysUn;
  std::vector<float> mvuRight;
  cv::Mat mDescriptors;
  DBoW2::FeatureVector mFeatVec;
  std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
  float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0;
  int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;  // include/KeyFrame.h (const int in the reference)
  float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
  int mnScaleLevels = 0;
  float mfLogScaleFactor = 0;
  MapPoint* GetMapPoint(const size_t& idx) { return mvpMapPoints[idx]; }
  std::vector<MapPoint*> GetMapPointMatches() { return mvpMapPoints; }
  std::set<MapPoint*> GetMapPoints() {
    std::set<MapPoint*> s;
    for (size_t i = 0; i < mvpMapPoints.size(); ++i)
      if (mvpMapPoints[i] && !mvpMapPoints[i]->isBad()) s.insert(mvpMapPoints[i]);  // src/KeyFrame.cc:232-245
    return s;
  }
  void AddMapPoint(MapPoint* pMP, const size_t& idx) { mvpMapPoints[idx] = pMP; }
  cv::Mat GetCameraCenter() { return Ow; }
  cv::Mat GetRotation() { return Rcw; }
  cv::Mat GetTranslation() { return tcw; }
  // test-side state
  std::vector<MapPoint*> mvpMapPoints;
  cv::Mat Ow, Rcw, tcw;  // 3x1, 3x3, 3x1 CV_32F
};

}  // namespace ORB_SLAM2
