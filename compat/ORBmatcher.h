// ORBmatcher.h — drop-in replacement of the reference's include/ORBmatcher.h:45-114.  Same namespace, class name,
// signatures and constants; all eleven members are defined in compat/ORBmatcher.cc on top of liborbfe.so (SURVEY.md
// 8(a) rows a10-a13 and 8(f)-1), so the reference's src/ORBmatcher.cc drops out of the build.
#ifndef ORBMATCHER_H
#define ORBMATCHER_H

#include <set>
#include <utility>
#include <vector>

#ifdef ORBFE_HAVE_OPENCV
#include "Frame.h"
#include "KeyFrame.h"
#include "MapPoint.h"
#else
#include "slam_min.h"
#endif

namespace ORB_SLAM2 {

class ORBmatcher {
 public:
  ORBmatcher(float nnratio = 0.6, bool checkOri = true);

  // src/ORBmatcher.cc:1715-1731
  static int DescriptorDistance(const cv::Mat& a, const cv::Mat& b);
  // src/ORBmatcher.cc:46-132 (Tracking::SearchLocalPoints, Tracking.cc:1277-1284)
  int SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th = 3);
  // src/ORBmatcher.cc:1392-1538 (Tracking::TrackWithMotionModel, Tracking.cc:917-946)
  int SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono);
  // src/ORBmatcher.cc:702-872 (LocalMapping::CreateNewMapPoints, LocalMapping.cc:276)
  int SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                             std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo);

  // SURVEY.md 8(f)-1 (src/ORBmatcher.cc:1540, :321, :182, :563, :446, :1158, :879, :1033)
  int SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                         const int ORBdist);
  int SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                         std::vector<MapPoint*>& vpMatched, int th);
  int SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches);
  int SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12);
  int SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                              std::vector<int>& vnMatches12, int windowSize = 10);
  int SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
                   const cv::Mat& R12, const cv::Mat& t12, const float th);
  int Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th = 3.0);
  int Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
           std::vector<MapPoint*>& vpReplacePoint);

  static const int TH_LOW;
  static const int TH_HIGH;
  static const int HISTO_LENGTH;

  static void SetDevice(int device) { sDevice = device; }

 protected:
  float mfNNratio;
  bool mbCheckOrientation;

 private:
  static int sDevice;
};

}  // namespace ORB_SLAM2

#endif
