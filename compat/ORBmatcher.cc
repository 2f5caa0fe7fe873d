// ORBmatcher.cc — the four GPU-backed members of ORB_SLAM2::ORBmatcher on top of the C ABI (include/orbfe.h).
// Each method snapshots the fields it reads into structure-of-arrays host buffers ON THE CALLING THREAD (the MapPoint
// / KeyFrame accessors take mutexes, src/MapPoint.cc:320-324), calls liborbfe.so on a per-thread matcher context
// (stack-local ORBmatcher objects are used concurrently by Tracking, LocalMapping and LoopClosing), and applies the
// result to the caller's objects exactly where the reference does.  No matching is computed on the CPU here.
#include "ORBmatcher.h"

#include <cstdio>
#include <cstring>

#include "../include/orbfe.h"

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = ORBFE_TH_HIGH;            // src/ORBmatcher.cc:37-39
const int ORBmatcher::TH_LOW = ORBFE_TH_LOW;
const int ORBmatcher::HISTO_LENGTH = ORBFE_HISTO_LENGTH;
int ORBmatcher::sDevice = 0;

namespace {

// one matcher context (stream + workspace) per host thread, created on first use, destroyed with the thread
struct ThreadMatcher {
  orbfe_matcher* h = nullptr;
  int device = -1;
  ~ThreadMatcher() { orbfe_matcher_destroy(h); }
  orbfe_matcher* get(int dev) {
    if (h && device == dev) return h;
    orbfe_matcher_destroy(h);
    h = nullptr;
    device = dev;
    if (orbfe_matcher_create(dev, &h) != ORBFE_OK) {
      std::fprintf(stderr, "ORBmatcher: %s\n", orbfe_last_error());
      h = nullptr;
    }
    return h;
  }
};
thread_local ThreadMatcher t_matcher;

static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint must be 28 bytes");

// Frame members read by both SearchByProjection overloads (src/ORBmatcher.cc:46-132, :1392-1538; src/Frame.cc:351-417)
struct FrameSnapshot {
  std::vector<int32_t> mp_obs;
  orbfe_frame_view v;
  explicit FrameSnapshot(Frame& F) {
    mp_obs.resize(F.N);
    for (int i = 0; i < F.N; ++i) {
      MapPoint* p = F.mvpMapPoints[i];
      mp_obs[i] = p ? p->Observations() : -1;  // :90-92, :1471-1473
    }
    v.n = F.N;
    v.keys_un = reinterpret_cast<const orbfe_keypoint*>(F.mvKeysUn.data());
    v.u_right = F.mvuRight.data();
    v.desc = F.mDescriptors.ptr(0);
    v.mp_obs = mp_obs.data();
    v.min_x = Frame::mnMinX; v.min_y = Frame::mnMinY; v.max_x = Frame::mnMaxX; v.max_y = Frame::mnMaxY;
    v.grid_inv_w = Frame::mfGridElementWidthInv;
    v.grid_inv_h = Frame::mfGridElementHeightInv;
    v.nlevels = (int)F.mvScaleFactors.size();
    v.scale_factors = F.mvScaleFactors.data();
  }
};

void copy_pose(const cv::Mat& Tcw, float* out12) {
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) out12[r * 4 + c] = Tcw.at<float>(r, c);
}

}  // namespace

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  int32_t d = 256;
  if (!m || orbfe_descriptor_distance(m, a.ptr(0), b.ptr(0), 1, &d) != ORBFE_OK)
    std::fprintf(stderr, "ORBmatcher::DescriptorDistance: %s\n", orbfe_last_error());
  return d;
}

int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  const int M = (int)vpMapPoints.size();
  if (!m || M == 0 || F.N == 0) return 0;
  FrameSnapshot fs(F);
  std::vector<float> px(M), py(M), pxr(M), vc(M);
  std::vector<int32_t> lvl(M), obs(M);
  std::vector<uint8_t> inview(M), desc((size_t)M * 32);
  for (int i = 0; i < M; ++i) {
    MapPoint* p = vpMapPoints[i];
    inview[i] = (p->mbTrackInView && !p->isBad()) ? 1 : 0;  // :52-56
    px[i] = p->mTrackProjX; py[i] = p->mTrackProjY; pxr[i] = p->mTrackProjXR;
    lvl[i] = p->mnTrackScaleLevel; vc[i] = p->mTrackViewCos;
    obs[i] = p->Observations();
    if (inview[i]) {
      const cv::Mat d = p->GetDescriptor();  // :73 (a clone under the point's mutex)
      std::memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
    }
  }
  orbfe_mappoints_view mv;
  mv.m = M; mv.proj_x = px.data(); mv.proj_y = py.data(); mv.proj_xr = pxr.data(); mv.level = lvl.data();
  mv.view_cos = vc.data(); mv.in_view = inview.data(); mv.obs = obs.data(); mv.desc = desc.data();
  std::vector<int32_t> assigned(F.N, -1);
  int32_t n = 0;
  if (orbfe_search_by_projection_map(m, &fs.v, &mv, th, mfNNratio, assigned.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchByProjection: %s\n", orbfe_last_error());
    return 0;
  }
  for (int f = 0; f < F.N; ++f)
    if (assigned[f] >= 0) F.mvpMapPoints[f] = vpMapPoints[assigned[f]];  // :126
  return n;
}

int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  if (!m || CurrentFrame.N == 0 || LastFrame.N == 0) return 0;
  FrameSnapshot fs(CurrentFrame);
  const int n = LastFrame.N;
  std::vector<uint8_t> valid(n), desc((size_t)n * 32);
  std::vector<float> pos((size_t)n * 3), angle(n);
  std::vector<int32_t> obs(n), octave(n);
  for (int i = 0; i < n; ++i) {
    MapPoint* p = LastFrame.mvpMapPoints[i];
    valid[i] = (p && !LastFrame.mvbOutlier[i]) ? 1 : 0;  // :1417-1421
    octave[i] = LastFrame.mvKeys[i].octave;              // :1443
    angle[i] = LastFrame.mvKeysUn[i].angle;              // :1501
    if (valid[i]) {
      const cv::Mat x = p->GetWorldPos();
      pos[3 * i] = x.at<float>(0); pos[3 * i + 1] = x.at<float>(1); pos[3 * i + 2] = x.at<float>(2);
      obs[i] = p->Observations();
      const cv::Mat d = p->GetDescriptor();
      std::memcpy(&desc[(size_t)i * 32], d.ptr(0), 32);
    }
  }
  orbfe_lastframe_view lv;
  lv.n = n; lv.valid = valid.data(); lv.world_pos = pos.data(); lv.obs = obs.data(); lv.desc = desc.data();
  lv.octave = octave.data(); lv.angle = angle.data();
  copy_pose(LastFrame.mTcw, lv.tcw);
  orbfe_camera_pose cam;
  copy_pose(CurrentFrame.mTcw, cam.tcw);
  cam.fx = Frame::fx; cam.fy = Frame::fy; cam.cx = Frame::cx; cam.cy = Frame::cy;
  cam.mbf = CurrentFrame.mbf; cam.mb = CurrentFrame.mb;
  std::vector<int32_t> assigned(CurrentFrame.N, -1);
  int32_t nm = 0;
  if (orbfe_search_by_projection_last(m, &fs.v, &cam, &lv, th, bMono ? 1 : 0, mbCheckOrientation ? 1 : 0,
                                      assigned.data(), &nm) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchByProjection: %s\n", orbfe_last_error());
    return 0;
  }
  for (int f = 0; f < CurrentFrame.N; ++f) {
    if (assigned[f] >= 0) CurrentFrame.mvpMapPoints[f] = LastFrame.mvpMapPoints[assigned[f]];  // :1496
    else if (assigned[f] == -2) CurrentFrame.mvpMapPoints[f] = static_cast<MapPoint*>(NULL);  // :1527
  }
  return nm;
}

int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                       std::vector<std::pair<size_t, size_t> >& vMatchedPairs,
                                       const bool bOnlyStereo) {
  vMatchedPairs.clear();
  orbfe_matcher* m = t_matcher.get(sDevice);
  if (!m || pKF1->N == 0 || pKF2->N == 0) return 0;
  // epipole in the second image (:708-716)
  const cv::Mat Cw = pKF1->GetCameraCenter(), R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
  float C2[3];
  for (int r = 0; r < 3; ++r)
    C2[r] = R2w.at<float>(r, 0) * Cw.at<float>(0) + R2w.at<float>(r, 1) * Cw.at<float>(1) +
            R2w.at<float>(r, 2) * Cw.at<float>(2) + t2w.at<float>(r);
  const float invz = 1.0f / C2[2];
  const float ex = pKF2->fx * C2[0] * invz + pKF2->cx;
  const float ey = pKF2->fy * C2[1] * invz + pKF2->cy;

  struct Snap {
    std::vector<uint8_t> has_mp;
    std::vector<uint32_t> node_ids;
    std::vector<int32_t> offsets, features;
    orbfe_keyframe_view v;
    explicit Snap(KeyFrame* kf) {
      has_mp.resize(kf->N);
      for (int i = 0; i < kf->N; ++i) has_mp[i] = kf->GetMapPoint(i) ? 1 : 0;  // :751, :772
      offsets.push_back(0);
      for (DBoW2::FeatureVector::const_iterator it = kf->mFeatVec.begin(); it != kf->mFeatVec.end(); ++it) {
        node_ids.push_back(it->first);
        for (size_t k = 0; k < it->second.size(); ++k) features.push_back((int32_t)it->second[k]);
        offsets.push_back((int32_t)features.size());
      }
      v.n = kf->N;
      v.keys_un = reinterpret_cast<const orbfe_keypoint*>(kf->mvKeysUn.data());
      v.u_right = kf->mvuRight.data();
      v.desc = kf->mDescriptors.ptr(0);
      v.has_mp = has_mp.data();
      v.n_nodes = (int)node_ids.size();
      v.node_ids = node_ids.data();
      v.node_offsets = offsets.data();
      v.node_features = features.data();
      v.nlevels = (int)kf->mvScaleFactors.size();
      v.scale_factors = kf->mvScaleFactors.data();
      v.level_sigma2 = kf->mvLevelSigma2.data();
    }
  };
  Snap s1(pKF1), s2(pKF2);
  float F[9];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) F[r * 3 + c] = F12.at<float>(r, c);
  std::vector<int32_t> pairs((size_t)pKF1->N * 2);
  int32_t np = 0;
  if (orbfe_search_for_triangulation(m, &s1.v, &s2.v, F, ex, ey, bOnlyStereo ? 1 : 0, mbCheckOrientation ? 1 : 0,
                                     pairs.data(), pKF1->N, &np) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchForTriangulation: %s\n", orbfe_last_error());
    return 0;
  }
  vMatchedPairs.reserve(np);
  for (int i = 0; i < np; ++i) vMatchedPairs.push_back(std::make_pair((size_t)pairs[2 * i], (size_t)pairs[2 * i + 1]));  // :861-869
  return np;
}

}  // namespace ORB_SLAM2
