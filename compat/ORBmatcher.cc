// ORBmatcher.cc — all eleven members of ORB_SLAM2::ORBmatcher on top of the C ABI (include/orbfe.h).
// Each method snapshots the fields it reads into structure-of-arrays host buffers ON THE CALLING THREAD (the MapPoint
// / KeyFrame accessors take mutexes, src/MapPoint.cc:320-324), calls liborbfe.so on a per-thread matcher context
// (stack-local ORBmatcher objects are used concurrently by Tracking, LocalMapping and LoopClosing), and applies the
// result to the caller's objects exactly where the reference does.  No matching is computed on the CPU here.
#include "ORBmatcher.h"

#include <cmath>
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

// a DBoW2::FeatureVector as the CSR the C ABI takes
struct FeatVecCsr {
  std::vector<uint32_t> node_ids;
  std::vector<int32_t> offsets, features;
  explicit FeatVecCsr(const DBoW2::FeatureVector& fv) {
    offsets.push_back(0);
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
      node_ids.push_back(it->first);
      for (size_t k = 0; k < it->second.size(); ++k) features.push_back((int32_t)it->second[k]);
      offsets.push_back((int32_t)features.size());
    }
  }
  template <class V> void fill(V& v) const {
    v.n_nodes = (int)node_ids.size();
    v.node_ids = node_ids.data();
    v.node_offsets = offsets.data();
    v.node_features = features.data();
  }
};

// KeyFrame members read by SearchForTriangulation / SearchByBoW.  good_only: has_mp = "map point present and not
// bad" (SearchByBoW, :224-228, :600-603, :617-621) instead of "present" (SearchForTriangulation, :751).
struct KeyFrameSnap {
  std::vector<uint8_t> has_mp;
  std::vector<MapPoint*> mps;
  FeatVecCsr csr;
  orbfe_keyframe_view v;
  KeyFrameSnap(KeyFrame* kf, bool good_only) : mps(kf->GetMapPointMatches()), csr(kf->mFeatVec) {
    has_mp.resize(kf->N);
    for (int i = 0; i < kf->N; ++i) has_mp[i] = (mps[i] && !(good_only && mps[i]->isBad())) ? 1 : 0;
    v.n = kf->N;
    v.keys_un = reinterpret_cast<const orbfe_keypoint*>(kf->mvKeysUn.data());
    v.u_right = kf->mvuRight.data();
    v.desc = kf->mDescriptors.ptr(0);
    v.has_mp = has_mp.data();
    csr.fill(v);
    v.nlevels = (int)kf->mvScaleFactors.size();
    v.scale_factors = kf->mvScaleFactors.data();
    v.level_sigma2 = kf->mvLevelSigma2.data();
  }
};

// KeyFrame members read by the projection searches (grid bounds are per-keyframe ints, include/KeyFrame.h).
// occupied[i] >= 0 marks slots the method must skip (vpMatched[i] != NULL, :406).
struct KeyFrameGridSnap {
  std::vector<int32_t> occupied;
  orbfe_frame_view v;
  explicit KeyFrameGridSnap(KeyFrame* kf, const std::vector<MapPoint*>* slots = NULL) {
    occupied.assign(kf->N, -1);
    if (slots)
      for (int i = 0; i < kf->N && i < (int)slots->size(); ++i) occupied[i] = (*slots)[i] ? 1 : -1;
    v.n = kf->N;
    v.keys_un = reinterpret_cast<const orbfe_keypoint*>(kf->mvKeysUn.data());
    v.u_right = kf->mvuRight.data();
    v.desc = kf->mDescriptors.ptr(0);
    v.mp_obs = occupied.data();
    v.min_x = (float)kf->mnMinX; v.min_y = (float)kf->mnMinY; v.max_x = (float)kf->mnMaxX; v.max_y = (float)kf->mnMaxY;
    v.grid_inv_w = kf->mfGridElementWidthInv;
    v.grid_inv_h = kf->mfGridElementHeightInv;
    v.nlevels = (int)kf->mvScaleFactors.size();
    v.scale_factors = kf->mvScaleFactors.data();
  }
};

inline float raw_max_distance(MapPoint* p) {
#ifdef ORBFE_MAPPOINT_RAW_DISTANCE
  return p->GetMaxDistance();  // `float GetMaxDistance()` added next to GetMaxDistanceInvariance
#else
  // Without the accessor mfMaxDistance is recovered from GetMaxDistanceInvariance() = 1.2f * mfMaxDistance.  Two
  // neighbouring floats can share one product when 1.2x crosses a binade; the smaller is taken, which moves
  // MapPoint::PredictScale only when log(ratio)/logScaleFactor lies within an ulp of an integer.
  const float inv = p->GetMaxDistanceInvariance();
  float x = inv / 1.2f;
  while (1.2f * x > inv) x = std::nextafter(x, 0.0f);
  while (1.2f * x < inv) x = std::nextafter(x, 2.0f * inv);
  return x;
#endif
}

// map points given by position: GetWorldPos / GetNormal / distance invariance / GetDescriptor, snapshotted under the
// points' own mutexes on the calling thread
struct PointsSnap {
  std::vector<uint8_t> valid, desc;
  std::vector<float> pos, normal, max_inv, min_inv, max_d, angle;
  orbfe_points3d_view v;
  explicit PointsSnap(size_t m) : valid(m, 0), desc(m * 32), pos(m * 3), normal(m * 3), max_inv(m), min_inv(m), max_d(m),
                                  angle(m) {}
  void set(size_t i, MapPoint* p) {
    valid[i] = 1;
    const cv::Mat x = p->GetWorldPos(), nrm = p->GetNormal();
    for (int k = 0; k < 3; ++k) { pos[3 * i + k] = x.at<float>(k); normal[3 * i + k] = nrm.at<float>(k); }
    max_inv[i] = p->GetMaxDistanceInvariance();
    min_inv[i] = p->GetMinDistanceInvariance();
    max_d[i] = raw_max_distance(p);
    const cv::Mat d = p->GetDescriptor();
    std::memcpy(&desc[i * 32], d.ptr(0), 32);
  }
  const orbfe_points3d_view* view() {
    v.m = (int)valid.size();
    v.valid = valid.data(); v.world_pos = pos.data(); v.normal = normal.data();
    v.max_dist_inv = max_inv.data(); v.min_dist_inv = min_inv.data(); v.max_distance = max_d.data();
    v.desc = desc.data(); v.angle = angle.data();
    return &v;
  }
};

void fill_projection(orbfe_projection& pr, const cv::Mat& T34, bool is_sim3, const cv::Mat* ow, float fx, float fy,
                     float cx, float cy, float bf, float log_sf) {
  copy_pose(T34, pr.tcw);
  pr.is_sim3 = is_sim3 ? 1 : 0;
  pr.has_ow = ow ? 1 : 0;
  for (int k = 0; k < 3; ++k) pr.ow[k] = ow ? ow->at<float>(k) : 0.f;
  pr.fx = fx; pr.fy = fy; pr.cx = cx; pr.cy = cy; pr.bf = bf;
  pr.log_scale_factor = log_sf;
}

}  // namespace

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

// src/ORBmatcher.cc:1715-1731.  The static single-pair helper is a SCALAR function (one 32-byte pair): callers such as
// the unchanged MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:291) call it ~N^2/2 times per map point, so a
// device round trip per pair (tens of microseconds) would stall LocalMapping.  It is therefore eight popcounts inline,
// exactly the integer the device kernels compute; every data-parallel use (the matchers, the brute force, the
// batched orbfe_descriptor_distance / orbfe_compute_distinctive_descriptors) runs on the GPU.
int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
  const unsigned char* pa = a.ptr(0);
  const unsigned char* pb = b.ptr(0);
  int dist = 0;
  for (int i = 0; i < 4; ++i) {
    uint64_t x, y;
    std::memcpy(&x, pa + 8 * i, 8);
    std::memcpy(&y, pb + 8 * i, 8);
    dist += __builtin_popcountll(x ^ y);
  }
  return dist;
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

  KeyFrameSnap s1(pKF1, false), s2(pKF2, false);  // :751, :772: any map point counts
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

// ---------------------------------------------------------------------------------------------------------------
// The seven members of SURVEY.md 8(f)-1
// ---------------------------------------------------------------------------------------------------------------
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
  vpMapPointMatches = std::vector<MapPoint*>(F.N, static_cast<MapPoint*>(NULL));  // :188
  orbfe_matcher* m = t_matcher.get(sDevice);
  if (!m || pKF->N == 0 || F.N == 0) return 0;
  KeyFrameSnap s1(pKF, true);
  FeatVecCsr csr(F.mFeatVec);
  std::vector<uint8_t> none(F.N, 0);
  orbfe_keyframe_view fv;
  fv.n = F.N;
  fv.keys_un = reinterpret_cast<const orbfe_keypoint*>(F.mvKeys.data());  // F.mvKeys[bestIdxF].angle (:269)
  fv.u_right = F.mvuRight.data();
  fv.desc = F.mDescriptors.ptr(0);
  fv.has_mp = none.data();
  csr.fill(fv);
  fv.nlevels = (int)F.mvScaleFactors.size();
  fv.scale_factors = F.mvScaleFactors.data();
  fv.level_sigma2 = F.mvLevelSigma2.data();
  std::vector<int32_t> matches(F.N, -1);
  int32_t n = 0;
  if (orbfe_search_by_bow(m, &s1.v, &fv, 0, mfNNratio, mbCheckOrientation ? 1 : 0, matches.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchByBoW: %s\n", orbfe_last_error());
    return 0;
  }
  for (int f = 0; f < F.N; ++f)
    if (matches[f] >= 0) vpMapPointMatches[f] = s1.mps[matches[f]];  // :263
  return n;
}

int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
  KeyFrameSnap s1(pKF1, true), s2(pKF2, true);
  vpMatches12 = std::vector<MapPoint*>(s1.mps.size(), static_cast<MapPoint*>(NULL));  // :575
  orbfe_matcher* m = t_matcher.get(sDevice);
  if (!m || pKF1->N == 0 || pKF2->N == 0) return 0;
  std::vector<int32_t> matches(pKF1->N, -1);
  int32_t n = 0;
  if (orbfe_search_by_bow(m, &s1.v, &s2.v, 1, mfNNratio, mbCheckOrientation ? 1 : 0, matches.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchByBoW: %s\n", orbfe_last_error());
    return 0;
  }
  for (int i = 0; i < pKF1->N; ++i)
    if (matches[i] >= 0) vpMatches12[i] = s2.mps[matches[i]];  // :643
  return n;
}

int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
                                        std::vector<int>& vnMatches12, int windowSize) {
  vnMatches12 = std::vector<int>(F1.mvKeysUn.size(), -1);  // :449
  orbfe_matcher* m = t_matcher.get(sDevice);
  if (!m || F1.N == 0 || F2.N == 0) return 0;
  FrameSnapshot s1(F1), s2(F2);
  static_assert(sizeof(cv::Point2f) == 8, "cv::Point2f must be two floats");
  std::vector<int32_t> m12(F1.N, -1);
  int32_t n = 0;
  if (orbfe_search_for_initialization(m, &s1.v, &s2.v, reinterpret_cast<float*>(vbPrevMatched.data()), windowSize,
                                      mfNNratio, mbCheckOrientation ? 1 : 0, m12.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchForInitialization: %s\n", orbfe_last_error());
    return 0;
  }
  for (int i = 0; i < F1.N; ++i) vnMatches12[i] = m12[i];
  return n;
}

int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound,
                                   const float th, const int ORBdist) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();  // :1554
  if (!m || CurrentFrame.N == 0 || vpMPs.empty()) return 0;
  FrameSnapshot fs(CurrentFrame);  // mp_obs >= 0  <=>  mvpMapPoints[i] != NULL (:1609)
  PointsSnap ps(vpMPs.size());
  for (size_t i = 0; i < vpMPs.size(); ++i) {
    MapPoint* p = vpMPs[i];
    if (p && !p->isBad() && !sAlreadyFound.count(p)) ps.set(i, p);  // :1560-1562
    ps.angle[i] = pKF->mvKeysUn[i].angle;                            // :1630
  }
  orbfe_projection pr;
  fill_projection(pr, CurrentFrame.mTcw, false, NULL, Frame::fx, Frame::fy, Frame::cx, Frame::cy, CurrentFrame.mbf,
                  CurrentFrame.mfLogScaleFactor);
  std::vector<int32_t> assigned(CurrentFrame.N, -1);
  int32_t n = 0;
  if (orbfe_search_by_projection_reloc(m, &fs.v, &pr, ps.view(), th, ORBdist, mbCheckOrientation ? 1 : 0,
                                       assigned.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchByProjection: %s\n", orbfe_last_error());
    return 0;
  }
  for (int f = 0; f < CurrentFrame.N; ++f) {
    if (assigned[f] >= 0) CurrentFrame.mvpMapPoints[f] = vpMPs[assigned[f]];                  // :1625
    else if (assigned[f] == -2) CurrentFrame.mvpMapPoints[f] = static_cast<MapPoint*>(NULL);  // :1659
  }
  return n;
}

int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                                   std::vector<MapPoint*>& vpMatched, int th) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  if (!m || pKF->N == 0 || vpPoints.empty()) return 0;
  std::set<MapPoint*> spAlreadyFound(vpMatched.begin(), vpMatched.end());  // :337-338
  spAlreadyFound.erase(static_cast<MapPoint*>(NULL));
  KeyFrameGridSnap ks(pKF, &vpMatched);
  PointsSnap ps(vpPoints.size());
  for (size_t i = 0; i < vpPoints.size(); ++i) {
    MapPoint* p = vpPoints[i];
    if (!p->isBad() && !spAlreadyFound.count(p)) ps.set(i, p);  // :348
  }
  orbfe_projection pr;
  fill_projection(pr, Scw, true, NULL, pKF->fx, pKF->fy, pKF->cx, pKF->cy, 0.f, pKF->mfLogScaleFactor);
  std::vector<int32_t> assigned(pKF->N, -1);
  int32_t n = 0;
  if (orbfe_search_by_projection_sim3(m, &ks.v, &pr, ps.view(), th, assigned.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchByProjection: %s\n", orbfe_last_error());
    return 0;
  }
  for (int f = 0; f < pKF->N; ++f)
    if (assigned[f] >= 0) vpMatched[f] = vpPoints[assigned[f]];  // :427
  return n;
}

int ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  const int nMPs = (int)vpMapPoints.size();
  if (!m || pKF->N == 0 || nMPs == 0) return 0;
  KeyFrameGridSnap ks(pKF);
  PointsSnap ps(nMPs);
  for (int i = 0; i < nMPs; ++i) {
    MapPoint* p = vpMapPoints[i];
    if (p && !p->isBad() && !p->IsInKeyFrame(pKF)) ps.set(i, p);  // :900-904
  }
  float T[12];
  const cv::Mat Rcw = pKF->GetRotation(), tcw = pKF->GetTranslation(), Ow = pKF->GetCameraCenter();
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) T[r * 4 + c] = Rcw.at<float>(r, c);
    T[r * 4 + 3] = tcw.at<float>(r);
  }
  orbfe_projection pr;
  fill_projection(pr, cv::Mat(3, 4, CV_32F, T), false, &Ow, pKF->fx, pKF->fy, pKF->cx, pKF->cy, pKF->mbf,
                  pKF->mfLogScaleFactor);
  std::vector<int32_t> best(nMPs, -1);
  int32_t n = 0;
  if (orbfe_fuse(m, &ks.v, &pr, ps.view(), th, pKF->mvInvLevelSigma2.data(), best.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::Fuse: %s\n", orbfe_last_error());
    return 0;
  }
  // the map surgery of :1005-1025, in row order; a point that an earlier row made bad or put into pKF is skipped
  // exactly where the reference tests it (:903)
  int nFused = 0;
  for (int i = 0; i < nMPs; ++i) {
    if (best[i] < 0) continue;
    MapPoint* pMP = vpMapPoints[i];
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
    MapPoint* pMPinKF = pKF->GetMapPoint(best[i]);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) {
        if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
        else pMPinKF->Replace(pMP);
      }
    } else {
      pMP->AddObservation(pKF, best[i]);
      pKF->AddMapPoint(pMP, best[i]);
    }
    nFused++;
  }
  return nFused;
}

int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
                     std::vector<MapPoint*>& vpReplacePoint) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  const int nPoints = (int)vpPoints.size();
  if (!m || pKF->N == 0 || nPoints == 0) return 0;
  const std::set<MapPoint*> spAlreadyFound = pKF->GetMapPoints();  // :1049
  KeyFrameGridSnap ks(pKF);
  PointsSnap ps(nPoints);
  for (int i = 0; i < nPoints; ++i) {
    MapPoint* p = vpPoints[i];
    if (!p->isBad() && !spAlreadyFound.count(p)) ps.set(i, p);  // :1061
  }
  orbfe_projection pr;
  fill_projection(pr, Scw, true, NULL, pKF->fx, pKF->fy, pKF->cx, pKF->cy, 0.f, pKF->mfLogScaleFactor);
  std::vector<int32_t> best(nPoints, -1);
  int32_t n = 0;
  if (orbfe_fuse(m, &ks.v, &pr, ps.view(), th, NULL, best.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::Fuse: %s\n", orbfe_last_error());
    return 0;
  }
  int nFused = 0;
  for (int i = 0; i < nPoints; ++i) {  // :1137-1152
    if (best[i] < 0) continue;
    MapPoint* pMP = vpPoints[i];
    MapPoint* pMPinKF = pKF->GetMapPoint(best[i]);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) vpReplacePoint[i] = pMPinKF;
    } else {
      pMP->AddObservation(pKF, best[i]);
      pKF->AddMapPoint(pMP, best[i]);
    }
    nFused++;
  }
  return nFused;
}

int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
                             const cv::Mat& R12, const cv::Mat& t12, const float th) {
  orbfe_matcher* m = t_matcher.get(sDevice);
  const std::vector<MapPoint*> vpMapPoints1 = pKF1->GetMapPointMatches(), vpMapPoints2 = pKF2->GetMapPointMatches();
  const int N1 = (int)vpMapPoints1.size(), N2 = (int)vpMapPoints2.size();
  if (!m || N1 == 0 || N2 == 0) return 0;
  std::vector<bool> vbAlreadyMatched1(N1, false), vbAlreadyMatched2(N2, false);  // :1185-1198
  for (int i = 0; i < N1; ++i) {
    MapPoint* pMP = vpMatches12[i];
    if (pMP) {
      vbAlreadyMatched1[i] = true;
      const int idx2 = pMP->GetIndexInKeyFrame(pKF2);
      if (idx2 >= 0 && idx2 < N2) vbAlreadyMatched2[idx2] = true;
    }
  }
  PointsSnap p1(N1), p2(N2);
  for (int i = 0; i < N1; ++i) {
    MapPoint* p = vpMapPoints1[i];
    if (p && !vbAlreadyMatched1[i] && !p->isBad()) p1.set(i, p);  // :1206-1212
  }
  for (int i = 0; i < N2; ++i) {
    MapPoint* p = vpMapPoints2[i];
    if (p && !vbAlreadyMatched2[i] && !p->isBad()) p2.set(i, p);  // :1286-1292
  }
  KeyFrameGridSnap k1(pKF1), k2(pKF2);
  orbfe_sim3_pair sp;
  const cv::Mat R1w = pKF1->GetRotation(), t1w = pKF1->GetTranslation(), R2w = pKF2->GetRotation(),
                t2w = pKF2->GetTranslation();
  for (int r = 0; r < 3; ++r) {
    for (int c = 0; c < 3; ++c) {
      sp.t1w[r * 4 + c] = R1w.at<float>(r, c);
      sp.t2w[r * 4 + c] = R2w.at<float>(r, c);
      sp.r12[r * 3 + c] = R12.at<float>(r, c);
    }
    sp.t1w[r * 4 + 3] = t1w.at<float>(r);
    sp.t2w[r * 4 + 3] = t2w.at<float>(r);
    sp.t12[r] = t12.at<float>(r);
  }
  sp.s12 = s12;
  sp.fx = pKF1->fx; sp.fy = pKF1->fy; sp.cx = pKF1->cx; sp.cy = pKF1->cy;  // :1161-1164
  sp.log_scale_factor1 = pKF1->mfLogScaleFactor;
  sp.log_scale_factor2 = pKF2->mfLogScaleFactor;
  std::vector<int32_t> m12(N1, -1);
  int32_t n = 0;
  if (orbfe_search_by_sim3(m, &k1.v, &k2.v, p1.view(), p2.view(), &sp, th, m12.data(), &n) != ORBFE_OK) {
    std::fprintf(stderr, "ORBmatcher::SearchBySim3: %s\n", orbfe_last_error());
    return 0;
  }
  for (int i = 0; i < N1; ++i)
    if (m12[i] >= 0) vpMatches12[i] = vpMapPoints2[m12[i]];  // :1375
  return n;
}

}  // namespace ORB_SLAM2
