// ref_driver_slam.cc — C API around the UNMODIFIED reference Frame / ORBmatcher / MapPoint / KeyFrame / Map sources
// (compiled in place from /root/reference/src, see oracle/Makefile).  TEST INFRASTRUCTURE ONLY.
//
// The entry points take the same struct-of-array views as the oracle (include/orbfe_types.h) and the C ABI, build the
// reference's own objects from them (Frame, KeyFrame, MapPoint with exactly the fields the views describe), call the
// reference's method, and translate what the method did to its objects back into the oracle's output convention.
// Nothing of the reference's text is copied: only constructors, public methods and member assignments.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <thread>
#include <vector>

#include "../include/orbfe_types.h"
#include <opencv2/core/core.hpp>  // every standard header the reference's headers pull in, BEFORE the access override

// the driver fills protected / private members (mWorldPos, mDescriptor, Ow, AssignFeaturesToGrid, ...) directly; the
// reference's own translation units are compiled without this
#define private public
#define protected public
#include "Frame.h"
#include "KeyFrame.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBmatcher.h"
#include "Converter.h"
#undef private
#undef protected

#include "ref_driver.h"

using namespace ORB_SLAM2;

namespace ORB_SLAM2 {
// src/Converter.cc needs Eigen/g2o (absent); the one member the compiled units call is a list of row headers
std::vector<cv::Mat> Converter::toDescriptorVector(const cv::Mat& Descriptors) {
  std::vector<cv::Mat> v;
  v.reserve(Descriptors.rows);
  for (int j = 0; j < Descriptors.rows; ++j) v.push_back(Descriptors.row(j));
  return v;
}
// src/KeyFrameDatabase.cc is not compiled (inverted-file bookkeeping, off the path); KeyFrame::SetBadFlag links to it
void KeyFrameDatabase::erase(KeyFrame*) {}
}  // namespace ORB_SLAM2

namespace {

cv::Mat mat_from(const float* p, int rows, int cols) {
  cv::Mat m(rows, cols, CV_32F);
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c) m.at<float>(r, c) = p[r * cols + c];
  return m;
}
cv::Mat pose44(const float* t34) {  // rows 0..2 given, last row 0 0 0 1
  cv::Mat T = cv::Mat::eye(4, 4, CV_32F);
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) T.at<float>(r, c) = t34[r * 4 + c];
  return T;
}
cv::Mat desc_mat(const uint8_t* d, int n) {
  cv::Mat m(n, 32, CV_8U);
  if (n) std::memcpy(m.data, d, (size_t)n * 32);
  return m;
}

// owns every object of one call
struct World {
  Map map;
  std::unique_ptr<Frame> dummy_frame;
  KeyFrame* dummy_kf = nullptr;
  std::vector<MapPoint*> points;
  std::vector<KeyFrame*> kfs;
  std::vector<Frame*> frames;

  World() {
    dummy_frame.reset(new Frame());
    Frame& F = *dummy_frame;
    F.N = 0;
    F.mnId = 0;
    F.mTimeStamp = 0;
    F.mbf = 0; F.mb = 0; F.mThDepth = 0;
    F.mnScaleLevels = 1; F.mfScaleFactor = 1.2f; F.mfLogScaleFactor = std::log(1.2f);
    F.mpORBvocabulary = nullptr;
    F.mTcw = cv::Mat::eye(4, 4, CV_32F);
    dummy_kf = new KeyFrame(F, &map, nullptr);
  }
  ~World() {
    for (MapPoint* p : points) delete p;
    for (KeyFrame* k : kfs) delete k;
    for (Frame* f : frames) delete f;
    delete dummy_kf;
  }
  MapPoint* point(const float* pos3, int obs, const uint8_t* desc32) {
    const float zero[3] = {0, 0, 0};
    MapPoint* p = new MapPoint(mat_from(pos3 ? pos3 : zero, 3, 1), dummy_kf, &map);
    p->nObs = obs;
    if (desc32) p->mDescriptor = desc_mat(desc32, 1);
    points.push_back(p);
    return p;
  }
  // a Frame with the members of an orbfe_frame_view (statics included), grid assigned by the reference's own code
  Frame* frame(const orbfe_frame_view* v) {
    Frame* F = new Frame();
    frames.push_back(F);
    F->mnId = Frame::nNextId++;
    F->mpORBvocabulary = nullptr;
    F->mpReferenceKF = nullptr;
    F->mTimeStamp = 0;
    F->N = v->n;
    F->mvKeysUn.resize(v->n);
    if (v->n) std::memcpy(F->mvKeysUn.data(), v->keys_un, (size_t)v->n * sizeof(cv::KeyPoint));
    F->mvKeys = F->mvKeysUn;
    F->mvuRight.assign(v->n, -1.f);
    if (v->u_right) F->mvuRight.assign(v->u_right, v->u_right + v->n);
    F->mvDepth.assign(v->n, -1.f);
    F->mDescriptors = desc_mat(v->desc, v->n);
    F->mvpMapPoints.assign(v->n, nullptr);
    if (v->mp_obs)
      for (int i = 0; i < v->n; ++i)
        if (v->mp_obs[i] >= 0) F->mvpMapPoints[i] = point(nullptr, v->mp_obs[i], nullptr);
    F->mvbOutlier.assign(v->n, false);
    Frame::mnMinX = v->min_x; Frame::mnMinY = v->min_y; Frame::mnMaxX = v->max_x; Frame::mnMaxY = v->max_y;
    Frame::mfGridElementWidthInv = v->grid_inv_w;
    Frame::mfGridElementHeightInv = v->grid_inv_h;
    F->mnScaleLevels = v->nlevels;
    F->mvScaleFactors.assign(v->scale_factors, v->scale_factors + v->nlevels);
    F->mvInvScaleFactors.resize(v->nlevels);
    F->mvLevelSigma2.resize(v->nlevels);
    F->mvInvLevelSigma2.resize(v->nlevels);
    for (int i = 0; i < v->nlevels; ++i) {
      F->mvInvScaleFactors[i] = 1.0f / F->mvScaleFactors[i];
      F->mvLevelSigma2[i] = F->mvScaleFactors[i] * F->mvScaleFactors[i];
      F->mvInvLevelSigma2[i] = 1.0f / F->mvLevelSigma2[i];
    }
    F->mfScaleFactor = v->nlevels > 1 ? F->mvScaleFactors[1] : 1.2f;
    F->mfLogScaleFactor = std::log(F->mfScaleFactor);
    F->mbf = 0; F->mb = 0; F->mThDepth = 0;
    F->mTcw = cv::Mat::eye(4, 4, CV_32F);
    F->AssignFeaturesToGrid();  // src/Frame.cc:243-258
    return F;
  }
};

}  // namespace

extern "C" {

// ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1715-1731), the literal SWAR code
int ref_descriptor_distance(const uint8_t* a, const uint8_t* b) {
  cv::Mat ma(1, 32, CV_8U, const_cast<uint8_t*>(a)), mb(1, 32, CV_8U, const_cast<uint8_t*>(b));
  return ORBmatcher::DescriptorDistance(ma, mb);
}

// Frame::ComputeStereoMatches (src/Frame.cc:496-686) on the results the two extractor handles hold from their last
// ref_extract calls.  `mb` is given explicitly (the reference's constructor reads it before it is set, :534 vs :127).
int ref_stereo_matches(ref_extractor* L, ref_extractor* R, float mbf, float mb, float* u_right, float* depth) {
  RefCallScope scope;
  std::unique_ptr<Frame> Fp(new Frame());
  Frame& F = *Fp;
  F.mpORBextractorLeft = L->ex;
  F.mpORBextractorRight = R->ex;
  F.mvKeys = L->kps;
  F.mvKeysRight = R->kps;
  F.mDescriptors = L->desc;
  F.mDescriptorsRight = R->desc;
  F.N = (int)F.mvKeys.size();
  F.mvScaleFactors = L->ex->GetScaleFactors();
  F.mvInvScaleFactors = L->ex->GetInverseScaleFactors();
  F.mbf = mbf;
  F.mb = mb;
  F.ComputeStereoMatches();
  int n = 0;
  for (int i = 0; i < F.N; ++i) {
    u_right[i] = F.mvuRight[i];
    depth[i] = F.mvDepth[i];
    n += F.mvuRight[i] >= 0;
  }
  return n;
}

// the stereo Frame constructor's extraction step as the reference runs it: two threads (src/Frame.cc:89-92)
void ref_extract_pair_threads(ref_extractor* L, ref_extractor* R, const uint8_t* left, const uint8_t* right, int width,
                              int height, int stride) {
  RefCallScope scope;
  std::unique_ptr<Frame> Fp(new Frame());
  Frame& F = *Fp;
  F.mpORBextractorLeft = L->ex;
  F.mpORBextractorRight = R->ex;
  cv::Mat il(height, width, CV_8UC1, const_cast<uint8_t*>(left), (size_t)stride);
  cv::Mat ir(height, width, CV_8UC1, const_cast<uint8_t*>(right), (size_t)stride);
  std::thread tl(&Frame::ExtractORB, &F, 0, il);
  std::thread tr(&Frame::ExtractORB, &F, 1, ir);
  tl.join();
  tr.join();
  L->kps = F.mvKeys; L->desc = F.mDescriptors;
  R->kps = F.mvKeysRight; R->desc = F.mDescriptorsRight;
}
int ref_extractor_results(ref_extractor* r, orbfe_keypoint* kps, uint8_t* desc, int capacity) {
  const int n = (int)r->kps.size();
  for (int i = 0; i < n && i < capacity; ++i) {
    std::memcpy(&kps[i], &r->kps[i], sizeof(cv::KeyPoint));
    std::memcpy(desc + (size_t)i * 32, r->desc.ptr(i), 32);
  }
  return n;
}

// ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th) (src/ORBmatcher.cc:46-132)
int ref_search_by_projection_map(const orbfe_frame_view* fv, const orbfe_mappoints_view* mv, float th, float nnratio,
                                 int32_t* assigned) {
  RefCallScope scope;
  World W;
  Frame* F = W.frame(fv);
  std::vector<MapPoint*> before = F->mvpMapPoints;
  std::vector<MapPoint*> mps(mv->m);
  std::map<MapPoint*, int> index;
  for (int i = 0; i < mv->m; ++i) {
    MapPoint* p = W.point(nullptr, mv->obs[i], mv->desc + (size_t)i * 32);
    p->mbTrackInView = mv->in_view[i] != 0;
    p->mTrackProjX = mv->proj_x[i];
    p->mTrackProjY = mv->proj_y[i];
    p->mTrackProjXR = mv->proj_xr[i];
    p->mnTrackScaleLevel = mv->level[i];
    p->mTrackViewCos = mv->view_cos[i];
    mps[i] = p;
    index[p] = i;
  }
  ORBmatcher matcher(nnratio);
  const int n = matcher.SearchByProjection(*F, mps, th);
  for (int f = 0; f < fv->n; ++f) assigned[f] = F->mvpMapPoints[f] != before[f] ? index.at(F->mvpMapPoints[f]) : -1;
  return n;
}

// ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) (src/ORBmatcher.cc:1392-1538).
// assigned[f]: row of `last`; -1 = Cur.mvpMapPoints[f] is what it was (or NULL again after the rotation check)
int ref_search_by_projection_last(const orbfe_frame_view* cur, const orbfe_camera_pose* cam,
                                  const orbfe_lastframe_view* last, float th, int mono, int check_ori,
                                  int32_t* assigned) {
  RefCallScope scope;
  World W;
  Frame* C = W.frame(cur);
  Frame::fx = cam->fx; Frame::fy = cam->fy; Frame::cx = cam->cx; Frame::cy = cam->cy;
  Frame::invfx = 1.0f / cam->fx; Frame::invfy = 1.0f / cam->fy;
  C->mbf = cam->mbf;
  C->mb = cam->mb;
  C->mTcw = pose44(cam->tcw);
  std::vector<MapPoint*> before = C->mvpMapPoints;
  std::unique_ptr<Frame> Lp(new Frame());
  Frame& L = *Lp;
  L.N = last->n;
  L.mTcw = pose44(last->tcw);
  L.mvKeys.resize(last->n);
  L.mvKeysUn.resize(last->n);
  L.mvpMapPoints.assign(last->n, nullptr);
  L.mvbOutlier.assign(last->n, false);
  std::map<MapPoint*, int> index;
  for (int i = 0; i < last->n; ++i) {
    L.mvKeys[i].octave = last->octave[i];
    L.mvKeysUn[i].octave = last->octave[i];
    L.mvKeysUn[i].angle = last->angle[i];
    if (!last->valid[i]) continue;
    MapPoint* p = W.point(last->world_pos + (size_t)i * 3, last->obs[i], last->desc + (size_t)i * 32);
    L.mvpMapPoints[i] = p;
    index[p] = i;
  }
  ORBmatcher matcher(0.9f, check_ori != 0);
  const int n = matcher.SearchByProjection(*C, L, th, mono != 0);
  for (int f = 0; f < cur->n; ++f) {
    MapPoint* p = C->mvpMapPoints[f];
    assigned[f] = (p != before[f] && p != nullptr) ? index.at(p) : (p == nullptr && before[f] != nullptr ? -2 : -1);
  }
  return n;
}

}  // extern "C"

namespace {
// a KeyFrame with the members of an orbfe_keyframe_view; pose = identity, fx = fy = 1, cx = cy = 0
KeyFrame* make_keyframe(World& W, const orbfe_keyframe_view* v) {
  orbfe_frame_view fv;
  std::memset(&fv, 0, sizeof(fv));
  fv.n = v->n; fv.keys_un = v->keys_un; fv.u_right = v->u_right; fv.desc = v->desc; fv.mp_obs = nullptr;
  fv.min_x = 0; fv.min_y = 0; fv.max_x = 1; fv.max_y = 1; fv.grid_inv_w = 0; fv.grid_inv_h = 0;
  fv.nlevels = v->nlevels; fv.scale_factors = v->scale_factors;
  Frame* F = W.frame(&fv);
  if (v->level_sigma2) F->mvLevelSigma2.assign(v->level_sigma2, v->level_sigma2 + v->nlevels);
  for (int i = 0; i < v->n; ++i)
    if (v->has_mp && v->has_mp[i]) F->mvpMapPoints[i] = W.point(nullptr, 1, nullptr);
  for (int a = 0; a < v->n_nodes; ++a) {
    std::vector<unsigned int>& feats = F->mFeatVec[v->node_ids[a]];
    for (int q = v->node_offsets[a]; q < v->node_offsets[a + 1]; ++q) feats.push_back((unsigned)v->node_features[q]);
  }
  Frame::fx = 1; Frame::fy = 1; Frame::cx = 0; Frame::cy = 0; Frame::invfx = 1; Frame::invfy = 1;
  KeyFrame* K = new KeyFrame(*F, &W.map, nullptr);
  W.kfs.push_back(K);
  return K;
}
}  // namespace

extern "C" {

// ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:702-872) + CheckDistEpipolarLine (:147-167).  The epipole
// (ex, ey) the method derives from the two poses (:709-716) is realised exactly: KF2 has the identity pose and unit
// intrinsics, KF1's camera centre is (ex, ey, 1).
int ref_search_for_triangulation(const orbfe_keyframe_view* k1, const orbfe_keyframe_view* k2, const float* F12,
                                 float ex, float ey, int only_stereo, int check_ori, int32_t* pairs, int cap) {
  RefCallScope scope;
  World W;
  KeyFrame* K1 = make_keyframe(W, k1);
  KeyFrame* K2 = make_keyframe(W, k2);
  const float c[3] = {ex, ey, 1.f};
  K1->Ow = mat_from(c, 3, 1);
  ORBmatcher matcher(0.6f, check_ori != 0);
  std::vector<std::pair<size_t, size_t> > out;
  const int n = matcher.SearchForTriangulation(K1, K2, mat_from(F12, 3, 3), out, only_stereo != 0);
  for (int i = 0; i < (int)out.size() && i < cap; ++i) {
    pairs[2 * i] = (int32_t)out[i].first;
    pairs[2 * i + 1] = (int32_t)out[i].second;
  }
  (void)n;
  return (int)out.size();
}

}  // extern "C"
