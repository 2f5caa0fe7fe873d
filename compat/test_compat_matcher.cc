// test_compat_matcher.cc — drives the drop-in ORB_SLAM2::ORBmatcher (compat/ORBmatcher.cc) on cases written by
// tests/test_gpu_compat.py as raw arrays in a directory, and writes what the reference's callers would observe
// (Frame::mvpMapPoints after the call, the return value, vMatchedPairs).  Usage: test_compat_matcher <mode> <dir>
#include <cstdio>
#include <cstdlib>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "ORBVocabulary.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM2;

float Frame::fx, Frame::fy, Frame::cx, Frame::cy;
float Frame::mnMinX, Frame::mnMaxX, Frame::mnMinY, Frame::mnMaxY;
float Frame::mfGridElementWidthInv, Frame::mfGridElementHeightInv;

template <class T>
static std::vector<T> rd(const std::string& dir, const char* name) {
  const std::string p = dir + "/" + name + ".bin";
  FILE* f = std::fopen(p.c_str(), "rb");
  if (!f) { std::fprintf(stderr, "cannot open %s\n", p.c_str()); std::exit(2); }
  std::fseek(f, 0, SEEK_END);
  const long n = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  std::vector<T> v(n / sizeof(T));
  if (n && std::fread(v.data(), 1, n, f) != (size_t)n) std::exit(2);
  std::fclose(f);
  return v;
}

static cv::Mat mat_f(const float* p, int r, int c) {
  cv::Mat m(r, c, CV_32F);
  for (int i = 0; i < r; ++i)
    for (int j = 0; j < c; ++j) m.at<float>(i, j) = p[i * c + j];
  return m;
}

static cv::Mat desc_row(const std::vector<uint8_t>& d, int i) {
  cv::Mat m(1, 32, CV_8U);
  for (int k = 0; k < 32; ++k) m.ptr(0)[k] = d[(size_t)i * 32 + k];
  return m;
}

// Frame from keys_un / u_right / desc / mp_obs / bounds / scale_factors; `owned` keeps the stand-in map points alive
static void load_frame(const std::string& dir, Frame& F, std::vector<MapPoint>& owned) {
  std::vector<cv::KeyPoint> k = rd<cv::KeyPoint>(dir, "keys_un");
  F.N = (int)k.size();
  F.mvKeysUn = k;
  F.mvKeys = k;
  F.mvuRight = rd<float>(dir, "u_right");
  std::vector<uint8_t> d = rd<uint8_t>(dir, "desc");
  F.mDescriptors.create(F.N, 32, CV_8U);
  std::copy(d.begin(), d.end(), F.mDescriptors.ptr(0));
  std::vector<int32_t> obs = rd<int32_t>(dir, "mp_obs");
  owned.resize(F.N);
  F.mvpMapPoints.assign(F.N, (MapPoint*)NULL);
  for (int i = 0; i < F.N; ++i)
    if (obs[i] >= 0) { owned[i].nObs = obs[i]; F.mvpMapPoints[i] = &owned[i]; }
  F.mvScaleFactors = rd<float>(dir, "scale_factors");
  std::vector<float> b = rd<float>(dir, "bounds");  // min_x, min_y, max_x, max_y
  Frame::mnMinX = b[0]; Frame::mnMinY = b[1]; Frame::mnMaxX = b[2]; Frame::mnMaxY = b[3];
  Frame::mfGridElementWidthInv = 64.0f / (b[2] - b[0]);   // src/Frame.cc:96-97
  Frame::mfGridElementHeightInv = 48.0f / (b[3] - b[1]);
}

static void write_i32(const std::string& dir, const char* name, const std::vector<int32_t>& v) {
  FILE* f = std::fopen((dir + "/" + name + ".bin").c_str(), "wb");
  std::fwrite(v.data(), 4, v.size(), f);
  std::fclose(f);
}

// ---- loaders for the seven members of SURVEY.md 8(f)-1 ----
static void set_desc(MapPoint& p, const std::vector<uint8_t>& d, int i) { p.mDescriptor = desc_row(d, i); }

// KeyFrame with the grid members from a frame dump (keys_un / u_right / desc / mp_obs / bounds / scale_factors);
// slots with mp_obs >= 0 get a stand-in map point with that many observations
static void load_keyframe(const std::string& dir, KeyFrame& K, std::vector<MapPoint>& owned) {
  std::vector<cv::KeyPoint> k = rd<cv::KeyPoint>(dir, "keys_un");
  K.N = (int)k.size();
  K.mvKeysUn = k;
  K.mvuRight = rd<float>(dir, "u_right");
  std::vector<uint8_t> d = rd<uint8_t>(dir, "desc");
  K.mDescriptors.create(K.N, 32, CV_8U);
  std::copy(d.begin(), d.end(), K.mDescriptors.ptr(0));
  std::vector<int32_t> obs = rd<int32_t>(dir, "mp_obs");
  owned.resize(K.N);
  K.mvpMapPoints.assign(K.N, (MapPoint*)NULL);
  for (int i = 0; i < K.N; ++i)
    if (obs[i] >= 0) { owned[i].nObs = obs[i]; K.mvpMapPoints[i] = &owned[i]; }
  K.mvScaleFactors = rd<float>(dir, "scale_factors");
  K.mnScaleLevels = (int)K.mvScaleFactors.size();
  for (size_t l = 0; l < K.mvScaleFactors.size(); ++l) {
    const float s2 = K.mvScaleFactors[l] * K.mvScaleFactors[l];
    K.mvLevelSigma2.push_back(s2);
    K.mvInvLevelSigma2.push_back(1.0f / s2);
  }
  std::vector<float> b = rd<float>(dir, "bounds");
  K.mnMinX = (int)b[0]; K.mnMinY = (int)b[1]; K.mnMaxX = (int)b[2]; K.mnMaxY = (int)b[3];
  K.mfGridElementWidthInv = 64.0f / (float)(K.mnMaxX - K.mnMinX);
  K.mfGridElementHeightInv = 48.0f / (float)(K.mnMaxY - K.mnMinY);
}

// points dump: valid / world_pos / normal / max_distance / min_distance / pt_desc.  invalid -> isBad()
static void load_points(const std::string& dir, std::vector<MapPoint>& pts, std::vector<uint8_t>& valid) {
  valid = rd<uint8_t>(dir, "valid");
  std::vector<float> pos = rd<float>(dir, "world_pos"), nrm = rd<float>(dir, "normal"), mx = rd<float>(dir, "max_distance"),
                     mn = rd<float>(dir, "min_distance");
  std::vector<uint8_t> d = rd<uint8_t>(dir, "pt_desc");
  pts.resize(valid.size());
  for (size_t i = 0; i < valid.size(); ++i) {
    pts[i].mbBad = !valid[i];
    pts[i].mWorldPos = mat_f(&pos[3 * i], 3, 1);
    pts[i].mNormalVector = nrm.empty() ? mat_f(&pos[3 * i], 3, 1) : mat_f(&nrm[3 * i], 3, 1);
    pts[i].mfMaxDistance = mx[i];
    pts[i].mfMinDistance = mn[i];
    pts[i].nObs = 1 + (int)(i % 5);
    set_desc(pts[i], d, (int)i);
  }
}

// proj dump: tcw[12], fx, fy, cx, cy, bf, log_scale_factor
struct Proj { float T[16]; float fx, fy, cx, cy, bf, lsf; };
static Proj load_proj(const std::string& dir) {
  std::vector<float> v = rd<float>(dir, "proj");
  Proj P;
  for (int i = 0; i < 12; ++i) P.T[i] = v[i];
  P.T[12] = P.T[13] = P.T[14] = 0.f; P.T[15] = 1.f;
  P.fx = v[12]; P.fy = v[13]; P.cx = v[14]; P.cy = v[15]; P.bf = v[16]; P.lsf = v[17];
  return P;
}

static void load_bow_keyframe(const std::string& sub, KeyFrame& K, std::vector<MapPoint>& owned) {
  std::vector<cv::KeyPoint> k = rd<cv::KeyPoint>(sub, "keys_un");
  K.N = (int)k.size();
  K.mvKeysUn = k;
  K.mvuRight = rd<float>(sub, "u_right");
  std::vector<uint8_t> d = rd<uint8_t>(sub, "desc"), has = rd<uint8_t>(sub, "has_mp");
  K.mDescriptors.create(K.N, 32, CV_8U);
  std::copy(d.begin(), d.end(), K.mDescriptors.ptr(0));
  owned.resize(K.N);
  K.mvpMapPoints.assign(K.N, (MapPoint*)NULL);
  for (int i = 0; i < K.N; ++i)
    if (has[i]) K.mvpMapPoints[i] = &owned[i];
  std::vector<uint32_t> ids = rd<uint32_t>(sub, "node_ids");
  std::vector<int32_t> off = rd<int32_t>(sub, "node_offsets"), feat = rd<int32_t>(sub, "node_features");
  for (size_t nd = 0; nd < ids.size(); ++nd)
    for (int p = off[nd]; p < off[nd + 1]; ++p) K.mFeatVec[ids[nd]].push_back((unsigned)feat[p]);
  K.mvScaleFactors = rd<float>(sub, "scale_factors");
  K.mvLevelSigma2 = rd<float>(sub, "level_sigma2");
}

template <class T>
static int index_in(const std::vector<T>& owned, const T* p) {
  return (p && !owned.empty() && p >= &owned[0] && p < &owned[0] + owned.size()) ? (int)(p - &owned[0]) : -1;
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  const std::string mode = argv[1], dir = argv[2];
  if (mode == "map" || mode == "map_threads") {
    // one SearchByProjection(F, MapPoints) from scratch: own Frame, own MapPoints, own (stack-local) matcher
    auto run_once = [&dir]() {
      Frame F;
      std::vector<MapPoint> owned;
      load_frame(dir, F, owned);
      std::vector<float> px = rd<float>(dir, "proj_x"), py = rd<float>(dir, "proj_y"), pxr = rd<float>(dir, "proj_xr"),
                         vc = rd<float>(dir, "view_cos");
      std::vector<int32_t> lvl = rd<int32_t>(dir, "level"), obs = rd<int32_t>(dir, "obs");
      std::vector<uint8_t> inv = rd<uint8_t>(dir, "in_view"), d = rd<uint8_t>(dir, "mp_desc");
      const int M = (int)px.size();
      std::vector<MapPoint> mps(M);
      std::vector<MapPoint*> vp(M);
      for (int i = 0; i < M; ++i) {
        mps[i].mTrackProjX = px[i]; mps[i].mTrackProjY = py[i]; mps[i].mTrackProjXR = pxr[i];
        mps[i].mnTrackScaleLevel = lvl[i]; mps[i].mTrackViewCos = vc[i]; mps[i].mbTrackInView = inv[i] != 0;
        mps[i].nObs = obs[i]; mps[i].mDescriptor = desc_row(d, i);
        vp[i] = &mps[i];
      }
      std::vector<float> par = rd<float>(dir, "params");  // th, nnratio
      ORBmatcher matcher(par[1], true);
      const int n = matcher.SearchByProjection(F, vp, par[0]);
      std::vector<int32_t> out(F.N + 1, -1);
      for (int f = 0; f < F.N; ++f)
        if (F.mvpMapPoints[f] >= &mps[0] && F.mvpMapPoints[f] < &mps[0] + M) out[f] = (int32_t)(F.mvpMapPoints[f] - &mps[0]);
      out[F.N] = n;
      return out;
    };
    const std::vector<int32_t> out = run_once();
    if (mode == "map_threads") {
      // Tracking, LocalMapping and LoopClosing use their own matcher objects at the same time:
      // three threads repeat the call concurrently; every result must equal the one computed alone
      int bad[3] = {0, 0, 0};
      std::vector<std::thread> ts;
      for (int t = 0; t < 3; ++t)
        ts.emplace_back([&, t] {
          for (int r = 0; r < 6; ++r)
            if (run_once() != out) ++bad[t];
        });
      for (std::thread& t : ts) t.join();
      if (bad[0] + bad[1] + bad[2]) {
        std::fprintf(stderr, "concurrent results differ: %d %d %d of 6\n", bad[0], bad[1], bad[2]);
        return 3;
      }
    }
    write_i32(dir, "result", out);
  } else if (mode == "last") {
    Frame Cur, Last;
    std::vector<MapPoint> owned;
    load_frame(dir, Cur, owned);
    const std::vector<int32_t> cur_obs = rd<int32_t>(dir, "mp_obs");
    std::vector<float> cam = rd<float>(dir, "cam");  // tcw[12], fx, fy, cx, cy, mbf, mb
    float T[16] = {0};
    for (int i = 0; i < 12; ++i) T[i] = cam[i];
    T[15] = 1.f;
    Cur.mTcw = mat_f(T, 4, 4);
    Frame::fx = cam[12]; Frame::fy = cam[13]; Frame::cx = cam[14]; Frame::cy = cam[15];
    Cur.mbf = cam[16]; Cur.mb = cam[17];
    std::vector<uint8_t> valid = rd<uint8_t>(dir, "valid"), d = rd<uint8_t>(dir, "last_desc");
    std::vector<float> pos = rd<float>(dir, "world_pos"), ang = rd<float>(dir, "angle"), ltcw = rd<float>(dir, "last_tcw");
    std::vector<int32_t> obs = rd<int32_t>(dir, "obs"), oct = rd<int32_t>(dir, "octave");
    const int n = (int)valid.size();
    Last.N = n;
    Last.mvKeys.resize(n);
    Last.mvKeysUn.resize(n);
    Last.mvbOutlier.assign(n, false);
    Last.mvpMapPoints.assign(n, (MapPoint*)NULL);
    std::vector<MapPoint> lmp(n);
    for (int i = 0; i < n; ++i) {
      Last.mvKeys[i].octave = oct[i];
      Last.mvKeysUn[i].angle = ang[i];
      if (valid[i]) {
        lmp[i].nObs = obs[i];
        lmp[i].mDescriptor = desc_row(d, i);
        lmp[i].mWorldPos = mat_f(&pos[3 * i], 3, 1);
        Last.mvpMapPoints[i] = &lmp[i];
      }
    }
    for (int i = 0; i < 12; ++i) T[i] = ltcw[i];
    Last.mTcw = mat_f(T, 4, 4);
    std::vector<float> par = rd<float>(dir, "params");  // th, bMono, checkOri
    ORBmatcher matcher(0.9f, par[2] != 0.f);
    const int nm = matcher.SearchByProjection(Cur, Last, par[0], par[1] != 0.f);
    std::vector<int32_t> out(Cur.N + 1, -1);
    for (int f = 0; f < Cur.N; ++f) {
      MapPoint* p = Cur.mvpMapPoints[f];
      if (p >= &lmp[0] && p < &lmp[0] + n) out[f] = (int32_t)(p - &lmp[0]);
      else if (p == NULL && cur_obs[f] >= 0) out[f] = -2;  // had a map point before the call, NULL after (:1527)
    }
    out[Cur.N] = nm;
    write_i32(dir, "result", out);
  } else if (mode == "tri") {
    KeyFrame kf[2];
    std::vector<std::vector<MapPoint> > owned(2);
    for (int q = 0; q < 2; ++q) {
      const std::string sub = dir + (q ? "/kf2" : "/kf1");
      std::vector<cv::KeyPoint> k = rd<cv::KeyPoint>(sub, "keys_un");
      kf[q].N = (int)k.size();
      kf[q].mvKeysUn = k;
      kf[q].mvuRight = rd<float>(sub, "u_right");
      std::vector<uint8_t> d = rd<uint8_t>(sub, "desc"), has = rd<uint8_t>(sub, "has_mp");
      kf[q].mDescriptors.create(kf[q].N, 32, CV_8U);
      std::copy(d.begin(), d.end(), kf[q].mDescriptors.ptr(0));
      owned[q].resize(kf[q].N);
      kf[q].mvpMapPoints.assign(kf[q].N, (MapPoint*)NULL);
      for (int i = 0; i < kf[q].N; ++i)
        if (has[i]) kf[q].mvpMapPoints[i] = &owned[q][i];
      std::vector<uint32_t> ids = rd<uint32_t>(sub, "node_ids");
      std::vector<int32_t> off = rd<int32_t>(sub, "node_offsets"), feat = rd<int32_t>(sub, "node_features");
      for (size_t nd = 0; nd < ids.size(); ++nd)
        for (int p = off[nd]; p < off[nd + 1]; ++p) kf[q].mFeatVec[ids[nd]].push_back((unsigned)feat[p]);
      kf[q].mvScaleFactors = rd<float>(sub, "scale_factors");
      kf[q].mvLevelSigma2 = rd<float>(sub, "level_sigma2");
    }
    // geometry chosen so that the epipole computed at src/ORBmatcher.cc:708-716 is the (ex, ey) of the case
    std::vector<float> g = rd<float>(dir, "geom");  // F12[9], ex, ey, bOnlyStereo, checkOri
    const float I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, z3[3] = {0, 0, 0};
    kf[0].Ow = mat_f(z3, 3, 1);
    kf[1].Rcw = mat_f(I3, 3, 3);
    const float t[3] = {g[9], g[10], 1.0f};  // C2 = (ex, ey, 1) with fx = fy = 1, cx = cy = 0
    kf[1].tcw = mat_f(t, 3, 1);
    kf[1].fx = kf[1].fy = 1.f;
    kf[1].cx = kf[1].cy = 0.f;
    ORBmatcher matcher(0.6f, g[12] != 0.f);
    std::vector<std::pair<size_t, size_t> > pairs;
    const int n = matcher.SearchForTriangulation(&kf[0], &kf[1], mat_f(g.data(), 3, 3), pairs, g[11] != 0.f);
    std::vector<int32_t> out;
    out.push_back(n);
    for (size_t i = 0; i < pairs.size(); ++i) { out.push_back((int32_t)pairs[i].first); out.push_back((int32_t)pairs[i].second); }
    write_i32(dir, "result", out);
  } else if (mode == "dist") {
    std::vector<uint8_t> a = rd<uint8_t>(dir, "a"), b = rd<uint8_t>(dir, "b");
    std::vector<int32_t> out;
    for (size_t i = 0; i + 32 <= a.size(); i += 32)
      out.push_back(ORBmatcher::DescriptorDistance(desc_row(a, (int)(i / 32)), desc_row(b, (int)(i / 32))));
    write_i32(dir, "result", out);
  } else if (mode == "bow0" || mode == "bow1") {
    KeyFrame k1, k2;
    std::vector<MapPoint> o1, o2;
    load_bow_keyframe(dir + "/kf1", k1, o1);
    load_bow_keyframe(dir + "/kf2", k2, o2);
    std::vector<float> par = rd<float>(dir, "params");  // nnratio, checkOri
    ORBmatcher matcher(par[0], par[1] != 0.f);
    std::vector<MapPoint*> vm;
    std::vector<int32_t> out;
    int n;
    if (mode == "bow0") {
      Frame F;
      F.N = k2.N; F.mvKeys = k2.mvKeysUn; F.mvKeysUn = k2.mvKeysUn; F.mvuRight = k2.mvuRight;
      F.mDescriptors = k2.mDescriptors; F.mFeatVec = k2.mFeatVec; F.mvScaleFactors = k2.mvScaleFactors;
      F.mvLevelSigma2 = k2.mvLevelSigma2;
      n = matcher.SearchByBoW(&k1, F, vm);
      for (int f = 0; f < F.N; ++f) out.push_back(index_in(o1, vm[f]));
    } else {
      n = matcher.SearchByBoW(&k1, &k2, vm);
      for (int i = 0; i < k1.N; ++i) out.push_back(index_in(o2, vm[i]));
    }
    out.push_back(n);
    write_i32(dir, "result", out);
  } else if (mode == "init") {
    Frame F1, F2;
    std::vector<MapPoint> o1, o2;
    load_frame(dir + "/f1", F1, o1);
    load_frame(dir + "/f2", F2, o2);
    std::vector<float> prev = rd<float>(dir, "prev"), par = rd<float>(dir, "params");  // window, nnratio, checkOri
    std::vector<cv::Point2f> vb(F1.N);
    for (int i = 0; i < F1.N; ++i) { vb[i].x = prev[2 * i]; vb[i].y = prev[2 * i + 1]; }
    std::vector<int> m12;
    ORBmatcher matcher(par[1], par[2] != 0.f);
    const int n = matcher.SearchForInitialization(F1, F2, vb, m12, (int)par[0]);
    std::vector<int32_t> out(m12.begin(), m12.end());
    out.push_back(n);
    write_i32(dir, "result", out);
    FILE* f = std::fopen((dir + "/prev_out.bin").c_str(), "wb");
    std::fwrite(vb.data(), 8, vb.size(), f);
    std::fclose(f);
  } else if (mode == "reloc") {
    Frame Cur;
    std::vector<MapPoint> owned, pts;
    std::vector<uint8_t> valid;
    load_frame(dir, Cur, owned);
    load_points(dir, pts, valid);
    const Proj P = load_proj(dir);
    Cur.mTcw = mat_f(P.T, 4, 4);
    Frame::fx = P.fx; Frame::fy = P.fy; Frame::cx = P.cx; Frame::cy = P.cy;
    Cur.mbf = P.bf; Cur.mfLogScaleFactor = P.lsf; Cur.mnScaleLevels = (int)Cur.mvScaleFactors.size();
    std::vector<float> ang = rd<float>(dir, "pt_angle"), par = rd<float>(dir, "params");  // th, ORBdist, checkOri
    KeyFrame K;
    K.N = (int)pts.size();
    K.mvKeysUn.resize(K.N);
    K.mvpMapPoints.assign(K.N, (MapPoint*)NULL);
    std::set<MapPoint*> found;
    for (int i = 0; i < K.N; ++i) {
      K.mvKeysUn[i].angle = ang[i];
      pts[i].mbBad = false;
      if (valid[i]) K.mvpMapPoints[i] = &pts[i];
      else if (i % 3 == 0) { K.mvpMapPoints[i] = &pts[i]; found.insert(&pts[i]); }  // skipped through sAlreadyFound
      else if (i % 3 == 1) { K.mvpMapPoints[i] = &pts[i]; pts[i].mbBad = true; }      // skipped through isBad()
    }
    ORBmatcher matcher(0.9f, par[2] != 0.f);
    const int n = matcher.SearchByProjection(Cur, &K, found, par[0], (int)par[1]);
    std::vector<int32_t> out;
    for (int f = 0; f < Cur.N; ++f) out.push_back(index_in(pts, Cur.mvpMapPoints[f]));
    out.push_back(n);
    write_i32(dir, "result", out);
  } else if (mode == "sim3proj" || mode == "fuse" || mode == "fusesim3") {
    KeyFrame K;
    std::vector<MapPoint> owned, pts;
    std::vector<uint8_t> valid;
    load_keyframe(dir, K, owned);
    load_points(dir, pts, valid);
    const Proj P = load_proj(dir);
    K.fx = P.fx; K.fy = P.fy; K.cx = P.cx; K.cy = P.cy; K.mbf = P.bf; K.mfLogScaleFactor = P.lsf;
    std::vector<float> par = rd<float>(dir, "params");  // th
    std::vector<MapPoint*> vp(pts.size());
    for (size_t i = 0; i < pts.size(); ++i) vp[i] = &pts[i];
    ORBmatcher matcher(0.8f, true);
    std::vector<int32_t> out;
    int n;
    if (mode == "sim3proj") {
      std::vector<MapPoint*> vpMatched = K.mvpMapPoints;
      n = matcher.SearchByProjection(&K, mat_f(P.T, 4, 4), vp, vpMatched, (int)par[0]);
      for (int f = 0; f < K.N; ++f) out.push_back(index_in(pts, vpMatched[f]));
    } else if (mode == "fusesim3") {
      std::vector<MapPoint*> vpReplace(pts.size(), (MapPoint*)NULL);
      n = matcher.Fuse(&K, mat_f(P.T, 4, 4), vp, par[0], vpReplace);
      for (size_t i = 0; i < pts.size(); ++i) {
        int idx = index_in(owned, vpReplace[i]);  // meets an existing point of the keyframe ...
        const int k = index_in(pts, vpReplace[i]);  // ... or one that an earlier row added ...
        if (idx < 0 && k >= 0) idx = (int)pts[k].mObservations[&K];
        if (idx < 0 && pts[i].mObservations.count(&K)) idx = (int)pts[i].mObservations[&K];  // ... or was added itself
        out.push_back(idx);
      }
    } else {
      float R[9], t[3];
      for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) R[r * 3 + c] = P.T[r * 4 + c]; t[r] = P.T[r * 4 + 3]; }
      K.Rcw = mat_f(R, 3, 3);
      K.tcw = mat_f(t, 3, 1);
      std::vector<float> ow = rd<float>(dir, "ow");
      K.Ow = mat_f(ow.data(), 3, 1);
      n = matcher.Fuse(&K, vp, par[0]);
      for (size_t i = 0; i < pts.size(); ++i) {
        int idx = -1;
        if (pts[i].mObservations.count(&K)) idx = (int)pts[i].mObservations[&K];  // added to an empty slot
        else if (pts[i].mpReplaced) {  // replaced BY the point sitting in the slot
          for (int f = 0; f < K.N && idx < 0; ++f)
            if (K.mvpMapPoints[f] == pts[i].mpReplaced) idx = f;
        } else {  // replaced the point sitting in the slot
          for (int f = 0; f < K.N && idx < 0; ++f)
            if (K.mvpMapPoints[f] && K.mvpMapPoints[f]->mpReplaced == &pts[i]) idx = f;
        }
        out.push_back(idx);
      }
    }
    out.push_back(n);
    write_i32(dir, "result", out);
  } else if (mode == "bysim3") {
    KeyFrame k1, k2;
    std::vector<MapPoint> o1, o2, p1, p2;
    std::vector<uint8_t> v1, v2;
    load_keyframe(dir + "/f1", k1, o1);
    load_keyframe(dir + "/f2", k2, o2);
    load_points(dir + "/p1", p1, v1);
    load_points(dir + "/p2", p2, v2);
    std::vector<float> g = rd<float>(dir, "pair");  // t1w[12], t2w[12], s12, r12[9], t12[3], fx, fy, cx, cy, lsf1, lsf2
    for (int q = 0; q < 2; ++q) {
      KeyFrame& K = q ? k2 : k1;
      const float* T = &g[q * 12];
      float R[9], t[3];
      for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) R[r * 3 + c] = T[r * 4 + c]; t[r] = T[r * 4 + 3]; }
      K.Rcw = mat_f(R, 3, 3);
      K.tcw = mat_f(t, 3, 1);
      K.fx = g[37]; K.fy = g[38]; K.cx = g[39]; K.cy = g[40];
      K.mfLogScaleFactor = g[41 + q];
      std::vector<MapPoint>& pts = q ? p2 : p1;
      const std::vector<uint8_t>& val = q ? v2 : v1;
      for (int i = 0; i < K.N; ++i) { pts[i].mbBad = false; K.mvpMapPoints[i] = val[i] ? &pts[i] : (MapPoint*)NULL; }
    }
    std::vector<MapPoint*> vm(k1.N, (MapPoint*)NULL);
    std::vector<float> par = rd<float>(dir, "params");  // th
    ORBmatcher matcher(0.75f, true);
    const float s12 = g[24];
    const int n = matcher.SearchBySim3(&k1, &k2, vm, s12, mat_f(&g[25], 3, 3), mat_f(&g[34], 3, 1), par[0]);
    std::vector<int32_t> out;
    for (int i = 0; i < k1.N; ++i) out.push_back(index_in(p2, vm[i]));
    out.push_back(n);
    write_i32(dir, "result", out);
  } else if (mode == "voc") {
    ORBVocabulary voc;
    if (!voc.loadFromTextFile(dir + "/voc.txt")) return 3;
    std::vector<uint8_t> d = rd<uint8_t>(dir, "desc");
    std::vector<float> par = rd<float>(dir, "params");  // levelsup
    std::vector<cv::Mat> feats;
    for (size_t i = 0; i + 32 <= d.size(); i += 32) feats.push_back(desc_row(d, (int)(i / 32)));
    DBoW2::BowVector v;
    DBoW2::FeatureVector fv;
    voc.transform(feats, v, fv, (int)par[0]);
    std::vector<int32_t> out;
    out.push_back((int32_t)voc.size());
    out.push_back((int32_t)v.size());
    FILE* f = std::fopen((dir + "/bow_vals.bin").c_str(), "wb");
    for (DBoW2::BowVector::const_iterator it = v.begin(); it != v.end(); ++it) {
      out.push_back((int32_t)it->first);
      std::fwrite(&it->second, 8, 1, f);
    }
    std::fclose(f);
    out.push_back((int32_t)fv.size());
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
      out.push_back((int32_t)it->first);
      out.push_back((int32_t)it->second.size());
      for (size_t k = 0; k < it->second.size(); ++k) out.push_back((int32_t)it->second[k]);
    }
    write_i32(dir, "result", out);
  } else {
    return 2;
  }
  return 0;
}
