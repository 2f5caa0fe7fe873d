// test_compat_matcher.cc — drives the drop-in ORB_SLAM2::ORBmatcher (compat/ORBmatcher.cc) on cases written by
// tests/test_gpu_compat.py as raw arrays in a directory, and writes what the reference's callers would observe
// (Frame::mvpMapPoints after the call, the return value, vMatchedPairs).  Usage: test_compat_matcher <mode> <dir>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

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

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  const std::string mode = argv[1], dir = argv[2];
  if (mode == "map") {
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
  } else {
    return 2;
  }
  return 0;
}
