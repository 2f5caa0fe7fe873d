// bench_dropin.cc — latency of the drop-in call sequence Tracking would execute for one stereo frame
// (src/Frame.cc:89-102): compat::ORBextractor::operator() for the left and the right image from TWO THREADS on pageable
// cv::Mat buffers, join, then ComputeStereoMatches.  Prints one JSON object with p50 / p95 / mean in milliseconds.
// Used by bench.py (rank 0, outside its timed regions).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "FrameStereo.h"
#include "ORBextractor.h"

static cv::Mat read_raw(FILE* f, int w, int h) {
  cv::Mat m(h, w, CV_8UC1);
  if (std::fread(m.ptr(0), 1, (size_t)w * h, f) != (size_t)w * h) { std::fprintf(stderr, "short read\n"); std::exit(2); }
  return m;
}

int main(int argc, char** argv) {
  if (argc < 10) {
    std::fprintf(stderr, "usage: bench_dropin pairs.raw n_pairs w h nfeatures bf fx frames warmup\n"
                         "  pairs.raw = n_pairs x (left, right) u8 images back to back\n");
    return 2;
  }
  const int np = std::atoi(argv[2]), w = std::atoi(argv[3]), h = std::atoi(argv[4]), nf = std::atoi(argv[5]);
  const float bf = (float)std::atof(argv[6]), fx = (float)std::atof(argv[7]);
  const int frames = std::atoi(argv[8]), warm = std::atoi(argv[9]);
  FILE* f = std::fopen(argv[1], "rb");
  if (!f) { std::fprintf(stderr, "cannot open %s\n", argv[1]); return 2; }
  std::vector<cv::Mat> L, R;
  for (int i = 0; i < np; ++i) { L.push_back(read_raw(f, w, h)); R.push_back(read_raw(f, w, h)); }
  std::fclose(f);
  ORB_SLAM2::ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);  // Tracking.cc:136,139
  std::vector<cv::KeyPoint> kL, kR;
  cv::Mat dL, dR;
  std::vector<float> uR, depth;
  std::vector<double> ms;
  long kp = 0, st = 0;
  for (int i = 0; i < frames + warm; ++i) {
    const cv::Mat& il = L[i % np];
    const cv::Mat& ir = R[i % np];
    const auto t0 = std::chrono::steady_clock::now();
    std::thread tl([&] { exL(il, cv::Mat(), kL, dL); });  // Frame.cc:89
    std::thread tr([&] { exR(ir, cv::Mat(), kR, dR); });  // Frame.cc:90
    tl.join();
    tr.join();
    const int n = ORB_SLAM2::compat::ComputeStereoMatches(exL, exR, (int)kL.size(), bf, bf / fx, uR, depth);  // :102
    const auto t1 = std::chrono::steady_clock::now();
    if (!exL.ok() || !exR.ok()) { std::fprintf(stderr, "device failure\n"); return 3; }
    if (i >= warm) { ms.push_back(std::chrono::duration<double, std::milli>(t1 - t0).count()); kp += (long)kL.size(); st += n; }
  }
  std::sort(ms.begin(), ms.end());
  double mean = 0;
  for (double v : ms) mean += v;
  mean /= (double)ms.size();
  std::printf("{\"frames\": %d, \"ms_p50\": %.4f, \"ms_p95\": %.4f, \"ms_mean\": %.4f, \"keypoints_per_image\": %.1f, "
              "\"stereo_matches_per_frame\": %.1f}\n",
              frames, ms[ms.size() / 2], ms[(size_t)(ms.size() * 0.95)], mean, (double)kp / frames, (double)st / frames);
  return 0;
}
