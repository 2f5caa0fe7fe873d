// test_compat.cc — links the drop-in ORB_SLAM2::ORBextractor + ComputeStereoMatches against liborbfe.so and prints
// a digest of the results of one stereo pair read from raw files; tests/test_gpu_compat.py compares it with the oracle.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "FrameStereo.h"
#include "ORBextractor.h"

static cv::Mat read_raw(const char* path, int w, int h) {
  cv::Mat m(h, w, CV_8UC1);
  FILE* f = std::fopen(path, "rb");
  if (!f || std::fread(m.ptr(0), 1, (size_t)w * h, f) != (size_t)w * h) { std::fprintf(stderr, "cannot read %s\n", path); std::exit(2); }
  std::fclose(f);
  return m;
}

int main(int argc, char** argv) {
  if (argc < 8) { std::fprintf(stderr, "usage: test_compat left.raw right.raw w h nfeatures bf fx out.bin\n"); return 2; }
  const int w = std::atoi(argv[3]), h = std::atoi(argv[4]), nf = std::atoi(argv[5]);
  const float bf = (float)std::atof(argv[6]), fx = (float)std::atof(argv[7]);
  cv::Mat left = read_raw(argv[1], w, h), right = read_raw(argv[2], w, h);
  ORB_SLAM2::ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);  // Tracking.cc:136,139
  std::vector<cv::KeyPoint> kL, kR;
  cv::Mat dL, dR;
  exL(left, cv::Mat(), kL, dL);   // Frame::ExtractORB(0, imLeft)  (Frame.cc:263)
  exR(right, cv::Mat(), kR, dR);  // Frame::ExtractORB(1, imRight) (Frame.cc:265)
  std::vector<float> uR, depth;
  const int n = ORB_SLAM2::compat::ComputeStereoMatches(exL, exR, (int)kL.size(), bf, bf / fx, uR, depth);
  FILE* f = std::fopen(argv[8], "wb");
  int32_t hdr[3] = {(int32_t)kL.size(), (int32_t)kR.size(), n};
  std::fwrite(hdr, 4, 3, f);
  std::fwrite(kL.data(), sizeof(cv::KeyPoint), kL.size(), f);
  std::fwrite(dL.ptr(0), 32, kL.size(), f);
  std::fwrite(uR.data(), 4, uR.size(), f);
  std::fwrite(depth.data(), 4, depth.size(), f);
  std::fclose(f);
  std::printf("levels %d scale %.3f left %zu right %zu stereo %d\n", exL.GetLevels(), exL.GetScaleFactor(), kL.size(), kR.size(), n);
  return 0;
}
