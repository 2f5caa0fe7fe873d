// test_compat_mono.cc — the drop-in bodies of Frame::UndistortKeyPoints / ComputeImageBounds / ComputeStereoFromRGBD
// (compat/FrameMono.h) driven the way the RGB-D Frame constructor does (src/Frame.cc:118-173): ExtractORB, then
// UndistortKeyPoints, ComputeStereoFromRGBD.  Writes mvKeys, mvKeysUn, mvuRight, mvDepth and the image bounds for
// tests/test_gpu_compat.py to compare with the oracle.
//   test_compat_mono <image.raw> <depth.f32> <w> <h> <nfeatures> <mbf> <out.bin>      (TUM1 camera)
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "FrameMono.h"

int main(int argc, char** argv) {
  if (argc < 8) return 2;
  const int w = std::atoi(argv[3]), h = std::atoi(argv[4]), nf = std::atoi(argv[5]);
  const float mbf = (float)std::atof(argv[6]);
  cv::Mat im(h, w, CV_8UC1);
  std::vector<float> depth((size_t)w * h);
  FILE* f = std::fopen(argv[1], "rb");
  if (!f || std::fread(im.ptr(0), 1, (size_t)w * h, f) != (size_t)w * h) return 3;
  std::fclose(f);
  f = std::fopen(argv[2], "rb");
  if (!f || std::fread(depth.data(), 4, (size_t)w * h, f) != (size_t)w * h) return 3;
  std::fclose(f);
  orbfe_camera cam = {517.306408f, 516.469215f, 318.643040f, 255.313989f,
                      {0.262383f, -0.953104f, -0.005358f, 0.002628f, 1.163314f}};  // Examples/Monocular/TUM1.yaml
  ORB_SLAM2::ORBextractor ex(nf, 1.2f, 8, 20, 7);
  std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
  cv::Mat desc;
  ex(im, cv::Mat(), mvKeys, desc);
  if (!ORB_SLAM2::compat::UndistortKeyPoints(ex, mvKeys, cam, mvKeysUn)) return 4;
  std::vector<float> mvuRight, mvDepth;
  const int N = (int)mvKeys.size();
  if (!ORB_SLAM2::compat::ComputeStereoFromRGBD(ex, N, depth.data(), w, mbf, mvuRight, mvDepth)) return 5;
  float bounds[4];
  if (orbfe_image_bounds(&cam, w, h, bounds) != ORBFE_OK) return 6;
  f = std::fopen(argv[7], "wb");
  std::fwrite(&N, 4, 1, f);
  std::fwrite(bounds, 4, 4, f);
  std::fwrite(mvKeys.data(), sizeof(cv::KeyPoint), N, f);
  std::fwrite(mvKeysUn.data(), sizeof(cv::KeyPoint), N, f);
  std::fwrite(mvuRight.data(), 4, N, f);
  std::fwrite(mvDepth.data(), 4, N, f);
  std::fclose(f);
  std::printf("mono drop-in: %d keypoints\n", N);
  return 0;
}
