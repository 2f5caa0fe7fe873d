// cv_min.h — the few OpenCV types the drop-in ORBextractor signature needs, for machines WITHOUT OpenCV C++
// (this image).  With OpenCV present, compile with -DORBFE_HAVE_OPENCV and the real headers are used instead.
// cv::KeyPoint here has the same field order and size (28 bytes) as OpenCV's, which is what orbfe_keypoint mirrors.
#pragma once
#ifdef ORBFE_HAVE_OPENCV
#include <opencv2/core/core.hpp>
#include <opencv2/features2d/features2d.hpp>
#else
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>
namespace cv {
enum { CV_8U_ = 0 };
#ifndef CV_8UC1
#define CV_8UC1 0
#define CV_8U 0
#endif
struct Point2f { float x = 0, y = 0; };
struct KeyPoint {
  Point2f pt;
  float size = 0, angle = -1, response = 0;
  int octave = 0, class_id = -1;
};
// minimal continuous 8-bit matrix with shared ownership (enough for images, descriptors and pyramid levels)
class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;
  uint8_t* data = nullptr;
  Mat() {}
  Mat(int r, int c, int /*type*/) { create(r, c, 0); }
  Mat(int r, int c, int /*type*/, void* ext, size_t stp = 0) : rows(r), cols(c), step(stp ? stp : (size_t)c), data((uint8_t*)ext) {}
  void create(int r, int c, int /*type*/) {
    if (r == rows && c == cols && own_) return;
    rows = r; cols = c; step = (size_t)c;
    own_ = std::shared_ptr<uint8_t>(new uint8_t[(size_t)r * c + 1], std::default_delete<uint8_t[]>());
    data = own_.get();
  }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return CV_8UC1; }
  void release() { own_.reset(); data = nullptr; rows = cols = 0; }
  uint8_t* ptr(int r = 0) { return data + (size_t)r * step; }
  const uint8_t* ptr(int r = 0) const { return data + (size_t)r * step; }
  Mat row(int r) const { Mat m; m.rows = 1; m.cols = cols; m.step = step; m.data = data + (size_t)r * step; m.own_ = own_; return m; }
 private:
  std::shared_ptr<uint8_t> own_;
};
typedef const Mat& InputArray;
typedef Mat& OutputArray;
}  // namespace cv
#endif
