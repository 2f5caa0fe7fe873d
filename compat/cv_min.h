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
#ifndef CV_32F
#define CV_32F 5
#endif
struct Point2f { float x = 0, y = 0; };
struct KeyPoint {
  Point2f pt;
  float size = 0, angle = -1, response = 0;
  int octave = 0, class_id = -1;
};
// minimal continuous matrix (CV_8U or CV_32F, one channel) with shared ownership: enough for images, descriptors,
// pyramid levels and the small float matrices (poses, F12) the matcher signatures carry
class Mat {
 public:
  int rows = 0, cols = 0;
  size_t step = 0;
  uint8_t* data = nullptr;
  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(int r, int c, int type, void* ext, size_t stp = 0)
      : rows(r), cols(c), step(stp ? stp : (size_t)c * esz(type)), data((uint8_t*)ext), type_(type) {}
  void create(int r, int c, int type) {
    if (r == rows && c == cols && type == type_ && own_) return;
    rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
    own_ = std::shared_ptr<uint8_t>(new uint8_t[(size_t)r * step + 4](), std::default_delete<uint8_t[]>());
    data = own_.get();
  }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  int type() const { return type_; }
  void release() { own_.reset(); data = nullptr; rows = cols = 0; }
  uint8_t* ptr(int r = 0) { return data + (size_t)r * step; }
  const uint8_t* ptr(int r = 0) const { return data + (size_t)r * step; }
  template <class T> T* ptr(int r = 0) { return reinterpret_cast<T*>(data + (size_t)r * step); }
  template <class T> const T* ptr(int r = 0) const { return reinterpret_cast<const T*>(data + (size_t)r * step); }
  template <class T> T& at(int r, int c = 0) { return ptr<T>(r)[c]; }
  template <class T> const T& at(int r, int c = 0) const { return ptr<T>(r)[c]; }
  Mat row(int r) const { Mat m; m.rows = 1; m.cols = cols; m.step = step; m.type_ = type_; m.data = data + (size_t)r * step; m.own_ = own_; return m; }
 private:
  static size_t esz(int type) { return type == CV_32F ? 4 : 1; }
  int type_ = CV_8UC1;
  std::shared_ptr<uint8_t> own_;
};
typedef const Mat& InputArray;
typedef Mat& OutputArray;
}  // namespace cv
#endif
