// cvshim.hpp — a minimal stand-in for the OpenCV C++ API, written for ONE purpose: to compile the UNMODIFIED sources of
// the reference (/root/reference/src/ORBextractor.cc, Frame.cc, ORBmatcher.cc, MapPoint.cc, KeyFrame.cc, Map.cc and the
// vendored DBoW2) in a container that has no OpenCV C++ headers or libraries, so that the oracle (and through it the CUDA
// path) can be pinned to the reference's literal code.
//
// TEST INFRASTRUCTURE ONLY (like everything under oracle/): never linked into the product.
//
// What is here is original code, not OpenCV source: cv::Mat (reference-counted, 2-D, ROI views), the lazy matrix
// expressions the reference uses (A*B+C, -A.t()*b, s*A, A/s, A-B) with OpenCV's evaluation rules restated from its
// documented behaviour (see cvshim.cc for each rule), KeyPoint / Point_ / Size_ / Rect_ / Scalar_, the Input/OutputArray
// proxies, and declarations of the five pixel primitives (resize, copyMakeBorder, FAST, GaussianBlur, fastAtan2), which
// cvshim.cc routes to the cv2-4.13-pinned integer restatements of oracle/orb_prims.cc.
#ifndef ORACLE_REF_SHIM_CVSHIM_HPP_
#define ORACLE_REF_SHIM_CVSHIM_HPP_

#include <algorithm>
#include <cassert>
#include <climits>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <fstream>
#include <limits>
#include <list>
#include <map>
#include <memory>
#include <numeric>
#include <set>
#include <sstream>
#include <string>
#include <vector>

typedef unsigned char uchar;
typedef unsigned short ushort;
typedef int64_t int64;
typedef uint64_t uint64;

#define CV_PI 3.1415926535897932384626433832795

#define CV_CN_SHIFT 3
#define CV_DEPTH_MAX (1 << CV_CN_SHIFT)
#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_MAT_DEPTH_MASK (CV_DEPTH_MAX - 1)
#define CV_MAT_DEPTH(flags) ((flags) & CV_MAT_DEPTH_MASK)
#define CV_MAKETYPE(depth, cn) (CV_MAT_DEPTH(depth) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_MAT_CN(flags) ((((flags) >> CV_CN_SHIFT) & 511) + 1)
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_8UC4 CV_MAKETYPE(CV_8U, 4)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)

#define CV_Assert(expr) do { if (!(expr)) ::cv::shim_fail(#expr, __FILE__, __LINE__); } while (0)

// cvRound: round half to even (OpenCV uses cvtsd2si / lrint in the default rounding mode)
static inline int cvRound(double v) { return (int)std::nearbyint(v); }
static inline int cvRound(float v) { return (int)std::nearbyintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

[[noreturn]] void shim_fail(const char* what, const char* file, int line);

enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };
enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
enum { GEMM_1_T = 1, GEMM_2_T = 2, GEMM_3_T = 4 };

template <typename T> static inline T saturate_cast(double v) { return (T)v; }
template <> inline uchar saturate_cast<uchar>(double v) { int i = cvRound(v); return (uchar)(i < 0 ? 0 : i > 255 ? 255 : i); }
template <> inline int saturate_cast<int>(double v) { return cvRound(v); }

// ---------------------------------------------------------------- small value types
template <typename T> struct Point_ {
  T x, y;
  Point_() : x(0), y(0) {}
  Point_(T _x, T _y) : x(_x), y(_y) {}
  template <typename U> explicit Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
  Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }   // fp32 product for Point2f (a6/:1120)
  Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
  Point_& operator-=(const Point_& o) { x -= o.x; y -= o.y; return *this; }
};
template <typename T> static inline Point_<T> operator+(Point_<T> a, const Point_<T>& b) { a += b; return a; }
template <typename T> static inline Point_<T> operator-(Point_<T> a, const Point_<T>& b) { a -= b; return a; }
template <typename T> static inline bool operator==(const Point_<T>& a, const Point_<T>& b) { return a.x == b.x && a.y == b.y; }
typedef Point_<int> Point2i;
typedef Point_<int> Point;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;

template <typename T> struct Point3_ {
  T x, y, z;
  Point3_() : x(0), y(0), z(0) {}
  Point3_(T _x, T _y, T _z) : x(_x), y(_y), z(_z) {}
};
typedef Point3_<float> Point3f;
typedef Point3_<double> Point3d;

template <typename T> struct Size_ {
  T width, height;
  Size_() : width(0), height(0) {}
  Size_(T w, T h) : width(w), height(h) {}
  T area() const { return width * height; }
};
template <typename T> static inline bool operator==(const Size_<T>& a, const Size_<T>& b) { return a.width == b.width && a.height == b.height; }
template <typename T> static inline bool operator!=(const Size_<T>& a, const Size_<T>& b) { return !(a == b); }
typedef Size_<int> Size2i;
typedef Size_<int> Size;

template <typename T> struct Rect_ {
  T x, y, width, height;
  Rect_() : x(0), y(0), width(0), height(0) {}
  Rect_(T _x, T _y, T w, T h) : x(_x), y(_y), width(w), height(h) {}
};
typedef Rect_<int> Rect;

template <typename T> struct Scalar_ {
  T val[4];
  Scalar_() { val[0] = val[1] = val[2] = val[3] = 0; }
  Scalar_(T v0, T v1 = 0, T v2 = 0, T v3 = 0) { val[0] = v0; val[1] = v1; val[2] = v2; val[3] = v3; }
  static Scalar_ all(T v) { return Scalar_(v, v, v, v); }
  T& operator[](int i) { return val[i]; }
  const T& operator[](int i) const { return val[i]; }
  bool isReal() const { return val[1] == 0 && val[2] == 0 && val[3] == 0; }
  bool isZero() const { return val[0] == 0 && isReal(); }
};
typedef Scalar_<double> Scalar;

struct Range {
  int start, end;
  Range() : start(0), end(0) {}
  Range(int s, int e) : start(s), end(e) {}
  static Range all() { return Range(INT32_MIN, INT32_MAX); }
};

class KeyPoint {
 public:
  KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
  KeyPoint(Point2f _pt, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
      : pt(_pt), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
  KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
      : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
  Point2f pt;
  float size;
  float angle;
  float response;
  int octave;
  int class_id;
};
static_assert(sizeof(KeyPoint) == 28, "cv::KeyPoint layout (a16)");

template <typename T> struct DataType;
template <> struct DataType<uchar> { enum { type = CV_8UC1, depth = CV_8U }; };
template <> struct DataType<short> { enum { type = CV_MAKETYPE(CV_16S, 1), depth = CV_16S }; };
template <> struct DataType<int> { enum { type = CV_32SC1, depth = CV_32S }; };
template <> struct DataType<float> { enum { type = CV_32FC1, depth = CV_32F }; };
template <> struct DataType<double> { enum { type = CV_64FC1, depth = CV_64F }; };

// ---------------------------------------------------------------- Mat
class MatExpr;
class Mat;

// `step` converts to size_t like cv::MatStep
struct MatStep {
  size_t v = 0;
  operator size_t() const { return v; }
  MatStep& operator=(size_t s) { v = s; return *this; }
};

class Mat {
 public:
  Mat() {}
  Mat(int r, int c, int type) { create(r, c, type); }
  Mat(Size sz, int type) { create(sz.height, sz.width, type); }
  Mat(int r, int c, int type, const Scalar& s) { create(r, c, type); setTo(s); }
  Mat(Size sz, int type, const Scalar& s) { create(sz.height, sz.width, type); setTo(s); }
  // user data, not owned
  Mat(int r, int c, int type, void* ptr, size_t stp = 0) {
    flags = type; rows = r; cols = c; data = (uchar*)ptr;
    step = stp ? stp : (size_t)c * elemSize();
  }
  Mat(const Mat& m, const Rect& roi) { *this = m; applyRoi(roi.y, roi.y + roi.height, roi.x, roi.x + roi.width); }
  Mat(const MatExpr& e);  // NOLINT: implicit, like OpenCV
  Mat& operator=(const MatExpr& e);

  static size_t elemSize1Of(int type) {
    switch (CV_MAT_DEPTH(type)) { case CV_8U: case CV_8S: return 1; case CV_16U: case CV_16S: return 2;
      case CV_32S: case CV_32F: return 4; default: return 8; }
  }
  void create(int r, int c, int type) {
    if (data && rows == r && cols == c && this->type() == type) return;
    flags = type; rows = r; cols = c;
    step = (size_t)c * elemSize();
    const size_t bytes = step * (size_t)r;
    buf = std::shared_ptr<uchar>(bytes ? (uchar*)std::malloc(bytes + 64) : nullptr, std::free);
    data = buf.get();
  }
  void create(Size sz, int type) { create(sz.height, sz.width, type); }
  void release() { buf.reset(); data = nullptr; rows = cols = 0; step = 0; }

  int type() const { return flags & 4095; }
  int depth() const { return CV_MAT_DEPTH(flags); }
  int channels() const { return CV_MAT_CN(flags); }
  size_t elemSize1() const { return elemSize1Of(flags); }
  size_t elemSize() const { return elemSize1() * channels(); }
  size_t step1() const { return step / elemSize1(); }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  size_t total() const { return (size_t)rows * cols; }
  Size size() const { return Size(cols, rows); }
  bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }

  uchar* ptr(int y = 0) { return data + step * (size_t)y; }
  const uchar* ptr(int y = 0) const { return data + step * (size_t)y; }
  template <typename T> T* ptr(int y = 0) { return (T*)(data + step * (size_t)y); }
  template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + step * (size_t)y); }
  template <typename T> T& at(int y, int x) { return ((T*)(data + step * (size_t)y))[x]; }
  template <typename T> const T& at(int y, int x) const { return ((const T*)(data + step * (size_t)y))[x]; }
  // single index: the i-th element of a vector (1 x n or n x 1), as OpenCV defines it
  template <typename T> T& at(int i) { return rows == 1 ? at<T>(0, i) : cols == 1 ? at<T>(i, 0) : at<T>(i / cols, i % cols); }
  template <typename T> const T& at(int i) const { return const_cast<Mat*>(this)->at<T>(i); }

  Mat rowRange(int s, int e) const { Mat m(*this); m.applyRoi(s, e, 0, cols); return m; }
  Mat colRange(int s, int e) const { Mat m(*this); m.applyRoi(0, rows, s, e); return m; }
  Mat rowRange(const Range& r) const { return rowRange(r.start, r.end); }
  Mat colRange(const Range& r) const { return colRange(r.start, r.end); }
  Mat row(int y) const { return rowRange(y, y + 1); }
  Mat col(int x) const { return colRange(x, x + 1); }
  Mat operator()(const Rect& r) const { return Mat(*this, r); }
  Mat operator()(Range rr, Range cr) const {
    Mat m(*this);
    m.applyRoi(rr.start == INT32_MIN ? 0 : rr.start, rr.end == INT32_MAX ? rows : rr.end,
               cr.start == INT32_MIN ? 0 : cr.start, cr.end == INT32_MAX ? cols : cr.end);
    return m;
  }

  Mat clone() const { Mat m; copyTo(m); return m; }
  void copyTo(Mat& dst) const;
  void copyTo(const class _OutputArray& dst) const;
  void convertTo(Mat& dst, int rtype, double alpha = 1, double beta = 0) const;
  void convertTo(const class _OutputArray& dst, int rtype, double alpha = 1, double beta = 0) const;
  Mat& setTo(const Scalar& s);
  Mat& operator=(const Scalar& s) { return setTo(s); }
  Mat reshape(int cn, int new_rows = 0) const;
  MatExpr t() const;
  MatExpr inv(int method = 0) const;
  MatExpr mul(const Mat& m, double scale = 1) const;
  double dot(const Mat& m) const;
  Mat cross(const Mat& m) const;
  void push_back(const Mat& m);
  template <typename T> void push_back(const T& v);

  // lazy initialisers, as in OpenCV: `m = Mat::zeros(r, c, t)` FILLS m in place when m already has that shape and type
  // (the reference relies on it: computeDescriptors() zeroes a row range of the caller's descriptor matrix, :1059)
  static MatExpr zeros(int r, int c, int type);
  static MatExpr zeros(Size sz, int type);
  static MatExpr ones(int r, int c, int type);
  static MatExpr ones(Size sz, int type);
  static MatExpr eye(int r, int c, int type);
  static MatExpr eye(Size sz, int type);

  int flags = 0;
  int dims = 2;
  int rows = 0, cols = 0;
  uchar* data = nullptr;
  MatStep step;
  std::shared_ptr<uchar> buf;  // null for user data

 private:
  void applyRoi(int r0, int r1, int c0, int c1) {
    if (r0 < 0 || r1 > rows || c0 < 0 || c1 > cols || r0 > r1 || c0 > c1) shim_fail("Mat ROI out of range", __FILE__, __LINE__);
    data += step * (size_t)r0 + (size_t)c0 * elemSize();
    rows = r1 - r0;
    cols = c1 - c0;
  }
};

template <typename T> void Mat::push_back(const T& v) {
  Mat one(1, 1, DataType<T>::type);
  one.at<T>(0, 0) = v;
  push_back(one);
}

template <typename T> class MatCommaInitializer_;

template <typename T> class Mat_ : public Mat {
 public:
  Mat_() {}
  Mat_(int r, int c) : Mat(r, c, DataType<T>::type) {}
  Mat_(const Mat& m) : Mat(m) { if (!m.empty() && m.type() != DataType<T>::type) m.convertTo(*this, DataType<T>::type); }  // NOLINT
  Mat_(const MatExpr& e);  // NOLINT
  T& operator()(int y, int x) { return this->template at<T>(y, x); }
  const T& operator()(int y, int x) const { return this->template at<T>(y, x); }
  T& operator()(int i) { return this->template at<T>(i); }
  const T& operator()(int i) const { return this->template at<T>(i); }
};

template <typename T> class MatCommaInitializer_ {
 public:
  explicit MatCommaInitializer_(Mat_<T>* m) : m_(m), i_(0) {}
  template <typename U> MatCommaInitializer_& operator,(U v) {
    const int n = (int)m_->total();
    if (i_ >= n) shim_fail("Mat_ comma initializer: too many values", __FILE__, __LINE__);
    m_->template at<T>(i_ / m_->cols, i_ % m_->cols) = (T)v;
    ++i_;
    return *this;
  }
  operator Mat_<T>() const { return *m_; }
  operator Mat() const { return *m_; }
  Mat_<T>* m_;
  int i_;
};
template <typename T, typename U> static inline MatCommaInitializer_<T> operator<<(const Mat_<T>& m, U v) {
  MatCommaInitializer_<T> ci(const_cast<Mat_<T>*>(&m));
  return (ci, v);
}

// ---------------------------------------------------------------- lazy expressions (rules restated in cvshim.cc)
class MatExpr {
 public:
  enum Kind { IDENT = 0, TRANSPOSE, GEMM, ADDEX, INV, INIT };
  MatExpr() {}
  MatExpr(const Mat& m) : kind(IDENT), a(m) {}  // NOLINT
  operator Mat() const { Mat m; assign(m); return m; }
  template <typename T> operator Mat_<T>() const { Mat m; assign(m); return Mat_<T>(m); }
  void assign(Mat& m) const;
  MatExpr t() const;
  double dot(const Mat& m) const { return ((Mat)*this).dot(m); }
  Mat row(int y) const { return ((Mat)*this).row(y); }
  Mat col(int x) const { return ((Mat)*this).col(x); }
  Size size() const { return ((Mat)*this).size(); }

  Kind kind = IDENT;
  int gemm_flags = 0;
  Mat a, b, c;
  double alpha = 1, beta = 0;
  Scalar s;
  int init_rows = 0, init_cols = 0, init_type = 0, init_what = 0;  // INIT: '0' zeros, '1' ones, 'I' eye (times alpha)
};

inline Mat::Mat(const MatExpr& e) { e.assign(*this); }
inline Mat& Mat::operator=(const MatExpr& e) { e.assign(*this); return *this; }
template <typename T> inline Mat_<T>::Mat_(const MatExpr& e) { Mat m; e.assign(m); *this = Mat_<T>(m); }

MatExpr operator+(const Mat& a, const Mat& b);
MatExpr operator+(const MatExpr& e, const Mat& b);
MatExpr operator+(const Mat& a, const MatExpr& e);
MatExpr operator+(const MatExpr& e1, const MatExpr& e2);
MatExpr operator-(const Mat& a, const Mat& b);
MatExpr operator-(const MatExpr& e, const Mat& b);
MatExpr operator-(const Mat& a, const MatExpr& e);
MatExpr operator-(const MatExpr& e1, const MatExpr& e2);
MatExpr operator-(const Mat& m);
MatExpr operator-(const MatExpr& e);
MatExpr operator*(const Mat& a, const Mat& b);
MatExpr operator*(const MatExpr& e, const Mat& b);
MatExpr operator*(const Mat& a, const MatExpr& e);
MatExpr operator*(const MatExpr& e1, const MatExpr& e2);
MatExpr operator*(const Mat& a, double s);
MatExpr operator*(double s, const Mat& a);
MatExpr operator*(const MatExpr& e, double s);
MatExpr operator*(double s, const MatExpr& e);
MatExpr operator/(const Mat& a, double s);
MatExpr operator/(const MatExpr& e, double s);

// ---------------------------------------------------------------- array proxies
class _InputArray {
 public:
  _InputArray() : m_(nullptr) {}
  _InputArray(const Mat& m) : m_(&m) {}  // NOLINT
  _InputArray(const MatExpr& e) : own_(new Mat(e)), m_(own_.get()) {}  // NOLINT
  Mat getMat() const { return m_ ? *m_ : Mat(); }
  bool empty() const { return !m_ || m_->empty(); }
  Size size() const { return m_ ? m_->size() : Size(); }
  int type() const { return m_ ? m_->type() : 0; }
 protected:
  std::shared_ptr<Mat> own_;
  const Mat* m_;
};
class _OutputArray : public _InputArray {
 public:
  _OutputArray() : w_(nullptr) {}
  _OutputArray(Mat& m) : _InputArray(m), w_(&m) {}  // NOLINT
  // a temporary header (e.g. `R.copyTo(T.rowRange(0,3).colRange(0,3))`): fixed size, written in place
  _OutputArray(const Mat& m) { own_.reset(new Mat(m)); m_ = own_.get(); w_ = own_.get(); }  // NOLINT
  void create(int r, int c, int type) const { if (w_) w_->create(r, c, type); }
  void create(Size sz, int type) const { if (w_) w_->create(sz, type); }
  void release() const { if (w_) w_->release(); }
  Mat& getMatRef() const { return *w_; }
  bool needed() const { return w_ != nullptr; }
 private:
  Mat* w_;
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
typedef const _OutputArray& InputOutputArray;
static inline InputArray noArray() { static _OutputArray none; return none; }

// ---------------------------------------------------------------- functions
double norm(InputArray a, int normType = NORM_L2);
double norm(InputArray a, InputArray b, int normType = NORM_L2);
void gemm(const Mat& a, const Mat& b, double alpha, const Mat& c, double beta, Mat& d, int flags = 0);
void transpose(const Mat& a, Mat& d);

// the pixel primitives of the hot path: implemented by the cv2-4.13-pinned restatements in oracle/orb_prims.cc
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType,
                    const Scalar& value = Scalar());
void FAST(InputArray image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression = true);
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0,
                  int borderType = BORDER_DEFAULT);
float fastAtan2(float y, float x);
// only reached with non-zero distortion coefficients (not on the rectified-stereo path): aborts
void undistortPoints(InputArray src, OutputArray dst, InputArray cameraMatrix, InputArray distCoeffs,
                     InputArray R = noArray(), InputArray P = noArray());

struct KeyPointsFilter {
  // dead code in the reference (ComputeKeyPointsOld); must link
  static void retainBest(std::vector<KeyPoint>& keypoints, int npoints);
};

// ---------------------------------------------------------------- persistence skeleton (DBoW2 save/load; never called)
class FileNode;
class FileNodeIterator;
class FileStorage {
 public:
  enum { READ = 0, WRITE = 1, APPEND = 2 };
  FileStorage() {}
  FileStorage(const std::string&, int) { shim_fail("cv::FileStorage is not available in the shim", __FILE__, __LINE__); }
  bool isOpened() const { return false; }
  void release() {}
  FileNode operator[](const std::string& name) const;
  FileNode operator[](const char* name) const;
};
class FileNode {
 public:
  enum { NONE = 0, INT = 1, REAL = 2, STR = 3, SEQ = 4, MAP = 5 };
  FileNode operator[](const std::string&) const { return FileNode(); }
  FileNode operator[](const char*) const { return FileNode(); }
  FileNode operator[](int) const { return FileNode(); }
  int type() const { return NONE; }
  size_t size() const { return 0; }
  bool empty() const { return true; }
  operator int() const { return 0; }
  operator float() const { return 0.f; }
  operator double() const { return 0.; }
  operator std::string() const { return std::string(); }
};
inline FileNode FileStorage::operator[](const std::string&) const { return FileNode(); }
inline FileNode FileStorage::operator[](const char*) const { return FileNode(); }
template <typename T> static inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }
template <typename T> static inline void operator>>(const FileNode&, T&) {}

}  // namespace cv

#endif  // ORACLE_REF_SHIM_CVSHIM_HPP_
