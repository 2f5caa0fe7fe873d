// cvshim.cc — implementation of the OpenCV stand-in declared in cvshim.hpp (TEST INFRASTRUCTURE ONLY).
//
// Arithmetic rules restated here (OpenCV 3.x / 4.x behaviour for CV_32F, as far as the reference's path depends on it):
//  * A*B (+C) with 2 <= inner length <= 4 and no transposition flag and (len == D.cols || len == D.rows): the
//    in-order fp32 dot product t = ((a0*b0 + a1*b1) + a2*b2) [+ a3*b3], then D = (float)((double)t*alpha +
//    (double)c*beta).  Every other product: double accumulation s += (double)a*(double)b in order (four interleaved
//    partial sums when B is transposed), D = (float)(s*alpha + (double)c*beta).
//  * -X for an expression X that is not already a scaled sum materialises X and scales by -1; (s*A), (A/s) scale
//    element-wise in fp32 by (float)s resp. (float)(1.0/s); (-A)*B and (s*A)*B fold the scale into the product's alpha.
//  * A +/- B element-wise in fp32; A - s*B = B*(float)(-s) + A (two roundings).
//  * Mat::dot and cv::norm accumulate in double in element order; NORM_L2 = sqrt of that.
//  * convertTo between u8 and f32 is exact; with a scale it is fp32 src*(float)alpha + (float)beta, saturated for u8.
// None of these use a fused multiply-add (compile with -ffp-contract=off).
#include "cvshim.hpp"

#include "../../include/orb_math_shared.h"
#include "../orb_internal.h"

namespace cv {

void shim_fail(const char* what, const char* file, int line) {
  std::fprintf(stderr, "cvshim: %s (%s:%d)\n", what, file, line);
  std::abort();
}

// ---------------------------------------------------------------- Mat basics
void Mat::copyTo(Mat& dst) const {
  if (empty()) { dst.release(); return; }
  if (dst.data == data && dst.rows == rows && dst.cols == cols && dst.type() == type() && (size_t)dst.step == (size_t)step) return;
  dst.create(rows, cols, type());
  const size_t row_bytes = (size_t)cols * elemSize();
  for (int y = 0; y < rows; ++y) std::memmove(dst.ptr(y), ptr(y), row_bytes);
}
void Mat::copyTo(const _OutputArray& dst) const { copyTo(dst.getMatRef()); }

Mat& Mat::setTo(const Scalar& s) {
  const int cn = channels();
  for (int y = 0; y < rows; ++y)
    for (int x = 0; x < cols; ++x)
      for (int c = 0; c < cn; ++c) {
        const double v = s[c < 4 ? c : 3];
        switch (depth()) {
          case CV_8U: ptr<uchar>(y)[x * cn + c] = saturate_cast<uchar>(v); break;
          case CV_32S: ptr<int>(y)[x * cn + c] = (int)v; break;
          case CV_32F: ptr<float>(y)[x * cn + c] = (float)v; break;
          case CV_64F: ptr<double>(y)[x * cn + c] = v; break;
          default: shim_fail("setTo: unsupported depth", __FILE__, __LINE__);
        }
      }
  return *this;
}

template <typename S, typename D> static void convert_rows(const Mat& src, Mat& dst, double alpha, double beta) {
  const int n = src.cols * src.channels();
  const bool scaled = !(alpha == 1.0 && beta == 0.0);
  for (int y = 0; y < src.rows; ++y) {
    const S* s = src.ptr<S>(y);
    D* d = dst.ptr<D>(y);
    for (int x = 0; x < n; ++x) {
      if (!scaled) {
        if (std::is_same<D, uchar>::value && !std::is_same<S, uchar>::value) d[x] = (D)saturate_cast<uchar>((double)s[x]);
        else d[x] = (D)s[x];
      } else if (std::is_same<S, double>::value || std::is_same<D, double>::value) {
        const double v = (double)s[x] * alpha + beta;
        if (std::is_same<D, uchar>::value) d[x] = (D)saturate_cast<uchar>(v); else d[x] = (D)v;
      } else {
        const float v = (float)s[x] * (float)alpha + (float)beta;  // fp32 working type
        if (std::is_same<D, uchar>::value) d[x] = (D)saturate_cast<uchar>((double)v); else d[x] = (D)v;
      }
    }
  }
}

void Mat::convertTo(Mat& dst, int rtype, double alpha, double beta) const {
  if (empty()) { dst.release(); return; }
  const int ddepth = rtype < 0 ? depth() : CV_MAT_DEPTH(rtype);
  const int dtype = CV_MAKETYPE(ddepth, channels());
  Mat src = *this;  // keeps the source alive when dst aliases it
  Mat out;
  if (dst.data && dst.rows == rows && dst.cols == cols && dst.type() == dtype && dst.data != data) out = dst;
  else if (dst.data == data && dtype == type()) out = dst;   // same type in place
  else out.create(rows, cols, dtype);
#define CVT(S, D) convert_rows<S, D>(src, out, alpha, beta)
  const int sd = depth();
  if (sd == CV_8U && ddepth == CV_8U) CVT(uchar, uchar);
  else if (sd == CV_8U && ddepth == CV_32F) CVT(uchar, float);
  else if (sd == CV_8U && ddepth == CV_64F) CVT(uchar, double);
  else if (sd == CV_8U && ddepth == CV_32S) CVT(uchar, int);
  else if (sd == CV_32F && ddepth == CV_8U) CVT(float, uchar);
  else if (sd == CV_32F && ddepth == CV_32F) CVT(float, float);
  else if (sd == CV_32F && ddepth == CV_64F) CVT(float, double);
  else if (sd == CV_64F && ddepth == CV_32F) CVT(double, float);
  else if (sd == CV_64F && ddepth == CV_64F) CVT(double, double);
  else if (sd == CV_32S && ddepth == CV_32F) CVT(int, float);
  else if (sd == CV_32S && ddepth == CV_32S) CVT(int, int);
  else shim_fail("convertTo: unsupported depth pair", __FILE__, __LINE__);
#undef CVT
  dst = out;
}
void Mat::convertTo(const _OutputArray& dst, int rtype, double alpha, double beta) const {
  convertTo(dst.getMatRef(), rtype, alpha, beta);
}

Mat Mat::reshape(int cn, int new_rows) const {
  if (!isContinuous() && rows > 1) shim_fail("reshape: needs a continuous matrix", __FILE__, __LINE__);
  if (cn == 0) cn = channels();
  Mat m(*this);
  const size_t total_scalars = (size_t)rows * cols * channels();
  const int r = new_rows ? new_rows : rows;
  if (total_scalars % ((size_t)r * cn)) shim_fail("reshape: size mismatch", __FILE__, __LINE__);
  m.flags = CV_MAKETYPE(depth(), cn);
  m.rows = r;
  m.cols = (int)(total_scalars / ((size_t)r * cn));
  m.step = (size_t)m.cols * m.elemSize();
  return m;
}

void Mat::push_back(const Mat& m) {
  if (empty()) { *this = m.clone(); return; }
  if (m.type() != type() || m.cols != cols) shim_fail("push_back: shape/type mismatch", __FILE__, __LINE__);
  Mat out(rows + m.rows, cols, type());
  const size_t rb = (size_t)cols * elemSize();
  for (int y = 0; y < rows; ++y) std::memcpy(out.ptr(y), ptr(y), rb);
  for (int y = 0; y < m.rows; ++y) std::memcpy(out.ptr(rows + y), m.ptr(y), rb);
  *this = out;
}

static MatExpr make_init(int r, int c, int type, int what) {
  MatExpr e;
  e.kind = MatExpr::INIT;
  e.init_rows = r; e.init_cols = c; e.init_type = type; e.init_what = what;
  return e;
}
MatExpr Mat::zeros(int r, int c, int type) { return make_init(r, c, type, '0'); }
MatExpr Mat::ones(int r, int c, int type) { return make_init(r, c, type, '1'); }
MatExpr Mat::eye(int r, int c, int type) { return make_init(r, c, type, 'I'); }
MatExpr Mat::zeros(Size sz, int type) { return zeros(sz.height, sz.width, type); }
MatExpr Mat::ones(Size sz, int type) { return ones(sz.height, sz.width, type); }
MatExpr Mat::eye(Size sz, int type) { return eye(sz.height, sz.width, type); }

void transpose(const Mat& a, Mat& d) {
  Mat out(a.cols, a.rows, a.type());
  const size_t es = a.elemSize();
  for (int y = 0; y < a.rows; ++y)
    for (int x = 0; x < a.cols; ++x) std::memcpy(out.ptr(x) + (size_t)y * es, a.ptr(y) + (size_t)x * es, es);
  d = out;
}

// element (i, k) of A or A^T as double / float
template <typename T> static inline T el(const Mat& m, bool tr, int i, int k) { return tr ? m.at<T>(k, i) : m.at<T>(i, k); }

template <typename T> static void gemm_t(const Mat& A, const Mat& B, double alpha, const Mat& C, double beta, Mat& D, int flags) {
  const bool ta = flags & GEMM_1_T, tb = flags & GEMM_2_T, tc = flags & GEMM_3_T;
  const int M = ta ? A.cols : A.rows, len = ta ? A.rows : A.cols;
  const int N = tb ? B.rows : B.cols, lenb = tb ? B.cols : B.rows;
  if (len != lenb) shim_fail("gemm: inner sizes differ", __FILE__, __LINE__);
  const bool hasC = !C.empty() && beta != 0;
  if (hasC && ((tc ? C.cols : C.rows) != M || (tc ? C.rows : C.cols) != N)) shim_fail("gemm: C size", __FILE__, __LINE__);
  Mat out(M, N, A.type());
  const bool small = flags == 0 && len >= 2 && len <= 4 && (len == N || len == M);
  for (int i = 0; i < M; ++i)
    for (int j = 0; j < N; ++j) {
      const double c = hasC ? (double)(tc ? C.at<T>(j, i) : C.at<T>(i, j)) : 0.0;
      double r;
      if (small) {
        T t = el<T>(A, false, i, 0) * el<T>(B, false, 0, j);
        for (int k = 1; k < len; ++k) t = t + el<T>(A, false, i, k) * el<T>(B, false, k, j);
        r = (double)t * alpha + c * beta;
      } else if (tb && len >= 4) {
        double s[4] = {0, 0, 0, 0};
        int k = 0;
        for (; k <= len - 4; k += 4)
          for (int u = 0; u < 4; ++u) s[u] += (double)el<T>(A, ta, i, k + u) * (double)el<T>(B, true, k + u, j);
        for (; k < len; ++k) s[0] += (double)el<T>(A, ta, i, k) * (double)el<T>(B, true, k, j);
        r = (s[0] + s[1] + s[2] + s[3]) * alpha;
        if (hasC) r = r + c * beta;
      } else {
        double s = 0;
        for (int k = 0; k < len; ++k) s += (double)el<T>(A, ta, i, k) * (double)el<T>(B, tb, k, j);
        r = s * alpha;
        if (hasC) r = r + c * beta;
      }
      out.at<T>(i, j) = (T)r;
    }
  D = out;
}

void gemm(const Mat& a, const Mat& b, double alpha, const Mat& c, double beta, Mat& d, int flags) {
  if (a.type() != b.type() || (a.type() != CV_32F && a.type() != CV_64F)) shim_fail("gemm: type", __FILE__, __LINE__);
  if (a.type() == CV_32F) gemm_t<float>(a, b, alpha, c, beta, d, flags); else gemm_t<double>(a, b, alpha, c, beta, d, flags);
}

// ---------------------------------------------------------------- expressions
static MatExpr addex(const Mat& a, const Mat& b, double alpha, double beta, const Scalar& s = Scalar()) {
  MatExpr e;
  e.kind = MatExpr::ADDEX; e.a = a; e.b = b; e.alpha = alpha; e.beta = beta; e.s = s;
  return e;
}
static MatExpr gemm_expr(int flags, const Mat& a, const Mat& b, double alpha, const Mat& c = Mat(), double beta = 1) {
  MatExpr e;
  e.kind = MatExpr::GEMM; e.gemm_flags = flags; e.a = a; e.b = b; e.alpha = alpha; e.c = c; e.beta = beta;
  return e;
}
static bool is_ident(const MatExpr& e) { return e.kind == MatExpr::IDENT; }
static bool is_t(const MatExpr& e) { return e.kind == MatExpr::TRANSPOSE; }
static bool is_addex(const MatExpr& e) { return e.kind == MatExpr::ADDEX; }
static bool is_scaled(const MatExpr& e) { return is_addex(e) && (!e.b.data || e.beta == 0) && e.s.isZero(); }
static bool is_matprod(const MatExpr& e) { return e.kind == MatExpr::GEMM && (!e.c.data || e.beta == 0); }

template <typename T> static void addex_eval(const MatExpr& e, Mat& out) {
  const Mat& a = e.a;
  out.create(a.rows, a.cols, a.type());
  const int n = a.cols * a.channels();
  const T s0 = (T)e.s[0];
  for (int y = 0; y < a.rows; ++y) {
    const T* pa = a.ptr<T>(y);
    const T* pb = e.b.data ? e.b.ptr<T>(y) : nullptr;
    T* d = out.ptr<T>(y);
    for (int x = 0; x < n; ++x) {
      if (pb) {
        T v;
        if (e.alpha == 1) {
          if (e.beta == 1) v = pa[x] + pb[x];
          else if (e.beta == -1) v = pa[x] - pb[x];
          else v = pb[x] * (T)e.beta + pa[x];               // scaleAdd(b, beta, a)
        } else if (e.beta == 1) {
          if (e.alpha == -1) v = pb[x] - pa[x];
          else v = pa[x] * (T)e.alpha + pb[x];              // scaleAdd(a, alpha, b)
        } else {
          v = pa[x] * (T)e.alpha + pb[x] * (T)e.beta;       // addWeighted (gamma added below)
        }
        d[x] = e.s.isZero() ? v : v + s0;
      } else {
        d[x] = pa[x] * (T)e.alpha + s0;                     // convertTo(alpha, s)
      }
    }
  }
}

void MatExpr::assign(Mat& m) const {
  if (kind == IDENT) { m = a; return; }
  Mat r;
  switch (kind) {
    case IDENT: break;
    case TRANSPOSE:
      transpose(a, r);
      if (alpha != 1) r.convertTo(r, -1, alpha);
      break;
    case GEMM: gemm(a, b, alpha, c, beta, r, gemm_flags); break;
    case ADDEX:
      if (b.data && (b.rows != a.rows || b.cols != a.cols || b.type() != a.type())) shim_fail("A+B: size/type", __FILE__, __LINE__);
      if (a.depth() == CV_32F) addex_eval<float>(*this, r);
      else if (a.depth() == CV_64F) addex_eval<double>(*this, r);
      else shim_fail("matrix expression on a non-floating-point Mat", __FILE__, __LINE__);
      break;
    case INIT: {
      r.create(init_rows, init_cols, init_type);
      r.setTo(Scalar(init_what == '1' ? alpha : 0.0));
      if (init_what == 'I') {
        Mat one(1, 1, init_type);
        one.setTo(Scalar(alpha));
        for (int i = 0; i < std::min(init_rows, init_cols); ++i)
          std::memcpy(r.ptr(i) + (size_t)i * r.elemSize(), one.ptr(0), r.elemSize());
      }
      break;
    }
    case INV: shim_fail("Mat::inv() is not available in the shim", __FILE__, __LINE__);
  }
  // like OpenCV (dst.create is a no-op for a matching header): write in place when m already has the shape and type
  if (m.data && m.data != r.data && m.rows == r.rows && m.cols == r.cols && m.type() == r.type()) r.copyTo(m);
  else m = r;
}

MatExpr Mat::t() const { MatExpr e; e.kind = MatExpr::TRANSPOSE; e.a = *this; e.alpha = 1; return e; }
MatExpr MatExpr::t() const {
  if (kind == TRANSPOSE && alpha == 1) return MatExpr(a);
  return ((Mat)*this).t();
}
MatExpr Mat::inv(int) const { MatExpr e; e.kind = MatExpr::INV; e.a = *this; return e; }
MatExpr Mat::mul(const Mat&, double) const { shim_fail("Mat::mul is not available in the shim", __FILE__, __LINE__); }
Mat Mat::cross(const Mat&) const { shim_fail("Mat::cross is not available in the shim", __FILE__, __LINE__); }

// scale an expression: the per-kind rules
static MatExpr scale_expr(const MatExpr& e, double s) {
  MatExpr r = e;
  switch (e.kind) {
    case MatExpr::ADDEX: r.alpha *= s; r.beta *= s; for (int i = 0; i < 4; ++i) r.s[i] *= s; return r;
    case MatExpr::TRANSPOSE: r.alpha *= s; return r;
    case MatExpr::GEMM: r.alpha *= s; r.beta *= s; return r;
    case MatExpr::INIT: r.alpha *= s; return r;
    default: return addex((Mat)e, Mat(), s, 0);
  }
}

// product: fold a transposition / scale of either side into the gemm, otherwise materialise the side
static MatExpr matmul(const MatExpr& e1, const MatExpr& e2) {
  double scale = 1;
  int flags = 0;
  Mat m1, m2;
  if (is_t(e1)) { flags |= GEMM_1_T; scale = e1.alpha; m1 = e1.a; }
  else if (is_scaled(e1)) { scale = e1.alpha; m1 = e1.a; }
  else m1 = (Mat)e1;
  if (is_t(e2)) { flags |= GEMM_2_T; scale *= e2.alpha; m2 = e2.a; }
  else if (is_scaled(e2)) { scale *= e2.alpha; m2 = e2.a; }
  else m2 = (Mat)e2;
  return gemm_expr(flags, m1, m2, scale);
}

// sum / difference (sign = +1 / -1 applied to e2)
static MatExpr addsub(const MatExpr& e1, const MatExpr& e2, double sign) {
  const bool i1 = is_ident(e1), i2 = is_ident(e2);
  const double alpha1 = i1 ? 1 : e1.alpha, alpha2 = (i2 ? 1 : e2.alpha) * sign;
  if (is_matprod(e1) && (i2 || is_scaled(e2) || is_t(e2)))
    return gemm_expr((e1.gemm_flags & ~GEMM_3_T) | (is_t(e2) ? GEMM_3_T : 0), e1.a, e1.b, alpha1, e2.a, alpha2);
  if (is_matprod(e2) && (i1 || is_scaled(e1) || is_t(e1)))
    return gemm_expr((e2.gemm_flags & ~GEMM_3_T) | (is_t(e1) ? GEMM_3_T : 0), e2.a, e2.b, alpha2, e1.a, alpha1);
  double alpha = 1, beta = sign;
  Scalar s;
  Mat m1, m2;
  if (is_addex(e1) && (!e1.b.data || e1.beta == 0)) { m1 = e1.a; alpha = e1.alpha; s = e1.s; } else m1 = (Mat)e1;
  if (is_addex(e2) && (!e2.b.data || e2.beta == 0)) {
    m2 = e2.a; beta = sign * e2.alpha;
    for (int i = 0; i < 4; ++i) s[i] += sign * e2.s[i];
  } else m2 = (Mat)e2;
  return addex(m1, m2, alpha, beta, s);
}

static MatExpr negate(const MatExpr& e) {
  if (is_addex(e)) {
    MatExpr r = e;
    r.alpha = -r.alpha; r.beta = -r.beta;
    for (int i = 0; i < 4; ++i) r.s[i] = -r.s[i];
    return r;
  }
  return addex((Mat)e, Mat(), -1, 0);  // materialise, then scale by -1
}

MatExpr operator+(const Mat& a, const Mat& b) { return addex(a, b, 1, 1); }
MatExpr operator+(const MatExpr& e, const Mat& b) { return addsub(e, MatExpr(b), 1); }
MatExpr operator+(const Mat& a, const MatExpr& e) { return addsub(e, MatExpr(a), 1); }
MatExpr operator+(const MatExpr& e1, const MatExpr& e2) { return addsub(e1, e2, 1); }
MatExpr operator-(const Mat& a, const Mat& b) { return addex(a, b, 1, -1); }
MatExpr operator-(const MatExpr& e, const Mat& b) { return addsub(e, MatExpr(b), -1); }
MatExpr operator-(const Mat& a, const MatExpr& e) { return addsub(MatExpr(a), e, -1); }
MatExpr operator-(const MatExpr& e1, const MatExpr& e2) { return addsub(e1, e2, -1); }
MatExpr operator-(const Mat& m) { return addex(m, Mat(), -1, 0); }
MatExpr operator-(const MatExpr& e) { return negate(e); }
MatExpr operator*(const Mat& a, const Mat& b) { return gemm_expr(0, a, b, 1); }
MatExpr operator*(const MatExpr& e, const Mat& b) { return matmul(e, MatExpr(b)); }
MatExpr operator*(const Mat& a, const MatExpr& e) { return matmul(MatExpr(a), e); }
MatExpr operator*(const MatExpr& e1, const MatExpr& e2) { return matmul(e1, e2); }
MatExpr operator*(const Mat& a, double s) { return addex(a, Mat(), s, 0); }
MatExpr operator*(double s, const Mat& a) { return addex(a, Mat(), s, 0); }
MatExpr operator*(const MatExpr& e, double s) { return scale_expr(e, s); }
MatExpr operator*(double s, const MatExpr& e) { return scale_expr(e, s); }
MatExpr operator/(const Mat& a, double s) { return addex(a, Mat(), 1. / s, 0); }
MatExpr operator/(const MatExpr& e, double s) { return scale_expr(e, 1. / s); }

// ---------------------------------------------------------------- reductions
template <typename T> static double dot_rows(const Mat& a, const Mat& b) {
  double r = 0;
  if (a.isContinuous() && b.isContinuous()) {
    const T* p = a.ptr<T>(0);
    const T* q = b.ptr<T>(0);
    const int n = (int)a.total() * a.channels();
    int i = 0;
    for (; i <= n - 4; i += 4)
      r += (double)p[i] * q[i] + (double)p[i + 1] * q[i + 1] + (double)p[i + 2] * q[i + 2] + (double)p[i + 3] * q[i + 3];
    for (; i < n; ++i) r += (double)p[i] * q[i];
    return r;
  }
  const int n = a.cols * a.channels();
  for (int y = 0; y < a.rows; ++y) {  // one call per row, partial results added in row order
    const T* p = a.ptr<T>(y);
    const T* q = b.ptr<T>(y);
    double rr = 0;
    int i = 0;
    for (; i <= n - 4; i += 4)
      rr += (double)p[i] * q[i] + (double)p[i + 1] * q[i + 1] + (double)p[i + 2] * q[i + 2] + (double)p[i + 3] * q[i + 3];
    for (; i < n; ++i) rr += (double)p[i] * q[i];
    r += rr;
  }
  return r;
}

double Mat::dot(const Mat& m) const {
  if (m.type() != type() || m.rows != rows || m.cols != cols) shim_fail("dot: size/type", __FILE__, __LINE__);
  if (depth() == CV_32F) return dot_rows<float>(*this, m);
  if (depth() == CV_64F) return dot_rows<double>(*this, m);
  shim_fail("dot: unsupported depth", __FILE__, __LINE__);
}

template <typename T> static double norm_rows(const Mat& a, const Mat* b, int normType) {
  double r = 0;
  const int n = a.cols * a.channels();
  for (int y = 0; y < a.rows; ++y) {
    const T* p = a.ptr<T>(y);
    const T* q = b ? b->ptr<T>(y) : nullptr;
    for (int x = 0; x < n; ++x) {
      const double v = q ? (double)p[x] - (double)q[x] : (double)p[x];
      if (normType == NORM_L1) r += std::fabs(v);
      else if (normType == NORM_L2) r += v * v;
      else r = std::max(r, std::fabs(v));
    }
  }
  return normType == NORM_L2 ? std::sqrt(r) : r;
}

static double norm_impl(const Mat& a, const Mat* b, int normType) {
  if (b && (b->type() != a.type() || b->rows != a.rows || b->cols != a.cols)) shim_fail("norm: size/type", __FILE__, __LINE__);
  switch (a.depth()) {
    case CV_8U: return norm_rows<uchar>(a, b, normType);
    case CV_32F: return norm_rows<float>(a, b, normType);
    case CV_64F: return norm_rows<double>(a, b, normType);
    default: shim_fail("norm: unsupported depth", __FILE__, __LINE__);
  }
}
double norm(InputArray a, int normType) { return norm_impl(a.getMat(), nullptr, normType); }
double norm(InputArray a, InputArray b, int normType) { Mat mb = b.getMat(); return norm_impl(a.getMat(), &mb, normType); }

// ---------------------------------------------------------------- the pixel primitives (oracle/orb_prims.cc)
void resize(InputArray _src, OutputArray _dst, Size dsize, double, double, int interpolation) {
  Mat src = _src.getMat();
  if (src.type() != CV_8UC1 || interpolation != INTER_LINEAR || dsize.area() == 0)
    shim_fail("resize: only CV_8UC1 INTER_LINEAR with an explicit size", __FILE__, __LINE__);
  _dst.create(dsize, CV_8UC1);
  Mat dst = _dst.getMatRef();
  orc::resize_linear_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

static inline int reflect101(int p, int len) {
  if (len == 1) return 0;
  while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
  return p;
}

void copyMakeBorder(InputArray _src, OutputArray _dst, int top, int bottom, int left, int right, int borderType,
                    const Scalar&) {
  Mat src = _src.getMat();
  // without BORDER_ISOLATED OpenCV would read real pixels around a sub-view; the shim's Mat does not track its parent,
  // and the reference only passes a whole image on that branch (:1149-1150)
  if ((borderType & ~BORDER_ISOLATED) != BORDER_REFLECT_101 || src.type() != CV_8UC1)
    shim_fail("copyMakeBorder: only CV_8UC1 BORDER_REFLECT_101", __FILE__, __LINE__);
  _dst.create(src.rows + top + bottom, src.cols + left + right, src.type());
  Mat dst = _dst.getMatRef();
  const int w = src.cols, h = src.rows;
  for (int y = 0; y < h; ++y) {  // interior (a no-op when src is already the interior view of dst) + left/right
    uchar* d = dst.ptr(y + top);
    if (d + left != src.ptr(y)) std::memmove(d + left, src.ptr(y), (size_t)w);
    for (int x = 0; x < left; ++x) d[x] = d[left + reflect101(x - left, w)];
    for (int x = 0; x < right; ++x) d[left + w + x] = d[left + reflect101(w + x, w)];
  }
  for (int y = 0; y < top; ++y) std::memcpy(dst.ptr(y), dst.ptr(top + reflect101(y - top, h)), (size_t)dst.cols);
  for (int y = 0; y < bottom; ++y) std::memcpy(dst.ptr(top + h + y), dst.ptr(top + reflect101(h + y, h)), (size_t)dst.cols);
}

void FAST(InputArray _image, std::vector<KeyPoint>& keypoints, int threshold, bool nonmaxSuppression) {
  Mat img = _image.getMat();
  if (img.type() != CV_8UC1) shim_fail("FAST: CV_8UC1 only", __FILE__, __LINE__);
  std::vector<orc::FastPoint> pts;
  orc::fast9_16(img.data, img.cols, img.rows, (int)img.step, threshold, nonmaxSuppression, pts);
  keypoints.clear();
  keypoints.reserve(pts.size());
  for (const orc::FastPoint& p : pts) keypoints.push_back(KeyPoint((float)p.x, (float)p.y, 7.f, -1, (float)p.score));
}

void GaussianBlur(InputArray _src, OutputArray _dst, Size ksize, double sigmaX, double sigmaY, int borderType) {
  Mat src = _src.getMat();
  if (src.type() != CV_8UC1 || ksize.width != 7 || ksize.height != 7 || sigmaX != 2 || sigmaY != 2 ||
      borderType != BORDER_REFLECT_101)
    shim_fail("GaussianBlur: only CV_8UC1 7x7 sigma 2 BORDER_REFLECT_101", __FILE__, __LINE__);
  _dst.create(src.rows, src.cols, src.type());
  Mat dst = _dst.getMatRef();
  orc::gaussian_blur7_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, (int)dst.step);  // safe in place
}

float fastAtan2(float y, float x) { return orb_fast_atan2_deg(y, x); }

void undistortPoints(InputArray, OutputArray, InputArray, InputArray, InputArray, InputArray) {
  shim_fail("cv::undistortPoints is not on the rectified-stereo path (zero distortion) and not in the shim", __FILE__, __LINE__);
}

void KeyPointsFilter::retainBest(std::vector<KeyPoint>&, int) {
  shim_fail("KeyPointsFilter::retainBest is only called by dead code (ComputeKeyPointsOld)", __FILE__, __LINE__);
}

}  // namespace cv
