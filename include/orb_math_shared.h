/* orb_math_shared.h — float/double arithmetic that must give the SAME BITS on
 * the host (g++ -ffp-contract=off) and on the device (nvcc).  Every operation
 * is an explicitly rounded IEEE add/mul/div so that neither compiler can fuse
 * a multiply-add.
 *
 *  - orb_fast_atan2_deg : the polynomial of cv::fastAtan2 (OpenCV 4.13
 *    semantics), called at src/ORBextractor.cc:103.
 *  - orb_sincos_deg     : a = cos(angle*pi/180), b = sin(...) for
 *    src/ORBextractor.cc:112-113.  Documented deviation 3:
 *    evaluated in double with a fixed range reduction + fixed polynomial and
 *    rounded to float, instead of libm cosf/sinf whose bits are machine
 *    dependent.
 *  - orb_log_f32        : logf for MapPoint::PredictScale (src/MapPoint.cc:406-444),
 *    same construction.
 */
#ifndef ORB_MATH_SHARED_H_
#define ORB_MATH_SHARED_H_

#if defined(__CUDA_ARCH__)
#define ORB_HD __host__ __device__ __forceinline__
#define ORB_FMUL(a, b) __fmul_rn((a), (b))
#define ORB_FADD(a, b) __fadd_rn((a), (b))
#define ORB_FSUB(a, b) __fsub_rn((a), (b))
#define ORB_FDIV(a, b) __fdiv_rn((a), (b))
#define ORB_DMUL(a, b) __dmul_rn((a), (b))
#define ORB_DADD(a, b) __dadd_rn((a), (b))
#define ORB_DSUB(a, b) __dsub_rn((a), (b))
#define ORB_DDIV(a, b) __ddiv_rn((a), (b))
#else
#if defined(__CUDACC__)
#define ORB_HD __host__ __device__ inline
#else
#define ORB_HD static inline
#endif
#define ORB_FMUL(a, b) ((float)(a) * (float)(b))
#define ORB_FADD(a, b) ((float)(a) + (float)(b))
#define ORB_FSUB(a, b) ((float)(a) - (float)(b))
#define ORB_FDIV(a, b) ((float)(a) / (float)(b))
#define ORB_DMUL(a, b) ((double)(a) * (double)(b))
#define ORB_DADD(a, b) ((double)(a) + (double)(b))
#define ORB_DSUB(a, b) ((double)(a) - (double)(b))
#define ORB_DDIV(a, b) ((double)(a) / (double)(b))
#endif

/* cv::fastAtan2(y, x) in degrees, fp32, no FMA. */
ORB_HD float orb_fast_atan2_deg(float y, float x) {
  const float k = (float)(180.0 / 3.14159265358979323846);
  const float p1 = ORB_FMUL(0.9997878412794807f, k);
  const float p3 = ORB_FMUL(-0.3258083974640975f, k);
  const float p5 = ORB_FMUL(0.1555786518463281f, k);
  const float p7 = ORB_FMUL(-0.04432655554792128f, k);
  const float eps = (float)2.2204460492503131e-16;
  float ax = x < 0.f ? -x : x;
  float ay = y < 0.f ? -y : y;
  float a, c, c2;
  if (ax >= ay) {
    c = ORB_FDIV(ay, ORB_FADD(ax, eps));
    c2 = ORB_FMUL(c, c);
    a = ORB_FMUL(ORB_FADD(ORB_FMUL(ORB_FADD(ORB_FMUL(ORB_FADD(ORB_FMUL(p7, c2), p5), c2), p3), c2), p1), c);
  } else {
    c = ORB_FDIV(ax, ORB_FADD(ay, eps));
    c2 = ORB_FMUL(c, c);
    a = ORB_FSUB(90.f,
                 ORB_FMUL(ORB_FADD(ORB_FMUL(ORB_FADD(ORB_FMUL(ORB_FADD(ORB_FMUL(p7, c2), p5), c2), p3), c2), p1), c));
  }
  if (x < 0.f) a = ORB_FSUB(180.f, a);
  if (y < 0.f) a = ORB_FSUB(360.f, a);
  return a;
}

/* sin and cos of a float angle given in DEGREES, as the reference computes
 * them: angle_rad = (float)angle_deg * factorPI (fp32 product,
 * src/ORBextractor.cc:107,112), then cos/sin of that float, rounded to float.
 * The cos/sin themselves are evaluated in double (Cody-Waite reduction by
 * pi/2 + degree-13/14 polynomials; |error| < 1e-15, i.e. the float result is
 * the correctly rounded one except with probability ~1e-8). */
ORB_HD void orb_sincos_rad(float ang, float* cos_out, float* sin_out);
ORB_HD void orb_sincos_deg(float angle_deg, float* cos_out, float* sin_out) {
  const float factorPI = (float)(3.14159265358979323846 / 180.0);
  orb_sincos_rad(ORB_FMUL(angle_deg, factorPI), cos_out, sin_out);
}
/* cos / sin of a float angle in RADIANS, rounded to float (the body orb_sincos_deg evaluates). */
ORB_HD void orb_sincos_rad(float ang, float* cos_out, float* sin_out) {
  double x = (double)ang;
  double ax = x < 0.0 ? -x : x;
  /* quadrant: k = round-to-nearest(ax * 2/pi) via truncation of +0.5 */
  int k = (int)ORB_DADD(ORB_DMUL(ax, 0.63661977236758134308), 0.5);
  double kd = (double)k;
  const double PIO2_HI = 1.57079632673412561417e+00; /* first 33 bits of pi/2 */
  const double PIO2_LO = 6.07710050650619224932e-11; /* pi/2 - PIO2_HI */
  double r = ORB_DSUB(ORB_DSUB(ax, ORB_DMUL(kd, PIO2_HI)), ORB_DMUL(kd, PIO2_LO));
  double z = ORB_DMUL(r, r);
  /* sin(r) = r + r*z*(S1 + z*(S2 + ... )) */
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
               S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
               S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  double ps = ORB_DADD(S5, ORB_DMUL(z, S6));
  ps = ORB_DADD(S4, ORB_DMUL(z, ps));
  ps = ORB_DADD(S3, ORB_DMUL(z, ps));
  ps = ORB_DADD(S2, ORB_DMUL(z, ps));
  ps = ORB_DADD(S1, ORB_DMUL(z, ps));
  double s = ORB_DADD(r, ORB_DMUL(ORB_DMUL(r, z), ps));
  /* cos(r) = 1 - z/2 + z*z*(C1 + z*(C2 + ...)) */
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
               C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
               C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  double pc = ORB_DADD(C5, ORB_DMUL(z, C6));
  pc = ORB_DADD(C4, ORB_DMUL(z, pc));
  pc = ORB_DADD(C3, ORB_DMUL(z, pc));
  pc = ORB_DADD(C2, ORB_DMUL(z, pc));
  pc = ORB_DADD(C1, ORB_DMUL(z, pc));
  double c = ORB_DADD(ORB_DSUB(1.0, ORB_DMUL(0.5, z)), ORB_DMUL(ORB_DMUL(z, z), pc));
  double sn, cs;
  switch (k & 3) {
    case 0: sn = s; cs = c; break;
    case 1: sn = c; cs = -s; break;
    case 2: sn = -s; cs = -c; break;
    default: sn = -c; cs = s; break;
  }
  if (x < 0.0) sn = -sn;
  *cos_out = (float)cs;
  *sin_out = (float)sn;
}

/* Natural logarithm of a positive finite float, rounded to float: the value
 * MapPoint::PredictScale needs (`log(ratio)` with a float argument is logf,
 * src/MapPoint.cc:414,437).  Evaluated in double (x = m*2^e, m in
 * [sqrt(1/2), sqrt(2)); log m = 2 atanh((m-1)/(m+1)), series to s^10;
 * |error| < 1e-16 relative), so the float result is the correctly rounded one
 * except with probability ~1e-8 and — the point — has the same bits on host
 * and device.  x <= 0, inf and nan are not handled (callers gate them). */
ORB_HD float orb_log_f32(float xf) {
  double x = (double)xf;
  long long bits;
#if defined(__CUDA_ARCH__)
  bits = __double_as_longlong(x);
#else
  __builtin_memcpy(&bits, &x, 8);
#endif
  int e = (int)((bits >> 52) & 0x7FF) - 1023;
  long long mb = (bits & 0x000FFFFFFFFFFFFFLL) | 0x3FF0000000000000LL;
  double m;
#if defined(__CUDA_ARCH__)
  m = __longlong_as_double(mb);
#else
  __builtin_memcpy(&m, &mb, 8);
#endif
  if (m > 1.41421356237309514547) { m = ORB_DMUL(m, 0.5); e += 1; }
  const double f = ORB_DDIV(ORB_DSUB(m, 1.0), ORB_DADD(m, 1.0));
  const double s = ORB_DMUL(f, f);
  double p = 1.0 / 21.0;
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 19.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 17.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 15.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 13.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 11.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 9.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 7.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 5.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0 / 3.0);
  p = ORB_DADD(ORB_DMUL(p, s), 1.0);
  const double r = ORB_DMUL(ORB_DMUL(2.0, f), p);
  const double ed = (double)e;
  const double LN2_HI = 6.93147180369123816490e-01, LN2_LO = 1.90821492927058770002e-10;
  return (float)ORB_DADD(ORB_DMUL(ed, LN2_HI), ORB_DADD(r, ORB_DMUL(ed, LN2_LO)));
}

/* MapPoint::PredictScale (src/MapPoint.cc:406-444): ratio = mfMaxDistance / dist,
 * nScale = ceil(log(ratio) / mfLogScaleFactor) clamped to [0, nlevels-1].  The clamp
 * is applied in float so that inf / nan / huge values convert identically on
 * host and device (ratio <= 0 or nan -> 0). */
ORB_HD int orb_predict_scale(float max_distance, float current_dist, float log_scale_factor, int nlevels) {
  const float ratio = ORB_FDIV(max_distance, current_dist);
  if (!(ratio > 0.f)) return 0;
  if (ratio > 3.0e38f) return nlevels - 1;
  const float q = ORB_FDIV(orb_log_f32(ratio), log_scale_factor);
  if (!(q > -1.f)) return 0;               /* ceil(q) <= -1, or nan */
  if (q > (float)(nlevels - 1)) return nlevels - 1;
  const int t = (int)q;                    /* trunc; q in (-1, nlevels-1] */
  const int c = ((float)t < q) ? t + 1 : t;
  return c < 0 ? 0 : c;
}


/* cv::undistortPoints(src, dst, K, D, noArray(), K) for ONE point (OpenCV 4.13 semantics: cvUndistortPointsInternal with
 * the default TermCriteria(MAX_ITER, 5, 0.01) = five fixed-point iterations, R = I, P = K), as Frame::UndistortKeyPoints
 * (src/Frame.cc:429-460) and Frame::ComputeImageBounds (:462-494) call it.  All arithmetic in double, in OpenCV's
 * order, no FMA; checked bit-for-bit against python-cv2 4.13 (tests/golden/undistort.npz).  k = {k1, k2, p1, p2, k3}
 * (mDistCoef: 4 or 5 coefficients, Tracking.cc:72-90; k3 = 0 when absent). */
ORB_HD void orb_undistort_point(float u, float v, float ffx, float ffy, float fcx, float fcy, const float* kf,
                                float* xu, float* yu) {
  const double fx = (double)ffx, fy = (double)ffy, cx = (double)fcx, cy = (double)fcy;
  const double k0 = (double)kf[0], k1 = (double)kf[1], k2 = (double)kf[2], k3 = (double)kf[3], k4 = (double)kf[4];
  const double ifx = ORB_DDIV(1.0, fx), ify = ORB_DDIV(1.0, fy);
  double x = ORB_DMUL(ORB_DSUB((double)u, cx), ifx), y = ORB_DMUL(ORB_DSUB((double)v, cy), ify);
  const double x0 = x, y0 = y;
  for (int j = 0; j < 5; ++j) {
    const double r2 = ORB_DADD(ORB_DMUL(x, x), ORB_DMUL(y, y));
    /* icdist = (1 + ((k[7]*r2 + k[6])*r2 + k[5])*r2) / (1 + ((k[4]*r2 + k[1])*r2 + k[0])*r2), k[5..7] = 0 */
    const double num = ORB_DADD(1.0, ORB_DMUL(ORB_DADD(ORB_DMUL(ORB_DADD(ORB_DMUL(0.0, r2), 0.0), r2), 0.0), r2));
    const double den = ORB_DADD(1.0, ORB_DMUL(ORB_DADD(ORB_DMUL(ORB_DADD(ORB_DMUL(k4, r2), k1), r2), k0), r2));
    const double icdist = ORB_DDIV(num, den);
    if (icdist < 0.0) {  /* "test: undistortPoints.regression_14583" */
      x = x0;
      y = y0;
      break;
    }
    /* deltaX = 2*k[2]*x*y + k[3]*(r2 + 2*x*x) + k[8]*r2 + k[9]*r2*r2 (k[8..11] = 0) */
    const double xy2a = ORB_DMUL(ORB_DMUL(ORB_DMUL(2.0, k2), x), y);
    const double dx = ORB_DADD(ORB_DADD(ORB_DADD(xy2a, ORB_DMUL(k3, ORB_DADD(r2, ORB_DMUL(ORB_DMUL(2.0, x), x)))),
                                        ORB_DMUL(0.0, r2)), ORB_DMUL(ORB_DMUL(0.0, r2), r2));
    const double xy2b = ORB_DMUL(ORB_DMUL(ORB_DMUL(2.0, k3), x), y);
    const double dy = ORB_DADD(ORB_DADD(ORB_DADD(ORB_DMUL(k2, ORB_DADD(r2, ORB_DMUL(ORB_DMUL(2.0, y), y))), xy2b),
                                        ORB_DMUL(0.0, r2)), ORB_DMUL(ORB_DMUL(0.0, r2), r2));
    x = ORB_DMUL(ORB_DSUB(x0, dx), icdist);
    y = ORB_DMUL(ORB_DSUB(y0, dy), icdist);
  }
  /* R = I, P = K: xx = fx*x + 0*y + cx, yy = 0*x + fy*y + cy, ww = 1 / (0*x + 0*y + 1) */
  const double xx = ORB_DADD(ORB_DADD(ORB_DMUL(fx, x), ORB_DMUL(0.0, y)), cx);
  const double yy = ORB_DADD(ORB_DADD(ORB_DMUL(0.0, x), ORB_DMUL(fy, y)), cy);
  const double ww = ORB_DDIV(1.0, ORB_DADD(ORB_DADD(ORB_DMUL(0.0, x), ORB_DMUL(0.0, y)), 1.0));
  *xu = (float)ORB_DMUL(xx, ww);
  *yu = (float)ORB_DMUL(yy, ww);
}

#endif /* ORB_MATH_SHARED_H_ */
