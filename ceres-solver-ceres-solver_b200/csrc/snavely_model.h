// Snavely / Bundler camera model with forward-mode derivatives, shared by the
// host evaluator (synthetic_bal.cc) and the device evaluator (lm_kernels.cuh).
//
// Reference: examples/snavely_reprojection_error.h:57-92 (camera model),
// include/ceres/rotation.h:792-850 (AngleAxisRotatePoint), and ceres::Jet
// (include/ceres/jet.h) whose role the 12-partial dual number below plays:
// 9 camera + 3 point parameters, exactly the partials AutoDiffCostFunction
// <SnavelyReprojectionError, 2, 9, 3> propagates.
#ifndef CB200_SNAVELY_MODEL_H_
#define CB200_SNAVELY_MODEL_H_

#include <cmath>
#include <cstring>

#ifdef __CUDACC__
#define CB200_HD __host__ __device__ __forceinline__
#else
#define CB200_HD inline
#endif

namespace cb200_model {

// Forward-mode dual number with 12 partials: 9 camera + 3 point parameters.
struct Dual {
  double v;
  double d[12];
};
CB200_HD Dual Const(double c) { Dual r; r.v = c; for (int i = 0; i < 12; ++i) r.d[i] = 0.0; return r; }
CB200_HD Dual Var(double c, int k) { Dual r = Const(c); r.d[k] = 1.0; return r; }
CB200_HD Dual operator+(const Dual& a, const Dual& b) {
  Dual r; r.v = a.v + b.v; for (int i = 0; i < 12; ++i) r.d[i] = a.d[i] + b.d[i]; return r;
}
CB200_HD Dual operator-(const Dual& a, const Dual& b) {
  Dual r; r.v = a.v - b.v; for (int i = 0; i < 12; ++i) r.d[i] = a.d[i] - b.d[i]; return r;
}
CB200_HD Dual operator-(const Dual& a) {
  Dual r; r.v = -a.v; for (int i = 0; i < 12; ++i) r.d[i] = -a.d[i]; return r;
}
CB200_HD Dual operator*(const Dual& a, const Dual& b) {
  Dual r; r.v = a.v * b.v;
  for (int i = 0; i < 12; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i];
  return r;
}
CB200_HD Dual operator/(const Dual& a, const Dual& b) {
  Dual r; const double inv = 1.0 / b.v; r.v = a.v * inv;
  for (int i = 0; i < 12; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * inv;
  return r;
}
CB200_HD Dual Scale(const Dual& a, double f, double df) {  // chain rule helper
  Dual r; r.v = f; for (int i = 0; i < 12; ++i) r.d[i] = df * a.d[i]; return r;
}
CB200_HD Dual Sqrt(const Dual& a) { const double s = sqrt(a.v); return Scale(a, s, 0.5 / s); }
CB200_HD Dual Sin(const Dual& a) { return Scale(a, sin(a.v), cos(a.v)); }
CB200_HD Dual Cos(const Dual& a) { return Scale(a, cos(a.v), -sin(a.v)); }

template <typename T> CB200_HD T Lift(double c);
template <> CB200_HD double Lift<double>(double c) { return c; }
template <> CB200_HD Dual Lift<Dual>(double c) { return Const(c); }
CB200_HD double Val(double a) { return a; }
CB200_HD double Val(const Dual& a) { return a.v; }
CB200_HD double Sqrt(double a) { return sqrt(a); }
CB200_HD double Sin(double a) { return sin(a); }
CB200_HD double Cos(double a) { return cos(a); }

// Rodrigues rotation of pt by the angle-axis vector aa, with the first-order
// branch near zero so derivatives stay finite (rotation.h:792-850).
template <typename T>
CB200_HD void RotateByAngleAxis(const T* aa, const T* pt, T* out) {
  const T theta2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
  if (Val(theta2) > 2.220446049250313e-16) {
    const T theta = Sqrt(theta2);
    const T c = Cos(theta), s = Sin(theta);
    const T inv = Lift<T>(1.0) / theta;
    const T w[3] = {aa[0] * inv, aa[1] * inv, aa[2] * inv};
    const T wxp[3] = {w[1] * pt[2] - w[2] * pt[1], w[2] * pt[0] - w[0] * pt[2],
                      w[0] * pt[1] - w[1] * pt[0]};
    const T k = (w[0] * pt[0] + w[1] * pt[1] + w[2] * pt[2]) * (Lift<T>(1.0) - c);
    for (int i = 0; i < 3; ++i) out[i] = pt[i] * c + wxp[i] * s + w[i] * k;
  } else {
    out[0] = pt[0] + (aa[1] * pt[2] - aa[2] * pt[1]);
    out[1] = pt[1] + (aa[2] * pt[0] - aa[0] * pt[2]);
    out[2] = pt[2] + (aa[0] * pt[1] - aa[1] * pt[0]);
  }
}

// Snavely / Bundler camera: 3 angle-axis, 3 translation, focal, k1, k2
// (snavely_reprojection_error.h:57-92). Returns the predicted image position.
template <typename T>
CB200_HD void Project(const T* cam, const T* pt, T* pred) {
  T p[3];
  RotateByAngleAxis(cam, pt, p);
  p[0] = p[0] + cam[3]; p[1] = p[1] + cam[4]; p[2] = p[2] + cam[5];
  const T xp = -(p[0] / p[2]);
  const T yp = -(p[1] / p[2]);
  const T r2 = xp * xp + yp * yp;
  const T distortion = Lift<T>(1.0) + r2 * (cam[7] + cam[8] * r2);
  pred[0] = cam[6] * distortion * xp;
  pred[1] = cam[6] * distortion * yp;
}


}  // namespace cb200_model

#endif  // CB200_SNAVELY_MODEL_H_
