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

// Forward-mode dual number with N partials (the role of ceres::Jet<double, N>).
template <int N>
struct DualN {
  double v;
  double d[N];
};
// 12 partials: 9 camera + 3 point parameters, what AutoDiffCostFunction<.., 2, 9, 3> propagates.
using Dual = DualN<12>;

template <int N> CB200_HD DualN<N> ConstN(double c) {
  DualN<N> r; r.v = c;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = 0.0;
  return r;
}
template <int N> CB200_HD DualN<N> VarN(double c, int k) {
  DualN<N> r; r.v = c;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = (i == k) ? 1.0 : 0.0;
  return r;
}
CB200_HD Dual Const(double c) { return ConstN<12>(c); }
CB200_HD Dual Var(double c, int k) { return VarN<12>(c, k); }
template <int N> CB200_HD DualN<N> operator+(const DualN<N>& a, const DualN<N>& b) {
  DualN<N> r; r.v = a.v + b.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.d[i] + b.d[i];
  return r;
}
template <int N> CB200_HD DualN<N> operator-(const DualN<N>& a, const DualN<N>& b) {
  DualN<N> r; r.v = a.v - b.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.d[i] - b.d[i];
  return r;
}
template <int N> CB200_HD DualN<N> operator-(const DualN<N>& a) {
  DualN<N> r; r.v = -a.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = -a.d[i];
  return r;
}
template <int N> CB200_HD DualN<N> operator*(const DualN<N>& a, const DualN<N>& b) {
  DualN<N> r; r.v = a.v * b.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b.v + a.v * b.d[i];
  return r;
}
template <int N> CB200_HD DualN<N> operator/(const DualN<N>& a, const DualN<N>& b) {
  DualN<N> r; const double inv = 1.0 / b.v; r.v = a.v * inv;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * inv;
  return r;
}
template <int N> CB200_HD DualN<N> Scale(const DualN<N>& a, double f, double df) {  // chain rule helper
  DualN<N> r; r.v = f;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = df * a.d[i];
  return r;
}
template <int N> CB200_HD DualN<N> Sqrt(const DualN<N>& a) { const double s = sqrt(a.v); return Scale(a, s, 0.5 / s); }
template <int N> CB200_HD DualN<N> Sin(const DualN<N>& a) { return Scale(a, sin(a.v), cos(a.v)); }
template <int N> CB200_HD DualN<N> Cos(const DualN<N>& a) { return Scale(a, cos(a.v), -sin(a.v)); }

template <typename T> struct LiftImpl;
template <> struct LiftImpl<double> { static CB200_HD double Get(double c) { return c; } };
template <int N> struct LiftImpl<DualN<N>> { static CB200_HD DualN<N> Get(double c) { return ConstN<N>(c); } };
template <typename T> CB200_HD T Lift(double c) { return LiftImpl<T>::Get(c); }
CB200_HD double Val(double a) { return a; }
template <int N> CB200_HD double Val(const DualN<N>& a) { return a.v; }
CB200_HD double Sqrt(double a) { return sqrt(a); }
CB200_HD double Sin(double a) { return sin(a); }
CB200_HD double Cos(double a) { return cos(a); }

// hypot(x, y, z) of three scalars / dual numbers (jet.h: value std::hypot, partials x/h dx + y/h dy + z/h dz).
CB200_HD double Hypot3(double x, double y, double z) {
#ifdef __CUDA_ARCH__
  return norm3d(x, y, z);
#else
  return std::hypot(x, y, z);
#endif
}
template <int N> CB200_HD DualN<N> Hypot3(const DualN<N>& x, const DualN<N>& y, const DualN<N>& z) {
  DualN<N> r; r.v = Hypot3(x.v, y.v, z.v);
  const double ax = x.v / r.v, ay = y.v / r.v, az = z.v / r.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = ax * x.d[i] + ay * y.d[i] + az * z.d[i];
  return r;
}

// Rodrigues rotation of pt by the angle-axis vector aa (rotation.h:792-850). As in the
// reference, the first-order branch is taken only when theta = hypot(aa) is EXACTLY zero
// (:799-801, fpclassify(theta) != FP_ZERO): it exists to give finite derivatives there.
template <typename T>
CB200_HD void RotateByAngleAxis(const T* aa, const T* pt, T* out) {
  if (Hypot3(Val(aa[0]), Val(aa[1]), Val(aa[2])) != 0.0) {
    const T theta = Hypot3(aa[0], aa[1], aa[2]);
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
// second half: from the point in the camera frame p = R X + t
template <typename T>
CB200_HD void ProjectFromCameraFrame(const T* p, const T& focal, const T& k1, const T& k2, T* pred) {
  const T xp = -(p[0] / p[2]);
  const T yp = -(p[1] / p[2]);
  const T r2 = xp * xp + yp * yp;
  const T distortion = Lift<T>(1.0) + r2 * (k1 + k2 * r2);
  pred[0] = focal * distortion * xp;
  pred[1] = focal * distortion * yp;
}
template <typename T>
CB200_HD void Project(const T* cam, const T* pt, T* pred) {
  T p[3];
  RotateByAngleAxis(cam, pt, p);
  p[0] = p[0] + cam[3]; p[1] = p[1] + cam[4]; p[2] = p[2] + cam[5];
  ProjectFromCameraFrame(p, cam[6], cam[7], cam[8], pred);
}

// The same residual Jacobian by the chain rule over the camera-frame point:
// stage 1 differentiates p = R(aa) X + t with respect to (aa, X) [6 partials],
// stage 2 differentiates the projection with respect to (p, focal, k1, k2)
// [6 partials]; d pred / d t = d pred / d p. Same operations per partial as the
// 12-partial evaluation at well under half the arithmetic and registers.
// J9 [2 x 9] (camera) and J3 [2 x 3] (point), row-major.
CB200_HD void ProjectWithJacobian(const double* cam, const double* pt, double* pred, double* J9, double* J3) {
  typedef DualN<6> D6;
  D6 aa[3], X[3], p[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) { aa[k] = VarN<6>(cam[k], k); X[k] = VarN<6>(pt[k], 3 + k); }
  RotateByAngleAxis(aa, X, p);
  double pv[3], dp[3][6];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    pv[i] = p[i].v + cam[3 + i];
#pragma unroll
    for (int k = 0; k < 6; ++k) dp[i][k] = p[i].d[k];
  }
  D6 q[3], out[2];
#pragma unroll
  for (int i = 0; i < 3; ++i) q[i] = VarN<6>(pv[i], i);
  ProjectFromCameraFrame(q, VarN<6>(cam[6], 3), VarN<6>(cam[7], 4), VarN<6>(cam[8], 5), out);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    pred[i] = out[i].v;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      // d/d aa_k and d/d X_k through p; d/d t_k = d/d p_k
      J9[9 * i + k] = out[i].d[0] * dp[0][k] + out[i].d[1] * dp[1][k] + out[i].d[2] * dp[2][k];
      J3[3 * i + k] = out[i].d[0] * dp[0][3 + k] + out[i].d[1] * dp[1][3 + k] + out[i].d[2] * dp[2][3 + k];
      J9[9 * i + 3 + k] = out[i].d[k];
      J9[9 * i + 6 + k] = out[i].d[3 + k];
    }
  }
}


}  // namespace cb200_model

#endif  // CB200_SNAVELY_MODEL_H_
