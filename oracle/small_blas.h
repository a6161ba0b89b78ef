// ORACLE — TEST INFRASTRUCTURE ONLY. Never linked, imported or executed by the
// product path (ceres_solver_ceres_solver_b200/); only tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs may use it.
//
// CPU restatement of the small dense kernels of the reference's
// ITERATIVE_SCHUR path. Follows (file:line relative to /root/reference):
//   internal/ceres/small_blas.h:385-466   MatrixVectorMultiply
//   internal/ceres/small_blas.h:471-555   MatrixTransposeVectorMultiply
//   internal/ceres/small_blas.h:265-370   MatrixTransposeMatrixMultiply
//   internal/ceres/small_blas.h:164-263   MatrixMatrixMultiply
//   internal/ceres/invert_psd_matrix.h:48-72                InvertPSDMatrix
//   internal/ceres/implicit_schur_complement.cc:201-202     LLT inverse (upper)
// Eigen (absent from this image) is replaced by its textbook meaning:
// unblocked Cholesky of the upper triangle + triangular solves against I, and
// the adjugate formula for Matrix3d::inverse().
#ifndef ORACLE_SMALL_BLAS_H_
#define ORACLE_SMALL_BLAS_H_

#include <cmath>

namespace oracle {

// c op= A b, A is (ra x ca) row-major. op: +1 add, -1 subtract, 0 assign.
// Each row's dot product is formed first (from 0.0) and then applied to c,
// which is the order of the reference's hand loops.
inline void MatVec(const double* A, int ra, int ca, const double* b, double* c,
                   int op) {
  for (int r = 0; r < ra; ++r) {
    double t = 0.0;
    for (int k = 0; k < ca; ++k) t += A[r * ca + k] * b[k];
    if (op > 0) c[r] += t; else if (op < 0) c[r] -= t; else c[r] = t;
  }
}

// c op= A' b, A is (ra x ca) row-major, b has ra entries, c has ca entries.
inline void MatTVec(const double* A, int ra, int ca, const double* b, double* c,
                    int op) {
  for (int k = 0; k < ca; ++k) {
    double t = 0.0;
    for (int r = 0; r < ra; ++r) t += A[r * ca + k] * b[r];
    if (op > 0) c[k] += t; else if (op < 0) c[k] -= t; else c[k] = t;
  }
}

// C[start_r.., start_c..] op= A' B ; A (ra x ca), B (ra x cb), C has row
// stride col_stride_c. Block written is (ca x cb).
inline void MatTMat(const double* A, int ra, int ca, const double* B, int rb,
                    int cb, double* C, int start_r, int start_c,
                    int /*row_stride_c*/, int col_stride_c, int op) {
  (void)rb;
  for (int i = 0; i < ca; ++i) {
    for (int j = 0; j < cb; ++j) {
      double t = 0.0;
      for (int k = 0; k < ra; ++k) t += A[k * ca + i] * B[k * cb + j];
      double& c = C[(start_r + i) * col_stride_c + start_c + j];
      if (op > 0) c += t; else if (op < 0) c -= t; else c = t;
    }
  }
}

// C op= A B ; A (ra x ca), B (ca x cb).
inline void MatMat(const double* A, int ra, int ca, const double* B, int /*rb*/,
                   int cb, double* C, int start_r, int start_c,
                   int /*row_stride_c*/, int col_stride_c, int op) {
  for (int i = 0; i < ra; ++i) {
    for (int j = 0; j < cb; ++j) {
      double t = 0.0;
      for (int k = 0; k < ca; ++k) t += A[i * ca + k] * B[k * cb + j];
      double& c = C[(start_r + i) * col_stride_c + start_c + j];
      if (op > 0) c += t; else if (op < 0) c -= t; else c = t;
    }
  }
}

// m <- inverse of the SPD matrix whose UPPER triangle is stored in m
// (n x n row-major, n <= 16). Meaning of Eigen's
//   m.selfadjointView<Upper>().llt().solve(Identity)
// i.e. A = U'U (U upper), then solve U'U X = I column by column.
// Returns false if a non-positive pivot is met (Eigen would report
// NumericalIssue; the reference ignores it, so the caller may too).
inline bool LltInverseUpper(double* m, int n) {
  double U[16 * 16];
  bool ok = true;
  for (int j = 0; j < n; ++j) {
    for (int i = 0; i <= j; ++i) {
      double s = m[i * n + j];
      for (int k = 0; k < i; ++k) s -= U[k * n + i] * U[k * n + j];
      if (i == j) {
        if (!(s > 0.0)) ok = false;
        U[j * n + j] = std::sqrt(s);
      } else {
        U[i * n + j] = s / U[i * n + i];
      }
    }
  }
  for (int c = 0; c < n; ++c) {
    double y[16];
    // forward: U' y = e_c
    for (int i = 0; i < n; ++i) {
      double s = (i == c) ? 1.0 : 0.0;
      for (int k = 0; k < i; ++k) s -= U[k * n + i] * y[k];
      y[i] = s / U[i * n + i];
    }
    // backward: U x = y
    for (int i = n - 1; i >= 0; --i) {
      double s = y[i];
      for (int k = i + 1; k < n; ++k) s -= U[i * n + k] * y[k];
      y[i] = s / U[i * n + i];
    }
    for (int i = 0; i < n; ++i) m[i * n + c] = y[i];
  }
  return ok;
}

// out = inverse(m) for a full 3x3 via the adjugate / determinant formula:
// the meaning of Eigen's fixed-size Matrix3d::inverse()
// (invert_psd_matrix.h:60-63, kSize < 5 and assume_full_rank).
inline void Inverse3x3(const double* m, double* out) {
  const double c00 = m[4] * m[8] - m[5] * m[7];
  const double c01 = m[5] * m[6] - m[3] * m[8];
  const double c02 = m[3] * m[7] - m[4] * m[6];
  const double det = m[0] * c00 + m[1] * c01 + m[2] * c02;
  const double inv = 1.0 / det;
  out[0] = c00 * inv;
  out[1] = (m[2] * m[7] - m[1] * m[8]) * inv;
  out[2] = (m[1] * m[5] - m[2] * m[4]) * inv;
  out[3] = c01 * inv;
  out[4] = (m[0] * m[8] - m[2] * m[6]) * inv;
  out[5] = (m[2] * m[3] - m[0] * m[5]) * inv;
  out[6] = c02 * inv;
  out[7] = (m[1] * m[6] - m[0] * m[7]) * inv;
  out[8] = (m[0] * m[4] - m[1] * m[3]) * inv;
}

// General small inverse by Gauss-Jordan with partial pivoting: stands in for
// Eigen's fixed-size inverse() for kSize in {1,2,4} (only used by the KATs,
// whose e-blocks are 1x1 or 2x2).
inline void InverseSmall(const double* m, int n, double* out) {
  if (n == 3) { Inverse3x3(m, out); return; }
  if (n == 1) { out[0] = 1.0 / m[0]; return; }
  if (n == 2) {
    const double det = m[0] * m[3] - m[1] * m[2];
    const double inv = 1.0 / det;
    out[0] = m[3] * inv; out[1] = -m[1] * inv;
    out[2] = -m[2] * inv; out[3] = m[0] * inv;
    return;
  }
  double a[16 * 32];
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      a[i * 2 * n + j] = m[i * n + j];
      a[i * 2 * n + n + j] = (i == j) ? 1.0 : 0.0;
    }
  for (int c = 0; c < n; ++c) {
    int p = c;
    for (int r = c + 1; r < n; ++r)
      if (std::fabs(a[r * 2 * n + c]) > std::fabs(a[p * 2 * n + c])) p = r;
    if (p != c)
      for (int j = 0; j < 2 * n; ++j) {
        double t = a[c * 2 * n + j]; a[c * 2 * n + j] = a[p * 2 * n + j];
        a[p * 2 * n + j] = t;
      }
    const double d = 1.0 / a[c * 2 * n + c];
    for (int j = 0; j < 2 * n; ++j) a[c * 2 * n + j] *= d;
    for (int r = 0; r < n; ++r) {
      if (r == c) continue;
      const double f = a[r * 2 * n + c];
      if (f == 0.0) continue;
      for (int j = 0; j < 2 * n; ++j) a[r * 2 * n + j] -= f * a[c * 2 * n + j];
    }
  }
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) out[i * n + j] = a[i * 2 * n + n + j];
}

}  // namespace oracle

#endif  // ORACLE_SMALL_BLAS_H_
