// Hand-written FP64 CUDA kernels (sm_100a) for the ITERATIVE_SCHUR path of a
// <2,3,9> bundle-adjustment Jacobian. See DESIGN.md for the data layout and the
// per-kernel byte counts. Reference functions each kernel replaces are cited at
// the kernel.
//
// Device layout of one point shard (all indices local to the shard):
//   E   [N x 6]   2x3 row-major point blocks, rows grouped by point
//   F   [N x 18]  2x9 row-major camera blocks, same row order
//   cam [N]       camera index of each row
//   row_start [P+1]  first row of each point
//   ete_inv [P x 9]  (E'E + D_e^2)^-1, row-major 3x3
// Camera-sized vectors (9C) are replicated on every GPU.
#ifndef CB200_KERNELS_CUH_
#define CB200_KERNELS_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "flatten.h"

namespace cb200 {

constexpr int kTileThreads = 256;
static_assert(kTileThreads == kTileObs, "one thread per observation of a tile");

// ----------------------------------------------------------------------------
// mbarrier + 1-D bulk async copy (TMA engine, cp.async.bulk) helpers
// ----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        " selp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(phase)
        : "memory");
  } while (!ok);
}

// ----------------------------------------------------------------------------
// small dense helpers
// ----------------------------------------------------------------------------
// In-place inverse of an SPD n x n row-major block via Cholesky of the UPPER
// triangle and two triangular solves per identity column: the meaning of Eigen's
// block.selfadjointView<Upper>().llt().solve(Identity)
// (implicit_schur_complement.cc:201-202, block_random_access_diagonal_matrix.cc:96-98).
template <int n>
__device__ __forceinline__ void LltInverseUpper(double* m) {
  double U[n * n];
#pragma unroll
  for (int j = 0; j < n; ++j) {
#pragma unroll
    for (int i = 0; i <= j; ++i) {
      double s = m[i * n + j];
#pragma unroll
      for (int k = 0; k < i; ++k) s -= U[k * n + i] * U[k * n + j];
      if (i == j) U[j * n + j] = sqrt(s);
      else U[i * n + j] = s / U[i * n + i];
    }
  }
#pragma unroll
  for (int c = 0; c < n; ++c) {
    double y[n];
#pragma unroll
    for (int i = 0; i < n; ++i) {
      double s = (i == c) ? 1.0 : 0.0;
#pragma unroll
      for (int k = 0; k < i; ++k) s -= U[k * n + i] * y[k];
      y[i] = s / U[i * n + i];
    }
#pragma unroll
    for (int i = n - 1; i >= 0; --i) {
      double s = y[i];
#pragma unroll
      for (int k = i + 1; k < n; ++k) s -= U[i * n + k] * y[k];
      y[i] = s / U[i * n + i];
    }
#pragma unroll
    for (int i = 0; i < n; ++i) m[i * n + c] = y[i];
  }
}

// Adjugate / determinant inverse of a 3x3: the meaning of Eigen's fixed-size
// Matrix3d::inverse() used by InvertPSDMatrix<3> (invert_psd_matrix.h:60-63).
__device__ __forceinline__ void Inverse3x3(const double* m, double* out) {
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

// ----------------------------------------------------------------------------
// Fused per-point kernels: one CTA per tile of whole points (<= kTileObs rows).
// ----------------------------------------------------------------------------
enum PointMode {
  kModeSx = 0,    // y += F'(t - E v),  t = F x          S*x without the D_f^2 term
  kModePseZ = 1,  // y += F'(E v),      t = F x          F'E (E'E)^-1 E'F x
  kModeRhs = 2,   // y += F'(t - E v),  t = b            reduced right-hand side
  kModeBack = 3   // y_e = v,           t = b - F x      back substitution
};                // with u = sum_rows E' t and v = (E'E + D_e^2)^-1 u per point

struct PointArgs {
  const double* E;
  const double* F;
  const int* cam;
  const int* row_start;
  const int* tile_point;
  const double* ete_inv;
  const double* x;   // 9C camera vector (unused for kModeRhs)
  const double* b;   // 2N rows (kModeRhs, kModeBack)
  double* yrep;      // nrep privatised camera accumulators, rep_stride apart
  int nrep;
  int rep_stride;
  double* y_e;       // 3P (kModeBack)
  const int* done;   // device convergence flag; the kernel is a no-op when set
};

constexpr int kPointSmemBytes =
    kTileObs * 6 * 8 + kTileObs * 18 * 8 + kTileObs * 3 * 8 + 8 * 288 * 8 + kTileObs * 4 + 16;

// Replaces the five SpMV passes of ImplicitSchurComplement::RightMultiplyAndAccumulate
// (implicit_schur_complement.cc:106-144: RightMultiplyAndAccumulateF,
// LeftMultiplyAndAccumulateE, block-diagonal multiply, RightMultiplyAndAccumulateE,
// LeftMultiplyAndAccumulateF; partitioned_matrix_view_impl.h:112-375) by one pass
// that streams each Jacobian block once. The same kernel, with a different
// definition of t and of the scattered quantity, is UpdateRhs (:251-276),
// InversePowerSeriesOperatorRightMultiplyAccumulate (:146-174) and
// BackSubstitute (:208-243).
template <int MODE>
__global__ void __launch_bounds__(kTileThreads)
point_tile_kernel(const PointArgs a) {
  if (a.done != nullptr && *a.done != 0) return;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  double* sE = reinterpret_cast<double*>(smem_raw);
  double* sF = sE + kTileObs * 6;
  double* sU = sF + kTileObs * 18;
  double* sX = sU + kTileObs * 3;
  int* sCam = reinterpret_cast<int*>(sX + 8 * 288);
  uint64_t* bar = reinterpret_cast<uint64_t*>(sCam + kTileObs);

  const int tid = threadIdx.x;
  const int lane = tid & 31;
  const int warp = tid >> 5;
  const int p0 = a.tile_point[blockIdx.x];
  const int p1 = a.tile_point[blockIdx.x + 1];
  const int r0 = a.row_start[p0];
  const int n = a.row_start[p1] - r0;
  const int np = p1 - p0;
  if (n > kTileObs) return;  // a long point: handled by point_long_kernel
  if (n == 0) {
    if (MODE == kModeBack && tid < np) {
      a.y_e[3 * (p0 + tid)] = 0.0; a.y_e[3 * (p0 + tid) + 1] = 0.0; a.y_e[3 * (p0 + tid) + 2] = 0.0;
    }
    return;
  }
  if (tid == 0) mbar_init(bar, 1);
  __syncthreads();
  if (tid == 0) {
    mbar_expect_tx(bar, (uint32_t)n * 192u);
    bulk_g2s(sE, a.E + (size_t)r0 * 6, (uint32_t)n * 48u, bar);
    bulk_g2s(sF, a.F + (size_t)r0 * 18, (uint32_t)n * 144u, bar);
  }
  // while the tile is in flight: camera ids, per-point metadata, x gather
  if (tid < n) sCam[tid] = a.cam[r0 + tid];
  int ps = 0, pe = 0;
  double inv[9];
  if (tid < np) {
    ps = a.row_start[p0 + tid] - r0;
    pe = a.row_start[p0 + tid + 1] - r0;
    const double* g = a.ete_inv + (size_t)(p0 + tid) * 9;
#pragma unroll
    for (int k = 0; k < 9; ++k) inv[k] = g[k];
  }
  double t0 = 0.0, t1 = 0.0;
  if (MODE == kModeRhs || MODE == kModeBack) {
    if (tid < n) {
      const double2 bb = reinterpret_cast<const double2*>(a.b)[r0 + tid];
      t0 = bb.x; t1 = bb.y;
    }
  }
  double* sXw = sX + warp * 288;
  const int wbase = warp * 32;
  const int nw = min(32, n - wbase);
  __syncwarp();
  if (MODE != kModeRhs) {
    // cooperative gather: consecutive lanes read consecutive doubles of the
    // 9-vectors of the warp's 32 cameras (few sectors per request)
#pragma unroll
    for (int j = 0; j < 9; ++j) {
      const int idx = j * 32 + lane;
      const int o = idx / 9;
      const int k = idx - 9 * o;
      if (o < nw) sXw[idx] = __ldg(a.x + 9 * sCam[wbase + o] + k);
    }
    __syncwarp();
  }
  mbar_wait(bar, 0);

  double Ei[6], Fi[18];
  if (tid < n) {
    const double2* e2 = reinterpret_cast<const double2*>(sE + tid * 6);
    const double2* f2 = reinterpret_cast<const double2*>(sF + tid * 18);
#pragma unroll
    for (int k = 0; k < 3; ++k) { const double2 v = e2[k]; Ei[2 * k] = v.x; Ei[2 * k + 1] = v.y; }
#pragma unroll
    for (int k = 0; k < 9; ++k) { const double2 v = f2[k]; Fi[2 * k] = v.x; Fi[2 * k + 1] = v.y; }
    if (MODE != kModeRhs) {
      double f0 = 0.0, f1 = 0.0;
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const double xk = sXw[lane * 9 + k];
        f0 += Fi[k] * xk;
        f1 += Fi[9 + k] * xk;
      }
      if (MODE == kModeBack) { t0 -= f0; t1 -= f1; } else { t0 = f0; t1 = f1; }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) sU[tid * 3 + k] = Ei[k] * t0 + Ei[3 + k] * t1;
  }
  __syncthreads();
  if (tid < np) {
    double u0 = 0.0, u1 = 0.0, u2 = 0.0;
    for (int i = ps; i < pe; ++i) { u0 += sU[i * 3]; u1 += sU[i * 3 + 1]; u2 += sU[i * 3 + 2]; }
    const double v0 = inv[0] * u0 + inv[1] * u1 + inv[2] * u2;
    const double v1 = inv[3] * u0 + inv[4] * u1 + inv[5] * u2;
    const double v2 = inv[6] * u0 + inv[7] * u1 + inv[8] * u2;
    if (MODE == kModeBack) {
      double* y = a.y_e + (size_t)(p0 + tid) * 3;
      y[0] = v0; y[1] = v1; y[2] = v2;
    } else {
      for (int i = ps; i < pe; ++i) { sU[i * 3] = v0; sU[i * 3 + 1] = v1; sU[i * 3 + 2] = v2; }
    }
  }
  if (MODE == kModeBack) return;
  __syncthreads();
  if (tid < n) {
    const double v0 = sU[tid * 3], v1 = sU[tid * 3 + 1], v2 = sU[tid * 3 + 2];
    const double ev0 = Ei[0] * v0 + Ei[1] * v1 + Ei[2] * v2;
    const double ev1 = Ei[3] * v0 + Ei[4] * v1 + Ei[5] * v2;
    const double w0 = (MODE == kModePseZ) ? ev0 : t0 - ev0;
    const double w1 = (MODE == kModePseZ) ? ev1 : t1 - ev1;
#pragma unroll
    for (int k = 0; k < 9; ++k) sXw[lane * 9 + k] = Fi[k] * w0 + Fi[9 + k] * w1;
  }
  __syncwarp();
  double* y = a.yrep + (size_t)(blockIdx.x % a.nrep) * a.rep_stride;
#pragma unroll
  for (int j = 0; j < 9; ++j) {
    const int idx = j * 32 + lane;
    const int o = idx / 9;
    const int k = idx - 9 * o;
    if (o < nw) atomicAdd(y + 9 * sCam[wbase + o] + k, sXw[idx]);
  }
}

// block-wide sum of one double per thread (kTileThreads threads); result valid
// in every thread. Fixed tree => deterministic.
__device__ __forceinline__ double BlockSum(double v, double* sbuf /* >= 8 */) {
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) sbuf[warp] = v;
  __syncthreads();
  double s = 0.0;
  const int nw = (blockDim.x + 31) >> 5;
  for (int w = 0; w < nw; ++w) s += sbuf[w];
  return s;
}

// A point with more rows than a tile: one CTA walks its rows twice (second pass
// hits L2). grid = number of long points.
template <int MODE>
__global__ void __launch_bounds__(kTileThreads)
point_long_kernel(const PointArgs a, const int* long_points) {
  if (a.done != nullptr && *a.done != 0) return;
  __shared__ double sbuf[8];
  const int p = long_points[blockIdx.x];
  const int r0 = a.row_start[p], r1 = a.row_start[p + 1];
  double u0 = 0.0, u1 = 0.0, u2 = 0.0;
  for (int r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
    const double* Ei = a.E + (size_t)r * 6;
    const double* Fi = a.F + (size_t)r * 18;
    double t0 = 0.0, t1 = 0.0;
    if (MODE == kModeRhs || MODE == kModeBack) { t0 = a.b[2 * (size_t)r]; t1 = a.b[2 * (size_t)r + 1]; }
    if (MODE != kModeRhs) {
      const double* x = a.x + 9 * a.cam[r];
      double f0 = 0.0, f1 = 0.0;
      for (int k = 0; k < 9; ++k) { f0 += Fi[k] * x[k]; f1 += Fi[9 + k] * x[k]; }
      if (MODE == kModeBack) { t0 -= f0; t1 -= f1; } else { t0 = f0; t1 = f1; }
    }
    u0 += Ei[0] * t0 + Ei[3] * t1;
    u1 += Ei[1] * t0 + Ei[4] * t1;
    u2 += Ei[2] * t0 + Ei[5] * t1;
  }
  u0 = BlockSum(u0, sbuf); u1 = BlockSum(u1, sbuf); u2 = BlockSum(u2, sbuf);
  const double* inv = a.ete_inv + (size_t)p * 9;
  const double v0 = inv[0] * u0 + inv[1] * u1 + inv[2] * u2;
  const double v1 = inv[3] * u0 + inv[4] * u1 + inv[5] * u2;
  const double v2 = inv[6] * u0 + inv[7] * u1 + inv[8] * u2;
  if (MODE == kModeBack) {
    if (threadIdx.x == 0) { a.y_e[3 * (size_t)p] = v0; a.y_e[3 * (size_t)p + 1] = v1; a.y_e[3 * (size_t)p + 2] = v2; }
    return;
  }
  double* y = a.yrep + (size_t)(blockIdx.x % a.nrep) * a.rep_stride;
  for (int r = r0 + threadIdx.x; r < r1; r += blockDim.x) {
    const double* Ei = a.E + (size_t)r * 6;
    const double* Fi = a.F + (size_t)r * 18;
    double t0 = 0.0, t1 = 0.0;
    if (MODE == kModeRhs) { t0 = a.b[2 * (size_t)r]; t1 = a.b[2 * (size_t)r + 1]; }
    else {
      const double* x = a.x + 9 * a.cam[r];
      for (int k = 0; k < 9; ++k) { t0 += Fi[k] * x[k]; t1 += Fi[9 + k] * x[k]; }
    }
    const double ev0 = Ei[0] * v0 + Ei[1] * v1 + Ei[2] * v2;
    const double ev1 = Ei[3] * v0 + Ei[4] * v1 + Ei[5] * v2;
    const double w0 = (MODE == kModePseZ) ? ev0 : t0 - ev0;
    const double w1 = (MODE == kModePseZ) ? ev1 : t1 - ev1;
    double* yc = y + 9 * a.cam[r];
    for (int k = 0; k < 9; ++k) atomicAdd(yc + k, Fi[k] * w0 + Fi[9 + k] * w1);
  }
}

// ----------------------------------------------------------------------------
// Per-point setup: E'E + D_e^2, its inverses, and the 2x2 row weights the
// Schur-Jacobi build needs. One thread per point.
//   UpdateBlockDiagonalEtE             partitioned_matrix_view_impl.h:446-523
//   AddDiagonalAndInvert               implicit_schur_complement.cc:179-204
//   ChunkDiagonalBlockAndGradient+InvertPSDMatrix<3>
//                                      schur_eliminator_impl.h:441-505, invert_psd_matrix.h:60-63
// weights (optional): W_r = I_2 - E_r inverse(ete) E_r' (3 doubles: w00,w01,w11)
// so that the camera block of the Schur complement is D_f^2 + sum_r F_r' W_r F_r.
// ----------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
point_ete_kernel(const double* __restrict__ E, const int* __restrict__ row_start,
                 const double* __restrict__ De /* 3P or null */, int P,
                 double* __restrict__ ete_inv, double* __restrict__ ete_raw /* or null */,
                 double* __restrict__ weights /* or null */) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int r0 = row_start[p], r1 = row_start[p + 1];
  double m[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) m[k] = 0.0;
  for (int r = r0; r < r1; ++r) {
    const double2* e2 = reinterpret_cast<const double2*>(E + (size_t)r * 6);
    const double2 a0 = e2[0], a1 = e2[1], a2 = e2[2];
    const double e[6] = {a0.x, a0.y, a1.x, a1.y, a2.x, a2.y};
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) m[i * 3 + j] += e[i] * e[j] + e[3 + i] * e[3 + j];
  }
  if (ete_raw != nullptr) {
#pragma unroll
    for (int k = 0; k < 9; ++k) ete_raw[(size_t)p * 9 + k] = m[k];
  }
  if (De != nullptr) {
#pragma unroll
    for (int i = 0; i < 3; ++i) { const double d = De[3 * (size_t)p + i]; m[i * 4] += d * d; }
  }
  if (weights != nullptr) {
    double c[9];
    Inverse3x3(m, c);
    for (int r = r0; r < r1; ++r) {
      const double* e = E + (size_t)r * 6;
      double g0[3], g1[3];  // rows of E_r * c
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        g0[j] = e[0] * c[j] + e[1] * c[3 + j] + e[2] * c[6 + j];
        g1[j] = e[3] * c[j] + e[4] * c[3 + j] + e[5] * c[6 + j];
      }
      double* w = weights + (size_t)r * 3;
      w[0] = 1.0 - (g0[0] * e[0] + g0[1] * e[1] + g0[2] * e[2]);
      w[1] = -(g0[0] * e[3] + g0[1] * e[4] + g0[2] * e[5]);
      w[2] = 1.0 - (g1[0] * e[3] + g1[1] * e[4] + g1[2] * e[5]);
    }
  }
  LltInverseUpper<3>(m);
#pragma unroll
  for (int k = 0; k < 9; ++k) ete_inv[(size_t)p * 9 + k] = m[k];
}

}  // namespace cb200
#endif  // CB200_KERNELS_CUH_
