// Device-side pieces of the Levenberg-Marquardt loop around the linear solve
// (SURVEY 8(f) rows 1 and 2): the Snavely residual / Jacobian evaluator writes
// straight into the solver's device-resident E / F / b arrays, so the Jacobian
// never crosses PCIe, and the full-Jacobian passes of the trust-region loop
// (column norms, Jacobi scaling, gradient, LM diagonal, candidate point) run
// next to it.
//
// Reference:
//   ProgramEvaluator::Evaluate + BlockJacobianWriter   program_evaluator.h:189-260,
//                                                      block_jacobian_writer.cc:102-165
//   SnavelyReprojectionError (autodiff 2,9,3)          examples/snavely_reprojection_error.h:57-92
//   SquaredColumnNorm / ScaleColumns                   block_sparse_matrix.cc:351-450
//   TrustRegionMinimizer::EvaluateGradientAndJacobian  trust_region_minimizer.cc:246-310
//   LevenbergMarquardtStrategy::ComputeStep (diagonal) levenberg_marquardt_strategy.cc:79-95
#ifndef CB200_LM_KERNELS_CUH_
#define CB200_LM_KERNELS_CUH_

#include "cg_kernels.cuh"
#include "snavely_model.h"

namespace cb200 {

constexpr int kEvalThreads = 128;

struct EvalArgs {
  const double* cameras;   // [9C]
  const double* points;    // [3 P_loc]
  const double* obs;       // [2 N_loc]
  const int* cam;          // [N_loc] camera of each row
  const int* row_point;    // [N_loc] local point of each row
  int N;
  const double* scale_e;   // [3 P_loc] Jacobi scaling or null (iteration 0: not known yet)
  const double* scale_f;   // [9C]
  double* b;               // [2 N_loc] residuals
  double* E;               // [N_loc x 6] (JAC) scaled point blocks
  double* F;               // [N_loc x 18] (JAC) scaled camera blocks
  double* norm_e;          // [3 P_loc] (JAC) squared column norms of the UNSCALED Jacobian, accumulated
  double* norm_f;          // [9C]
  double* grad_e;          // [3 P_loc] (JAC) J'r of the unscaled Jacobian, accumulated
  double* grad_f;          // [9C]
  double* cost_partial;    // [gridDim.x] per-CTA sums of r'r
};

// One thread per observation. JAC: residual + 2x3 / 2x9 Jacobian blocks through
// 12-partial dual numbers (what AutoDiffCostFunction<.., 2, 9, 3> propagates);
// the blocks are written already column-scaled, the column norms and the
// gradient of the unscaled Jacobian are accumulated on the fly.
template <bool JAC>
__global__ void __launch_bounds__(kEvalThreads)
bal_eval_kernel(const EvalArgs a) {
  __shared__ double sbuf[8];
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  double rr = 0.0;
  if (r < a.N) {
    const int c = a.cam[r];
    const int p = a.row_point[r];
    double cam[9], pt[3];
#pragma unroll
    for (int k = 0; k < 9; ++k) cam[k] = __ldg(a.cameras + 9 * (size_t)c + k);
#pragma unroll
    for (int k = 0; k < 3; ++k) pt[k] = __ldg(a.points + 3 * (size_t)p + k);
    const double2 ob = reinterpret_cast<const double2*>(a.obs)[r];
    if (!JAC) {
      double pred[2];
      cb200_model::Project<double>(cam, pt, pred);
      const double r0 = pred[0] - ob.x, r1 = pred[1] - ob.y;
      if (a.b != nullptr) reinterpret_cast<double2*>(a.b)[r] = make_double2(r0, r1);
      rr = r0 * r0 + r1 * r1;
    } else {
      using cb200_model::Dual;
      Dual dc[9], dp[3], pred[2];
#pragma unroll
      for (int k = 0; k < 9; ++k) dc[k] = cb200_model::Var(cam[k], k);
#pragma unroll
      for (int k = 0; k < 3; ++k) dp[k] = cb200_model::Var(pt[k], 9 + k);
      cb200_model::Project<Dual>(dc, dp, pred);
      const double r0 = pred[0].v - ob.x, r1 = pred[1].v - ob.y;
      reinterpret_cast<double2*>(a.b)[r] = make_double2(r0, r1);
      rr = r0 * r0 + r1 * r1;
      double* Er = a.E + (size_t)r * 6;
      double* Fr = a.F + (size_t)r * 18;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double d0 = pred[0].d[9 + k], d1 = pred[1].d[9 + k];
        const double s = a.scale_e != nullptr ? a.scale_e[3 * (size_t)p + k] : 1.0;
        Er[k] = d0 * s;
        Er[3 + k] = d1 * s;
        atomicAdd(a.norm_e + 3 * (size_t)p + k, d0 * d0 + d1 * d1);
        atomicAdd(a.grad_e + 3 * (size_t)p + k, d0 * r0 + d1 * r1);
      }
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        const double d0 = pred[0].d[k], d1 = pred[1].d[k];
        const double s = a.scale_f != nullptr ? a.scale_f[9 * (size_t)c + k] : 1.0;
        Fr[k] = d0 * s;
        Fr[9 + k] = d1 * s;
        atomicAdd(a.norm_f + 9 * (size_t)c + k, d0 * d0 + d1 * d1);
        atomicAdd(a.grad_f + 9 * (size_t)c + k, d0 * r0 + d1 * r1);
      }
    }
  }
  const double s = BlockSum(rr, sbuf);
  if (threadIdx.x == 0) a.cost_partial[blockIdx.x] = s;
}

// out[0] = sum of partial[0..n) in index order (deterministic), times `factor`
__global__ void __launch_bounds__(256)
ordered_sum_kernel(const double* __restrict__ partial, int n, double factor, double* __restrict__ out) {
  __shared__ double sbuf[8];
  // fixed assignment of entries to threads, fixed tree: run-to-run identical
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
  s = BlockSum(s, sbuf);
  if (threadIdx.x == 0) out[0] = s * factor;
}

// The Jacobian was written unscaled (scale unknown at iteration 0): scale it in
// place once (ScaleColumns, block_sparse_matrix.cc:425-450).
__global__ void scale_jacobian_kernel(double* __restrict__ E, double* __restrict__ F,
                                      const int* __restrict__ cam, const int* __restrict__ row_point,
                                      long long N, const double* __restrict__ scale_e,
                                      const double* __restrict__ scale_f) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 24 * N) return;
  const long long r = i / 24;
  const int k = (int)(i - 24 * r);
  if (k < 6) E[r * 6 + k] *= scale_e[3 * (size_t)row_point[r] + k % 3];
  else F[r * 18 + (k - 6)] *= scale_f[9 * (size_t)cam[r] + (k - 6) % 9];
}

// jacobian_scaling = 1 / (1 + sqrt(squared column norm))  (trust_region_minimizer.cc:263-274)
__global__ void jacobi_scale_kernel(const double* __restrict__ norm2, double* __restrict__ scale, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) scale[i] = 1.0 / (1.0 + sqrt(norm2[i]));
}

// LM diagonal D = sqrt(clamp(squared column norm of the SCALED Jacobian) / radius)
// (levenberg_marquardt_strategy.cc:79-95); the scaled norm is scale^2 * unscaled norm.
__global__ void lm_diagonal_kernel(const double* __restrict__ norm2, const double* __restrict__ scale,
                                   double min_diagonal, double max_diagonal, double radius,
                                   double* __restrict__ D, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double s = scale != nullptr ? scale[i] : 1.0;
  const double d = fmin(fmax(norm2[i] * s * s, min_diagonal), max_diagonal);
  D[i] = sqrt(d / radius);
}

// candidate = x + delta with delta = (-y) * scale: the step of the scaled problem
// J y = r, negated and un-scaled (levenberg_marquardt_strategy.cc:131-132,
// trust_region_minimizer.cc:448-451). Also the squared norms the parameter-
// tolerance test needs (:720-728): partial[0] += |x|^2, partial[1] += |delta|^2.
__global__ void __launch_bounds__(256)
candidate_kernel(const double* __restrict__ x, const double* __restrict__ y,
                 const double* __restrict__ scale, double* __restrict__ cand, int n,
                 double* __restrict__ norms /* [2], accumulated */) {
  __shared__ double sbuf[8];
  double xx = 0.0, dd = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double xi = x[i];
    const double d = -y[i] * (scale != nullptr ? scale[i] : 1.0);
    cand[i] = xi + d;
    xx += xi * xi;
    dd += d * d;
  }
  xx = BlockSum(xx, sbuf);
  dd = BlockSum(dd, sbuf);
  if (threadIdx.x == 0) { atomicAdd(norms, xx); atomicAdd(norms + 1, dd); }
}

// out[0] = max |g|, out[1] += |g|^2 (gradient_max_norm / gradient_norm, :306-309)
__global__ void __launch_bounds__(256)
gradient_norm_kernel(const double* __restrict__ g, int n, double* __restrict__ out) {
  __shared__ double sbuf[8];
  double m = 0.0, s = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double v = g[i];
    m = fmax(m, fabs(v));
    s += v * v;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, off));
  s = BlockSum(s, sbuf);
  // non-negative doubles order like their bit patterns
  if ((threadIdx.x & 31) == 0)
    atomicMax(reinterpret_cast<unsigned long long*>(out), (unsigned long long)__double_as_longlong(m));
  if (threadIdx.x == 0) atomicAdd(out + 1, s);
}

}  // namespace cb200
#endif  // CB200_LM_KERNELS_CUH_
