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
#include "stream_kernels.cuh"

namespace cb200 {

constexpr int kEvalThreads = 128;
constexpr int kEvalReplicas = 32;

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
  double* grad_e;          // [3 P_loc] (JAC) J'r of the unscaled Jacobian, accumulated
  double* cam_acc;         // [kEvalReplicas][C x 18] (JAC) per camera: 9 squared column norms, then 9
                           // gradient entries; privatised copies spread the ~2800 same-address
                           // atomics per camera entry, fold_camera_accumulators_kernel sums them
  int C;
  double* cost_partial;    // [gridDim.x] per-CTA sums of r'r
};

// One thread per observation. JAC: residual + 2x3 / 2x9 Jacobian blocks through
// forward-mode dual numbers (the role of ceres::Jet in AutoDiffCostFunction<.., 2, 9, 3>),
// chained over the camera-frame point (snavely_model.h: ProjectWithJacobian);
// the blocks are written already column-scaled, the column norms and the
// gradient of the unscaled Jacobian are accumulated on the fly.
template <bool JAC>
__global__ void __launch_bounds__(kEvalThreads, JAC ? 4 : 8)
bal_eval_kernel(const EvalArgs a) {
  __shared__ double sbuf[8];
  __shared__ __align__(16) double sT[JAC ? kEvalThreads / 32 : 1][JAC ? 32 * 19 : 2];  // per warp: 32 rows x 18 (+1 pad)
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool on = r < a.N;
  double rr = 0.0;
  // rows past the end re-evaluate the last row (keeps every lane in the shuffles) and add nothing
  const int rc = on ? r : a.N - 1;
  const int c = a.cam[rc];
  const int p = a.row_point[rc];
  double cam[9], pt[3];
#pragma unroll
  for (int k = 0; k < 9; ++k) cam[k] = __ldg(a.cameras + 9 * (size_t)c + k);
#pragma unroll
  for (int k = 0; k < 3; ++k) pt[k] = __ldg(a.points + 3 * (size_t)p + k);
  const double2 ob = reinterpret_cast<const double2*>(a.obs)[rc];
  if (!JAC) {
    double pred[2];
    cb200_model::Project<double>(cam, pt, pred);
    const double r0 = pred[0] - ob.x, r1 = pred[1] - ob.y;
    if (on) {
      if (a.b != nullptr) reinterpret_cast<double2*>(a.b)[r] = make_double2(r0, r1);
      rr = r0 * r0 + r1 * r1;
    }
  } else {
    double pred[2], J9[18], J3[6];
    cb200_model::ProjectWithJacobian(cam, pt, pred, J9, J3);
    const double r0 = pred[0] - ob.x, r1 = pred[1] - ob.y;
    const double live = on ? 1.0 : 0.0;
    if (on) {
      reinterpret_cast<double2*>(a.b)[r] = make_double2(r0, r1);
      rr = r0 * r0 + r1 * r1;
    }
    // ---- the warp's 32 consecutive rows of F (then E) are one contiguous range of global
    // memory: stage them in shared memory and store 16 bytes per lane, fully coalesced
    // (per-thread row stores would touch 32 sectors per instruction, 8 bytes each) ----
    double* t = sT[warp];
    const long long row0 = (long long)blockIdx.x * blockDim.x + warp * 32;  // first row of the warp
    const int rows = (int)min(32LL, (long long)a.N - row0);                // (<= 0: nothing to store)
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const double s = a.scale_f != nullptr ? a.scale_f[9 * (size_t)c + k] : 1.0;
      t[lane * 18 + k] = J9[k] * s;
      t[lane * 18 + 9 + k] = J9[9 + k] * s;
    }
    __syncwarp();
    {
      double2* dst = reinterpret_cast<double2*>(a.F + row0 * 18);
      const double2* src = reinterpret_cast<const double2*>(t);
#pragma unroll
      for (int j = 0; j < 9; ++j) {
        const int idx = j * 32 + lane;
        if (idx < rows * 9) dst[idx] = src[idx];
      }
    }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double s = a.scale_e != nullptr ? a.scale_e[3 * (size_t)p + k] : 1.0;
      t[lane * 6 + k] = J3[k] * s;
      t[lane * 6 + 3 + k] = J3[3 + k] * s;
    }
    __syncwarp();
    {
      double2* dst = reinterpret_cast<double2*>(a.E + row0 * 6);
      const double2* src = reinterpret_cast<const double2*>(t);
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const int idx = j * 32 + lane;
        if (idx < rows * 3) dst[idx] = src[idx];
      }
    }
    __syncwarp();
    // ---- point part: rows of a point are consecutive, so sum them along the lanes
    // (segments keyed on the point id) and let the last lane of a segment add ----
    const int p_up = __shfl_up_sync(0xffffffffu, p, 1);
    const unsigned heads = __ballot_sync(0xffffffffu, lane == 0 || p != p_up);
    const int head = 31 - __clz(heads & (0xffffffffu >> (31 - lane)));
    const bool is_tail = lane == 31 || ((heads >> (lane + 1)) & 1u);
    const SegMask km = MakeSegMask(lane, head);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double d0 = J3[k], d1 = J3[3 + k];
      const double n2 = SegScan(live * (d0 * d0 + d1 * d1), km);
      const double g = SegScan(live * (d0 * r0 + d1 * r1), km);
      if (is_tail) {
        atomicAdd(a.norm_e + 3 * (size_t)p + k, n2);
        atomicAdd(a.grad_e + 3 * (size_t)p + k, g);
      }
    }
    // ---- camera part: transpose the warp's 32 x 18 values through shared memory so that
    // consecutive lanes add to consecutive entries of a camera (few sectors per RED) ----
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const double d0 = J9[k], d1 = J9[9 + k];
      t[lane * 19 + k] = live * (d0 * d0 + d1 * d1);
      t[lane * 19 + 9 + k] = live * (d0 * r0 + d1 * r1);
    }
    __syncwarp();
    double* acc = a.cam_acc + (size_t)(blockIdx.x % kEvalReplicas) * a.C * 18;
#pragma unroll
    for (int j = 0; j < 18; ++j) {
      const int idx = j * 32 + lane;
      const int o = idx / 18;
      const int k = idx - 18 * o;
      const int co = __shfl_sync(0xffffffffu, c, o);
      atomicAdd(acc + (size_t)co * 18 + k, t[o * 19 + k]);
    }
  }
  const double s = BlockSum(rr, sbuf);
  if (threadIdx.x == 0) a.cost_partial[blockIdx.x] = s;
}

// camera column norms / gradient: sum of the privatised copies (which are re-zeroed)
__global__ void fold_camera_accumulators_kernel(double* __restrict__ cam_acc, int C,
                                                double* __restrict__ norm_f, double* __restrict__ grad_f) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (camera, entry of 18)
  if (i >= 18 * C) return;
  double s = 0.0;
  for (int rep = 0; rep < kEvalReplicas; ++rep) s += __ldcg(cam_acc + (size_t)rep * 18 * C + i);
  for (int rep = 0; rep < kEvalReplicas; ++rep) cam_acc[(size_t)rep * 18 * C + i] = 0.0;
  const int c = i / 18, k = i - 18 * c;
  if (k < 9) norm_f[9 * c + k] = s; else grad_f[9 * c + k - 9] = s;
}

// out[0] = sum of partial[0..n) in index order (deterministic), times `factor`
__global__ void __launch_bounds__(1024)
ordered_sum_kernel(const double* __restrict__ partial, int n, double factor, double* __restrict__ out) {
  __shared__ double sw[32];
  // fixed assignment of entries to threads, fixed tree: run-to-run identical
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
  if ((threadIdx.x & 31) == 0) sw[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = sw[threadIdx.x];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if (threadIdx.x == 0) out[0] = s * factor;
  }
}

// The Jacobian was written unscaled (scale unknown at iteration 0): scale it in
// place once (ScaleColumns, block_sparse_matrix.cc:425-450).
__global__ void __launch_bounds__(128)
scale_jacobian_kernel(double* __restrict__ E, double* __restrict__ F,
                      const int* __restrict__ cam, const int* __restrict__ row_point,
                      long long N, const double* __restrict__ scale_e,
                      const double* __restrict__ scale_f) {
  const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // one thread per row
  if (r >= N) return;
  const double* se = scale_e + 3 * (size_t)row_point[r];
  const double* sf = scale_f + 9 * (size_t)cam[r];
  double2* e2 = reinterpret_cast<double2*>(E + r * 6);
  double2* f2 = reinterpret_cast<double2*>(F + r * 18);
  double e[6], f[18], s3[3], s9[9];
#pragma unroll
  for (int k = 0; k < 3; ++k) { const double2 v = e2[k]; e[2 * k] = v.x; e[2 * k + 1] = v.y; s3[k] = __ldg(se + k); }
#pragma unroll
  for (int k = 0; k < 9; ++k) { const double2 v = f2[k]; f[2 * k] = v.x; f[2 * k + 1] = v.y; s9[k] = __ldg(sf + k); }
#pragma unroll
  for (int k = 0; k < 6; ++k) e[k] *= s3[k % 3];
#pragma unroll
  for (int k = 0; k < 18; ++k) f[k] *= s9[k % 9];
#pragma unroll
  for (int k = 0; k < 3; ++k) e2[k] = make_double2(e[2 * k], e[2 * k + 1]);
#pragma unroll
  for (int k = 0; k < 9; ++k) f2[k] = make_double2(f[2 * k], f[2 * k + 1]);
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
