// Device-resident preconditioned conjugate gradients on the reduced camera
// system: the vector updates, dot products and every scalar convergence test of
// ConjugateGradientsSolver (conjugate_gradients_solver.h:107-305) run on the GPU;
// the host only launches iterations and polls CgState::done.
//
// Dot products are reduced deterministically: each CTA writes its partial, the
// last CTA to finish (ticket counter) sums the partials in index order and runs
// the scalar logic.
#ifndef CB200_CG_KERNELS_CUH_
#define CB200_CG_KERNELS_CUH_

#include <math.h>

#include "kernels.cuh"

namespace cb200 {

constexpr int kVecThreads = 256;
constexpr int kMaxVecBlocks = 1024;    // CTAs of the stand-alone vector kernels (their last CTA sums the partials alone)
constexpr int kMaxStepBlocks = 4096;   // CTAs of the fused step kernel (14 cameras each: up to 57 344 cameras)
// PCG iterations whose per-point intermediate v_k = (E'E + D_e^2)^-1 E'F p_k is kept, so that a solve
// that ends within this many iterations gets its back substitution as y_e = v_rhs - sum alpha_k v_k
// (a 3P-sized combination) instead of one more pass over the Jacobian (stream_kernels.cuh).
constexpr int kHistCap = 8;

// message selectors (host formats the reference's strings from these)
enum CgReason {
  kReasonMaxIterations = 0,
  kReasonZeroRhs = 1,
  kReasonInitialResidual = 2,
  kReasonRhoFailure = 3,
  kReasonBetaFailure = 4,
  kReasonIndefinite = 5,
  kReasonAlphaFailure = 6,
  kReasonZetaConvergence = 7,
  kReasonResidualConvergence = 8
};

struct CgState {
  double rho, last_rho, beta, pq, alpha, Q0, Q1, zeta, norm_r, norm_rhs, tol_r;
  double norm_p, norm_q;       // only for the "indefinite" message
  double q_tolerance, r_tolerance;
  int iter;                    // current iteration (summary.num_iterations)
  int done;                    // 0 = running
  int termination;             // CB200_LINEAR_SOLVER_*
  int reason;                  // CgReason
  int min_iterations, max_iterations;
  int peer_timeout;            // a peer's partial sum never arrived (cg_step.cuh); fatal
  int hist_valid;              // 1: every x update so far has its alpha in alpha_hist (and its v_k stored)
  double alpha_hist[kHistCap]; // step lengths of iterations 1 .. kHistCap (0 where x was not updated)
  unsigned long long stamp[4]; // %globaltimer at: solve start, PCG start, PCG end (= done), spare
  // ---- everything above is the "head": reset from the host at the start of a solve, read back after ----
  unsigned int ticket[12];           // [0..3] FinishDots; [4 + 2c], [5 + 2c]: peer_allreduce_kernel on channel c
  unsigned int bar_count, bar_flag;  // grid barrier of cg_step_kernel: arrivals (0 between launches), barriers completed
  unsigned int peer_epoch[2];        // exchanges completed on peer channel 0 (main stream) / 1 (side stream)
  double partial[6][kMaxStepBlocks];
};

__device__ __forceinline__ unsigned long long GlobalTimerNs() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ bool IsZeroOrInfinity(double x) { return x == 0.0 || isinf(x); }

// Block partials -> ordered final sums in the last CTA. Returns true in every
// thread of the last CTA; sums[] then holds the K totals (shared memory).
template <int K>
__device__ __forceinline__ bool FinishDots(CgState* st, int slot, const double* vals, double* sums,
                                           double* sbuf) {
  __shared__ bool is_last;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    const double s = BlockSum(vals[k], sbuf);
    if (threadIdx.x == 0) st->partial[k][blockIdx.x] = s;
  }
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int t = atomicAdd(&st->ticket[slot], 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return false;
  __threadfence();
  if (threadIdx.x < K) {
    double s = 0.0;
    const volatile double* part = st->partial[threadIdx.x];
    for (unsigned int b = 0; b < gridDim.x; ++b) s += part[b];
    sums[threadIdx.x] = s;
  }
  if (threadIdx.x == 0) st->ticket[slot] = 0;
  __syncthreads();
  return true;
}

__device__ __forceinline__ void Finish(CgState* st, int termination, int reason) {
  st->termination = termination;
  st->reason = reason;
  st->stamp[2] = GlobalTimerNs();
  st->done = 1;
}

// Start of ConjugateGradientsSolver (:118-157): |rhs|, r = rhs - S x0 (r is given:
// the host computed it, or copied rhs when x0 = 0), |r|, Q0 = -x'(rhs + r).
__global__ void __launch_bounds__(kVecThreads)
cg_begin_kernel(CgState* st, const double* __restrict__ rhs, const double* __restrict__ r,
                double* __restrict__ x, int n) {
  __shared__ double sbuf[8];
  __shared__ double sums[3];
  double v[3] = {0.0, 0.0, 0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double b = rhs[i], ri = r[i], xi = x[i];
    v[0] += b * b; v[1] += ri * ri; v[2] += xi * (b + ri);
  }
  if (!FinishDots<3>(st, 0, v, sums, sbuf)) return;
  if (threadIdx.x == 0) {
    st->iter = 0;
    st->done = 0;
    st->termination = CB200_LINEAR_SOLVER_NO_CONVERGENCE;
    st->reason = kReasonMaxIterations;
    st->norm_rhs = sqrt(sums[0]);
    st->tol_r = st->r_tolerance * st->norm_rhs;
    st->norm_r = sqrt(sums[1]);
    st->rho = 1.0;
    st->Q0 = -sums[2];
    st->stamp[1] = GlobalTimerNs();
    if (st->norm_rhs == 0.0) {
      Finish(st, CB200_LINEAR_SOLVER_SUCCESS, kReasonZeroRhs);
    } else if (st->min_iterations == 0 && st->norm_r <= st->tol_r) {
      Finish(st, CB200_LINEAR_SOLVER_SUCCESS, kReasonInitialResidual);
    }
  }
}
// |rhs| = 0: the reference zeroes the solution (:126-131).
__global__ void cg_zero_solution_if_zero_rhs_kernel(const CgState* st, double* x, int n) {
  if (!(st->done && st->reason == kReasonZeroRhs)) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x[i] = 0.0;
}

// z = M^-1 r and rho = r'z (:160-171), then beta (:176-187).
// PRECOND 0: z = r (IDENTITY); 1: 9x9 block diagonal; 2: z was computed by
// other kernels (power series), only the dot product is taken.
template <int PRECOND>
__global__ void __launch_bounds__(kVecThreads)
cg_rho_kernel(CgState* st, const double* __restrict__ blocks, const double* __restrict__ r,
              double* __restrict__ z, int n) {
  if (st->done) return;
  __shared__ double sbuf[8];
  __shared__ double sums[1];
  double v[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double zi;
    if (PRECOND == 0) { zi = r[i]; z[i] = zi; }
    else if (PRECOND == 1) {
      const double* row = blocks + (size_t)i * 9;
      const double* rc = r + 9 * (i / 9);
      zi = 0.0;
#pragma unroll
      for (int k = 0; k < 9; ++k) zi += row[k] * rc[k];
      z[i] = zi;
    } else { zi = z[i]; }
    v[0] += r[i] * zi;
  }
  if (!FinishDots<1>(st, 1, v, sums, sbuf)) return;
  if (threadIdx.x == 0) {
    st->iter += 1;
    st->last_rho = st->rho;
    st->rho = sums[0];
    if (IsZeroOrInfinity(st->rho)) { Finish(st, CB200_LINEAR_SOLVER_FAILURE, kReasonRhoFailure); return; }
    if (st->iter > 1) {
      st->beta = st->rho / st->last_rho;
      if (IsZeroOrInfinity(st->beta)) Finish(st, CB200_LINEAR_SOLVER_FAILURE, kReasonBetaFailure);
    }
  }
}

// p = z (first iteration) or p = z + beta p (:173-188).
__global__ void cg_update_p_kernel(const CgState* st, const double* __restrict__ z,
                                   double* __restrict__ p, double* __restrict__ p10, int n) {
  if (st->done) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double pi = (st->iter == 1) ? z[i] : z[i] + st->beta * p[i];
  p[i] = pi;
  p10[i + i / 9] = pi;  // the padded copy S*p gathers from
}

// q = sum of the privatised accumulators (which are re-zeroed for the next use).
// Deliberately not fused with the D_f^2 term: with several GPUs the all-reduce of
// q sits between the two.
__global__ void reduce_replicas_kernel(const int* done, double* __restrict__ yrep, int nrep,
                                       int rep_stride, double* __restrict__ q, int n) {
  if (done != nullptr && *done) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = i / 9;
  double s = 0.0;
  for (int rep = 0; rep < nrep; ++rep) s += __ldcg(yrep + (size_t)rep * rep_stride + i + c);  // 10 per camera
  for (int rep = 0; rep < nrep; ++rep) yrep[(size_t)rep * rep_stride + i + c] = 0.0;
  q[i] = s;
}

// x [9C] -> x10 [10C]: camera vectors are gathered with 16-byte loads, which need
// 16-byte aligned 9-vectors (the 10th double is never read as data).
__global__ void pad10_kernel(const int* done, int run_when_done, const double* __restrict__ x,
                             double* __restrict__ x10, int n) {
  if (done != nullptr && ((*done != 0) != (run_when_done != 0))) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) x10[i + i / 9] = x[i];
}

// y = D_f^2 x + q  (implicit_schur_complement.cc:130-143), in place on q.
__global__ void add_diagonal_kernel(const double* __restrict__ Df, const double* __restrict__ x,
                                    double* __restrict__ q, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (Df != nullptr) q[i] += Df[i] * Df[i] * x[i];
}

// q += D_f^2 p; pq = p'q; alpha = rho / pq with the reference's failure tests (:190-214).
__global__ void __launch_bounds__(kVecThreads)
cg_pq_kernel(CgState* st, const double* __restrict__ Df, const double* __restrict__ p,
             double* __restrict__ q, int n) {
  if (st->done) return;
  __shared__ double sbuf[8];
  __shared__ double sums[3];
  double v[3] = {0.0, 0.0, 0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double pi = p[i];
    double qi = q[i];
    if (Df != nullptr) { const double d = Df[i]; qi += d * d * pi; q[i] = qi; }
    v[0] += pi * qi; v[1] += pi * pi; v[2] += qi * qi;
  }
  if (!FinishDots<3>(st, 2, v, sums, sbuf)) return;
  if (threadIdx.x == 0) {
    st->pq = sums[0];
    if (st->pq <= 0.0 || isinf(st->pq)) {
      st->norm_p = sqrt(sums[1]); st->norm_q = sqrt(sums[2]);
      Finish(st, CB200_LINEAR_SOLVER_NO_CONVERGENCE, kReasonIndefinite);
      return;
    }
    st->alpha = st->rho / st->pq;
    if (isinf(st->alpha)) { Finish(st, CB200_LINEAR_SOLVER_FAILURE, kReasonAlphaFailure); return; }
    if (st->iter <= kHistCap) st->alpha_hist[st->iter - 1] = st->alpha;  // x += alpha p follows
  }
}

// x += alpha p alone (residual-reset iterations; the fresh S x follows) (:216).
__global__ void cg_update_x_kernel(const CgState* st, const double* __restrict__ p,
                                   double* __restrict__ x, int n, double* trace, int max_trace) {
  if (st->done) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double xi = x[i] + st->alpha * p[i];
  x[i] = xi;
  if (trace != nullptr && st->iter <= max_trace) trace[(size_t)(st->iter - 1) * n + i] = xi;
}

// RESET = false: x += alpha p; r -= alpha q                       (:216, :227)
// RESET = true : x already updated; q holds F'(..) of x: r = rhs - (D_f^2 x + q)  (:221-225)
// then Q1 = -x'(rhs + r), zeta test, |r| test, max-iteration test (:229-300).
template <bool RESET>
__global__ void __launch_bounds__(kVecThreads)
cg_update_kernel(CgState* st, const double* __restrict__ rhs, const double* __restrict__ Df,
                 const double* __restrict__ p, const double* __restrict__ q, double* __restrict__ x,
                 double* __restrict__ r, int n, double* trace, int max_trace) {
  if (st->done) return;
  __shared__ double sbuf[8];
  __shared__ double sums[2];
  double v[2] = {0.0, 0.0};
  const double alpha = st->alpha;
  const int iter = st->iter;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double xi, ri;
    if (RESET) {
      xi = x[i];
      double sx = q[i];
      if (Df != nullptr) { const double d = Df[i]; sx += d * d * xi; }
      ri = rhs[i] - sx;
    } else {
      xi = x[i] + alpha * p[i];
      x[i] = xi;
      if (trace != nullptr && iter <= max_trace) trace[(size_t)(iter - 1) * n + i] = xi;
      ri = r[i] - alpha * q[i];
    }
    r[i] = ri;
    v[0] += xi * (rhs[i] + ri);
    v[1] += ri * ri;
  }
  if (!FinishDots<2>(st, 3, v, sums, sbuf)) return;
  if (threadIdx.x == 0) {
    st->Q1 = -sums[0];
    st->norm_r = sqrt(sums[1]);
    st->zeta = st->iter * (st->Q1 - st->Q0) / st->Q1;
    if (st->zeta < st->q_tolerance && st->iter >= st->min_iterations) {
      Finish(st, CB200_LINEAR_SOLVER_SUCCESS, kReasonZetaConvergence);
      return;
    }
    st->Q0 = st->Q1;
    if (st->norm_r <= st->tol_r && st->iter >= st->min_iterations) {
      Finish(st, CB200_LINEAR_SOLVER_SUCCESS, kReasonResidualConvergence);
      return;
    }
    if (st->iter >= st->max_iterations) Finish(st, CB200_LINEAR_SOLVER_NO_CONVERGENCE, kReasonMaxIterations);
  }
}

// small vector helpers for the power-series preconditioner and setup
__global__ void axpy_kernel(const int* done, double a, const double* __restrict__ x,
                            double* __restrict__ y, int n) {
  if (done != nullptr && *done) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] += a * x[i];
}
__global__ void sub_kernel(const double* __restrict__ a, const double* __restrict__ b,
                           double* __restrict__ out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = a[i] - b[i];
}
__global__ void __launch_bounds__(kVecThreads)
norm2_kernel(CgState* st, const double* __restrict__ x, int n, double* out) {
  __shared__ double sbuf[8];
  __shared__ double sums[1];
  double v[1] = {0.0};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) v[0] += x[i] * x[i];
  if (!FinishDots<1>(st, 0, v, sums, sbuf)) return;
  if (threadIdx.x == 0) *out = sums[0];
}

}  // namespace cb200
#endif  // CB200_CG_KERNELS_CUH_
