// One cooperative kernel per PCG iteration for the camera-sized vector work of
// ConjugateGradientsSolver (conjugate_gradients_solver.h:159-300) when the
// preconditioner is the identity or a 9x9 block diagonal (JACOBI, SCHUR_JACOBI):
// the fold of the S*p accumulators, the D_f^2 term, p'q, alpha, the x and r
// updates, Q1, |r|, every termination test, z = M^-1 r, rho, beta and the new p
// -- the "epilogue" of S*p fused with the "prologue" of the next S*p. One thread
// owns one camera (9 entries of every vector, one 9x9 block); the two scalar
// reductions per iteration go through per-CTA partials summed in index order by
// every CTA (deterministic, identical decisions everywhere) behind a grid
// barrier. Launched with cudaLaunchCooperativeKernel (all CTAs co-resident).
#ifndef CB200_CG_STEP_CUH_
#define CB200_CG_STEP_CUH_

#include "cg_kernels.cuh"

namespace cb200 {

constexpr int kStepThreads = 128;
constexpr int kStepCams = 14;  // cameras per CTA: 9 threads each
constexpr int kMaxPeers = 16;
constexpr unsigned long long kPeerTimeoutNs = 4000000000ull;  // a peer rank died or fell out of step

enum CgStepKind {
  kStepStart = 0,   // iteration 1 prologue: z = M^-1 r, rho, p = z          (:160-174)
  kStepNormal = 1,  // epilogue of iteration i + prologue of iteration i+1    (:190-300,160-188)
  kStepResetA = 2,  // residual-reset iteration, first half: pq, alpha, x     (:190-216)
  kStepResetB = 3   // second half: r = rhs - S x, tests, prologue of i+1     (:221-300,160-188)
};

struct CgStepArgs {
  CgState* st;
  const double* blocks;  // [C x 81] inverted preconditioner blocks (PRECOND 1)
  const double* Df;      // [9C] or null
  const double* rhs;
  double* x;
  double* r;
  double* p;
  double* x10;           // [10C] padded copy of the vector the next S* multiplies
  const double* q;       // [9C] S*p without D_f^2 when the fold was done elsewhere (multi-GPU), else null
  double* yrep;          // accumulators of the per-point kernel (10 per camera), folded and zeroed here
  int nrep;
  int rep_stride;
  int C;
  double* trace;
  int max_trace;
  // several GPUs, fused all-reduce over NVLink peer memory: every rank publishes its folded
  // partial in its own buffer (two slots, alternating) and reads all ranks' in rank order
  int world, rank;
  unsigned int* epoch;           // device counter of exchanges completed on this channel (CgState::peer_epoch):
                                 // this launch's exchange is number *epoch + 1, written back by CTA 0 -- kept
                                 // on the device so that launches replayed from a CUDA graph, and launches that
                                 // return early because the solve is over, stay in step on every rank
  size_t slot_stride;            // doubles between the two slots of a rank's buffer
  double* peer_q[kMaxPeers];     // [rank] -> that rank's buffer: 2 slots x slot_stride doubles
  unsigned int* peer_flag[kMaxPeers];  // [rank] -> that rank's 2 flags (one per slot, 128 bytes apart)
};

// Grid barrier on two ints of CgState: bar_flag counts completed barriers (it only
// grows), every CTA carries its own copy `gen` read at kernel entry. Safe because no
// barrier can complete before every CTA has arrived, i.e. has read its `gen`.
// Requires co-resident CTAs (cooperative launch).
__device__ __forceinline__ void GridBarrier(CgState* st, unsigned int& gen) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned int old = atomicAdd(&st->bar_count, 1u);
    if (old == gridDim.x - 1) {
      st->bar_count = 0;
      __threadfence();
      atomicExch(&st->bar_flag, gen + 1);
    } else {
      // (never hang the GPU: a CTA that is not co-resident -- impossible under a cooperative launch -- or a
      //  faulted sibling would otherwise spin here for ever; 4 s, checked every 4096 polls)
      unsigned long long t0 = 0;
      unsigned int polls = 0;
      while (*reinterpret_cast<volatile unsigned int*>(&st->bar_flag) == gen) {
        if ((++polls & 4095u) == 0u) {
          unsigned long long t1;
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
          if (t0 == 0) t0 = t1;
          else if (t1 - t0 > kPeerTimeoutNs) { atomicExch(&st->peer_timeout, 1); break; }
        }
      }
    }
    __threadfence();
  }
  ++gen;
  __syncthreads();
}

// Block partials -> grid barrier -> every CTA sums all partials in the same fixed order
// (thread t takes partials t, t + 128, ...; then the block tree): deterministic, and
// identical in every CTA, so all of them take the same decisions.
// `row0` selects the rows of CgState::partial: the two reductions of one launch
// use different rows, so a fast CTA cannot overwrite what a slow one still reads.
// K block-wide sums in one pass (fixed tree => deterministic); result valid in every thread.
template <int K>
__device__ __forceinline__ void BlockSums(double* v, double* sbuf /* >= 4 * K */) {
#pragma unroll
  for (int k = 0; k < K; ++k) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], off);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < K; ++k) sbuf[4 * k + warp] = v[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = (sbuf[4 * k] + sbuf[4 * k + 1]) + (sbuf[4 * k + 2] + sbuf[4 * k + 3]);
}

template <int K>
__device__ __forceinline__ void GridSums(CgState* st, unsigned int& gen, int row0, const double* vals,
                                         double* sums, double* sbuf) {
  double v[K];
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = vals[k];
  BlockSums<K>(v, sbuf);
#pragma unroll
  for (int k = 0; k < K; ++k)
    if (threadIdx.x == k) st->partial[row0 + k][blockIdx.x] = v[k];
  GridBarrier(st, gen);
#pragma unroll
  for (int k = 0; k < K; ++k) {
    double s = 0.0;
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += blockDim.x) s += __ldcg(&st->partial[row0 + k][b]);
    sums[k] = s;
  }
  BlockSums<K>(sums, sbuf);
}

// One thread per scalar entry: a CTA owns kStepCams cameras (9 threads each, 2 idle).
// Everything is coalesced: vectors at 9*camera + i, accumulators at 10*camera + i,
// preconditioner rows at 81*camera + 9*i.
template <int KIND, int PRECOND>
__global__ void __launch_bounds__(kStepThreads)
cg_step_kernel(const CgStepArgs a) {
  CgState* st = a.st;
  if (st->done || st->peer_timeout) return;
  __shared__ double sbuf[12];
  __shared__ double sR[kStepThreads];
  double sums[3];
  // scalars of the iteration, read before anything is written (CTA 0 writes after barrier 1)
  const int it = st->iter;
  double rho = st->rho;
  double Q0 = st->Q0;
  const double q_tol = st->q_tolerance, tol_r = st->tol_r;
  const int min_it = st->min_iterations, max_it = st->max_iterations;
  unsigned int gen = *reinterpret_cast<volatile unsigned int*>(&st->bar_flag);
  const bool writer = blockIdx.x == 0 && threadIdx.x == 0;
  const int local = threadIdx.x / 9;             // camera within the CTA
  const int i = threadIdx.x - 9 * local;         // entry within the camera
  const int c = blockIdx.x * kStepCams + local;
  const bool on = local < kStepCams && c < a.C;
  const size_t o9 = (size_t)c * 9 + i, o10 = (size_t)c * 10 + i;

  double p = 0.0, q = 0.0, x = 0.0, r = 0.0, z = 0.0;
  double alpha = 0.0;
  // everything this thread reads from global memory is requested up front: the loads are in flight
  // together instead of one L2 round trip after each grid barrier
  double r_old = 0.0, rhs_i = 0.0, df = 0.0;
  double Brow[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) Brow[k] = 0.0;
  if (on) {
    r_old = a.r[o9];
    rhs_i = a.rhs[o9];
    if (a.Df != nullptr) df = a.Df[o9];
    if (PRECOND != 0) {
      const double* B = a.blocks + (size_t)c * 81 + 9 * i;
#pragma unroll
      for (int k = 0; k < 9; ++k) Brow[k] = __ldg(B + k);
    }
  }

  if (KIND != kStepStart) {
    // ---- fold the accumulators: q = F'(..) of the multiplied vector ----
    if (on) {
      p = a.p[o9];
      x = a.x[o9];
      if (a.q != nullptr) {
        q = a.q[o9];
      } else {
        // all loads first (independent, pipelined), the zeroing stores afterwards
        for (int rep = 0; rep < a.nrep; ++rep) q += __ldcg(a.yrep + (size_t)rep * a.rep_stride + o10);
        for (int rep = 0; rep < a.nrep; ++rep) a.yrep[(size_t)rep * a.rep_stride + o10] = 0.0;
      }
    }
    if (a.world > 1 && a.q == nullptr) {
      // ---- all-reduce of q over the ranks through peer memory ----
      const unsigned int epoch = *reinterpret_cast<volatile unsigned int*>(a.epoch) + 1u;  // (read before barrier 1)
      const int slot = epoch & 1u;
      if (on) a.peer_q[a.rank][slot * a.slot_stride + o9] = q;
      GridBarrier(st, gen);  // every CTA of this rank has published (and has read the epoch counter)
      if (threadIdx.x == 0 && blockIdx.x == 0) {
        *a.epoch = epoch;
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(a.peer_flag[a.rank] + slot * 32), "r"(epoch) : "memory");
      }
      // one polling thread per peer: the NVLink round trips of the world - 1 flags overlap instead of adding up
      if (threadIdx.x < a.world && threadIdx.x != a.rank) {
        const int rk = threadIdx.x;
        unsigned int seen;
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        do {
          asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(a.peer_flag[rk] + slot * 32) : "memory");
          asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
          if (t1 - t0 > kPeerTimeoutNs) { atomicExch(&st->peer_timeout, 1); break; }  // never hang the GPU
        } while (seen != epoch);
      }
      __syncthreads();
      if (on) {
        // all peers' values requested at once (independent loads: one NVLink round trip, not world of them),
        // then added in rank order: the same sum, bit for bit, on every rank
        double v[kMaxPeers];
#pragma unroll
        for (int rk = 0; rk < kMaxPeers; ++rk)
          v[rk] = rk < a.world ? __ldcg(a.peer_q[rk] + slot * a.slot_stride + o9) : 0.0;
        q = v[0];
#pragma unroll
        for (int rk = 1; rk < kMaxPeers; ++rk)
          if (rk < a.world) q += v[rk];
      }
    }
    // + D_f^2 * (the multiplied vector): p, or x in the second half of a reset
    if (on) q += df * df * (KIND == kStepResetB ? x : p);
  }
  if (KIND == kStepNormal || KIND == kStepResetA) {
    // ---- pq, alpha (:190-214) ----
    const double v[3] = {p * q, p * p, q * q};
    GridSums<3>(st, gen, 0, v, sums, sbuf);
    const double pq = sums[0];
    if (pq <= 0.0 || isinf(pq)) {
      if (writer) {
        st->pq = pq; st->norm_p = sqrt(sums[1]); st->norm_q = sqrt(sums[2]);
        Finish(st, CB200_LINEAR_SOLVER_NO_CONVERGENCE, kReasonIndefinite);
      }
      return;
    }
    alpha = rho / pq;
    if (writer) { st->pq = pq; st->alpha = alpha; }
    if (isinf(alpha)) {
      if (writer) Finish(st, CB200_LINEAR_SOLVER_FAILURE, kReasonAlphaFailure);
      return;
    }
    if (writer && it <= kHistCap) st->alpha_hist[it - 1] = alpha;  // x += alpha p follows (stream_kernels.cuh: history)
    // ---- x += alpha p (:216) ----
    if (on) {
      x += alpha * p;
      a.x[o9] = x;
      if (a.trace != nullptr && it <= a.max_trace) a.trace[(size_t)(it - 1) * 9 * a.C + o9] = x;
      if (KIND == kStepResetA) a.x10[o10] = x;  // S x comes next
    }
    if (KIND == kStepResetA) return;
  }
  // ---- residual (:221-227) ----
  if (on) {
    if (KIND == kStepNormal) r = r_old - alpha * q;
    else if (KIND == kStepResetB) r = rhs_i - q;
    else r = r_old;
    if (KIND != kStepStart) a.r[o9] = r;
  }
  // ---- z = M^-1 r (:160-161): row i of the camera's block times the camera's r ----
  if (PRECOND == 0) {
    z = r;
  } else {
    sR[threadIdx.x] = r;
    __syncthreads();
    if (on) {
      const double* rc = sR + 9 * local;
      double s = 0.0;
#pragma unroll
      for (int k = 0; k < 9; ++k) s += Brow[k] * rc[k];
      z = s;
    }
  }
  double v[3] = {r * z, 0.0, 0.0};
  if (KIND != kStepStart && on) { v[1] = x * (rhs_i + r); v[2] = r * r; }
  GridSums<3>(st, gen, 3, v, sums, sbuf);
  int next_it = it;
  if (KIND != kStepStart) {
    // ---- Q1, zeta, |r|, iteration limit (:229-300) ----
    const double Q1 = -sums[1];
    const double norm_r = sqrt(sums[2]);
    const double zeta = it * (Q1 - Q0) / Q1;
    int term = -1, reason = 0;
    if (zeta < q_tol && it >= min_it) { term = CB200_LINEAR_SOLVER_SUCCESS; reason = kReasonZetaConvergence; }
    else if (norm_r <= tol_r && it >= min_it) { term = CB200_LINEAR_SOLVER_SUCCESS; reason = kReasonResidualConvergence; }
    else if (it >= max_it) { term = CB200_LINEAR_SOLVER_NO_CONVERGENCE; reason = kReasonMaxIterations; }
    if (writer) {
      st->Q1 = Q1; st->norm_r = norm_r; st->zeta = zeta;
      if (reason != kReasonZetaConvergence) st->Q0 = Q1;
      if (term >= 0) Finish(st, term, reason);
    }
    if (term >= 0) return;
  }
  // ---- prologue of the next iteration: rho, beta, p (:163-188) ----
  next_it = it + 1;
  const double last_rho = rho;
  rho = sums[0];
  double beta = 0.0;
  int fail = 0;
  if (IsZeroOrInfinity(rho)) fail = kReasonRhoFailure;
  else if (next_it > 1) {
    beta = rho / last_rho;
    if (IsZeroOrInfinity(beta)) fail = kReasonBetaFailure;
  }
  if (writer) {
    st->iter = next_it; st->last_rho = last_rho; st->rho = rho;
    if (next_it > 1 && fail != kReasonRhoFailure) st->beta = beta;
    if (fail) Finish(st, CB200_LINEAR_SOLVER_FAILURE, fail);
  }
  if (fail) return;
  if (on) {
    const double pk = (next_it == 1) ? z : z + beta * p;
    a.p[o9] = pk;
    a.x10[o10] = pk;  // S p comes next
  }
}

}  // namespace cb200

namespace cb200 {

// ----------------------------------------------------------------------------
// One term of the power-series preconditioner (power_series_expansion_preconditioner.cc:57-82) after its pass
// over the Jacobian: fold the accumulators (q = F'E (E'E)^-1 E'F * previous term), term = (F'F + D_f^2)^-1 q,
// y += term, and the padded copy the next pass multiplies -- one kernel instead of fold + block multiply + axpy +
// pad. FIRST: the zeroth term y = (F'F + D_f^2)^-1 x. Same thread layout as cg_step_kernel (a CTA owns 14
// cameras); no grid-wide dependency, so a plain launch. q_in non-null: q arrives folded (several GPUs,
// deterministic mode).
// ----------------------------------------------------------------------------
template <bool FIRST>
__global__ void __launch_bounds__(kStepThreads)
pse_term_kernel(const int* done, const double* __restrict__ blocks, const double* __restrict__ q_in,
                double* __restrict__ yrep, int nrep, int rep_stride, int C, double* __restrict__ y,
                double* __restrict__ term_out, double* __restrict__ x10) {
  if (done != nullptr && *done) return;
  __shared__ double sQ[kStepThreads];
  const int local = threadIdx.x / 9;
  const int i = threadIdx.x - 9 * local;
  const int c = blockIdx.x * kStepCams + local;
  const bool on = local < kStepCams && c < C;
  const size_t o9 = (size_t)c * 9 + i, o10 = (size_t)c * 10 + i;
  double q = 0.0;
  double Brow[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) Brow[k] = 0.0;
  if (on) {
    const double* B = blocks + (size_t)c * 81 + 9 * i;
#pragma unroll
    for (int k = 0; k < 9; ++k) Brow[k] = __ldg(B + k);
    if (q_in != nullptr) {
      q = q_in[o9];
    } else {
      for (int rep = 0; rep < nrep; ++rep) q += __ldcg(yrep + (size_t)rep * rep_stride + o10);
      for (int rep = 0; rep < nrep; ++rep) yrep[(size_t)rep * rep_stride + o10] = 0.0;
    }
  }
  sQ[threadIdx.x] = q;
  __syncthreads();
  if (!on) return;
  const double* qc = sQ + 9 * local;
  double t = 0.0;
#pragma unroll
  for (int k = 0; k < 9; ++k) t += Brow[k] * qc[k];
  y[o9] = FIRST ? t : y[o9] + t;
  term_out[o9] = t;
  x10[o10] = t;
}

// ----------------------------------------------------------------------------
// The same exchange for the per-solve sums (reduced rhs, preconditioner blocks, the
// scalars of the trust-region loop): buf[0..n) <- sum (or max) over ranks, in rank
// order, through the peer-mapped buffers. ONE plain launch of <= 64 CTAs: every CTA
// copies its part of the local partial into this rank's slot; the last one to do so
// releases the epoch flag; every CTA then waits for all peers' flags and adds the
// slots up. A waiting CTA only needs CTAs of OTHER GPUs (and its own grid's, all of
// which fit the device at once) to make progress, so no cooperative launch is needed.
// The exchange number lives on the device (CgState::peer_epoch[channel]) and is
// advanced by the last CTA to finish, i.e. after every CTA has read it.
// ----------------------------------------------------------------------------
struct PeerRefs {
  int world, rank;
  int channel;                   // 0: main stream, 1: side stream (own buffers, flags, counter, tickets)
  size_t slot_stride;
  unsigned int* epoch;           // &CgState::peer_epoch[channel]
  double* base[kMaxPeers];
  unsigned int* flag[kMaxPeers];
};

template <bool MAX>
__global__ void __launch_bounds__(256)
peer_allreduce_kernel(const PeerRefs p, double* __restrict__ buf, size_t n, CgState* st) {
  const unsigned int epoch = *reinterpret_cast<volatile unsigned int*>(p.epoch) + 1u;
  const int slot = epoch & 1u;
  double* mine = p.base[p.rank] + slot * p.slot_stride;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    mine[i] = buf[i];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* ticket = &st->ticket[4 + 2 * p.channel];
    const unsigned int t = atomicAdd(ticket, 1u);
    if (t == gridDim.x - 1) {
      *ticket = 0;
      __threadfence_system();
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p.flag[p.rank] + slot * 32), "r"(epoch) : "memory");
    }
  }
  // one polling thread per rank (own flag included: the sum below also reads this rank's slot, complete only
  // when every CTA of this grid has published): the NVLink round trips overlap instead of adding up
  if (threadIdx.x < p.world && (threadIdx.x != p.rank || gridDim.x > 1)) {
    const int r = threadIdx.x;
    unsigned int seen;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(p.flag[r] + slot * 32) : "memory");
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > kPeerTimeoutNs) { atomicExch(&st->peer_timeout, 1); break; }
    } while (seen != epoch);
  }
  __syncthreads();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double v[kMaxPeers];  // independent loads first, then the rank-ordered sum: bit-identical on every rank
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r) v[r] = r < p.world ? __ldcg(p.base[r] + slot * p.slot_stride + i) : 0.0;
    double s = v[0];
#pragma unroll
    for (int r = 1; r < kMaxPeers; ++r)
      if (r < p.world) s = MAX ? fmax(s, v[r]) : s + v[r];
    buf[i] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* ticket = &st->ticket[5 + 2 * p.channel];
    __threadfence();
    const unsigned int t = atomicAdd(ticket, 1u);
    if (t == gridDim.x - 1) { *ticket = 0; *p.epoch = epoch; }
  }
}

}  // namespace cb200
#endif  // CB200_CG_STEP_CUH_
