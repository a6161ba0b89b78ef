// Shared device helpers (mbarrier / bulk-copy wrappers, small dense inverses,
// block reductions) of the hand-written FP64 CUDA kernels (sm_100a) for the
// ITERATIVE_SCHUR path of a <2,3,9> bundle-adjustment Jacobian. See DESIGN.md for the data layout and the
// per-kernel byte counts. Reference functions each kernel replaces are cited at
// the kernel.
//
// Device layout of one point shard (all indices local to the shard):
//   E   [N x 6]   2x3 row-major point blocks, rows grouped by point
//   F   [N x 18]  2x9 row-major camera blocks, same row order
//   cam [N]       camera index of each row
//   row_start [P+1]  first row of each point
//   ete_inv [P x 6]  upper triangle of (E'E + D_e^2)^-1
//   tiles [T]        TileMeta side-band records (flatten.h)
// Camera-sized vectors (9C) are replicated on every GPU.
#ifndef CB200_KERNELS_CUH_
#define CB200_KERNELS_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "flatten.h"

namespace cb200 {


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
// The same copy with an L2 eviction-priority hint (createpolicy result).
__device__ __forceinline__ void bulk_g2s_hint(void* dst, const void* src, uint32_t bytes,
                                              uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
// Streamed-once data must not push the small re-used working set (camera
// vectors, accumulators, preconditioner blocks, tile side-band, (E'E)^-1) out of
// the 126 MB L2.
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
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
  double rd[n];  // 1 / U(i,i): one division per pivot, the solves multiply
#pragma unroll
  for (int j = 0; j < n; ++j) {
#pragma unroll
    for (int i = 0; i <= j; ++i) {
      double s = m[i * n + j];
#pragma unroll
      for (int k = 0; k < i; ++k) s -= U[k * n + i] * U[k * n + j];
      if (i == j) { U[j * n + j] = sqrt(s); rd[j] = 1.0 / U[j * n + j]; }
      else U[i * n + j] = s * rd[i];
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
      y[i] = s * rd[i];
    }
#pragma unroll
    for (int i = n - 1; i >= 0; --i) {
      double s = y[i];
#pragma unroll
      for (int k = i + 1; k < n; ++k) s -= U[i * n + k] * y[k];
      y[i] = s * rd[i];
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

// block-wide sum of one double per thread (<= 256 threads); result valid
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

}  // namespace cb200
#endif  // CB200_KERNELS_CUH_
