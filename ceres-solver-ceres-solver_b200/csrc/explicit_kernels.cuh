// Explicit reduced camera matrix (Solver::Options::use_explicit_schur_complement with
// ITERATIVE_SCHUR, linear_solver.cc:111-116): S is formed once per solve as a block-sparse
// matrix of 9x9 cells (upper block triangle, explicit_plan.h) and PCG multiplies by it.
//
//   schur_form_kernel      SchurEliminator<2,3,9>::Eliminate into a BlockRandomAccessSparseMatrix
//                          (schur_eliminator_impl.h:176-304: ChunkDiagonalBlockAndGradient :441-505,
//                          ChunkOuterProduct :511-561, EBlockRowOuterProduct :664-714); the reduced
//                          rhs comes from the fused rhs pass the implicit path already has.
//   explicit_diag_kernel   D_f^2 on the diagonal (:193-214) and the SCHUR_JACOBI preconditioner =
//                          inverted diagonal cells (schur_complement_solver.cc:361-383,
//                          block_random_access_diagonal_matrix.cc:90-100)
//   explicit_spmv_kernel   BlockRandomAccessSparseMatrix::SymmetricRightMultiplyAndAccumulate
//                          (block_random_access_sparse_matrix.cc:128-170)
#ifndef CB200_EXPLICIT_KERNELS_CUH_
#define CB200_EXPLICIT_KERNELS_CUH_

#include "explicit_plan.h"
#include "kernels.cuh"

namespace cb200 {

constexpr int kFormWarps = 4;
constexpr int kFormThreads = 32 * kFormWarps;

struct FormArgs {
  const double* E;
  const double* F;
  const double* ete_inv;        // [P_loc x 6] upper triangle of (E'E + D_e^2)^-1
  const int* row_start;         // P_loc + 1
  const int* srow;              // camera-sorted rows of every point
  const long long* pair_start;  // P_loc + 1
  const unsigned int* pair_cell;
  const FormItem* items;
  int num_items;
  int* counter;                 // work queue head (zeroed before the launch)
  double* S;                    // cells, 81 doubles each
};

// G_r = F_r' E_r (9x3, row-major) of `n` rows of a point, into shared memory.
__device__ __forceinline__ void StageG(const FormArgs& a, const int* rows, int n, double* sG, int lane) {
  for (int idx = lane; idx < n * 27; idx += 32) {
    const int row = idx / 27, e = idx - row * 27;
    const int j = e / 3, m = e - 3 * j;
    const size_t r = (size_t)rows[row];
    const double* f = a.F + r * 18;
    const double* ee = a.E + r * 6;
    sG[idx] = f[j] * ee[m] + f[9 + j] * ee[3 + m];
  }
}

// One warp per item = up to kFormRows rows u of one point against all its rows v >= u (camera
// order). Lane l < 27 owns the elements l, l + 27, l + 54 of a 9x9 cell: three runs of 27
// consecutive doubles per reduction instruction. Every pair is one
// -(G_u inv) G_v' (+ F_u'F_u when u == v) added to its cell with RED.ADD.F64.
__global__ void __launch_bounds__(kFormThreads) schur_form_kernel(FormArgs a) {
  __shared__ double sH_all[kFormWarps][kFormRows * 27];
  __shared__ double sG_all[kFormWarps][kFormRows * 27];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  double* sH = sH_all[warp];
  double* sG = sG_all[warp];
  const int j = lane % 9, ib = lane / 9;  // element rows ib, ib + 3, ib + 6; column j
  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(a.counter, 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= a.num_items) break;
    const FormItem it = a.items[item];
    const int r0 = a.row_start[it.point];
    const int k = a.row_start[it.point + 1] - r0;
    const int u0 = it.u0;
    const int nu = min(kFormRows, k - u0);
    const double* iv = a.ete_inv + (size_t)it.point * 6;
    const double i00 = iv[0], i01 = iv[1], i02 = iv[2], i11 = iv[3], i12 = iv[4], i22 = iv[5];
    __syncwarp();
    StageG(a, a.srow + r0 + u0, nu, sG, lane);
    __syncwarp();
    for (int idx = lane; idx < nu * 27; idx += 32) {  // H = G inv
      const int m = idx % 3;
      const double* g = sG + (idx - m);
      const double c0 = m == 0 ? i00 : (m == 1 ? i01 : i02);
      const double c1 = m == 0 ? i01 : (m == 1 ? i11 : i12);
      const double c2 = m == 0 ? i02 : (m == 1 ? i12 : i22);
      sH[idx] = g[0] * c0 + g[1] * c1 + g[2] * c2;
    }
    __syncwarp();
    const long long pair_base = a.pair_start[it.point];
    for (int v0 = u0; v0 < k; v0 += kFormRows) {
      const int nv = min(kFormRows, k - v0);
      if (v0 != u0) {
        __syncwarp();
        StageG(a, a.srow + r0 + v0, nv, sG, lane);
        __syncwarp();
      }
      for (int u = 0; u < nu; ++u) {
        const long long U = u0 + u;
        double h[3][3];
#pragma unroll
        for (int q = 0; q < 3; ++q)
#pragma unroll
          for (int m = 0; m < 3; ++m) h[q][m] = lane < 27 ? sH[u * 27 + (ib + 3 * q) * 3 + m] : 0.0;
        const unsigned int* cells = a.pair_cell + (pair_base + U * k - U * (U - 1) / 2 - U);
        for (int v = (v0 == u0 ? u : 0); v < nv; ++v) {
          const int V = v0 + v;
          const unsigned int cell = cells[V];
          if (lane < 27) {
            const double g0 = sG[v * 27 + j * 3], g1 = sG[v * 27 + j * 3 + 1], g2 = sG[v * 27 + j * 3 + 2];
            double out[3];
#pragma unroll
            for (int q = 0; q < 3; ++q) out[q] = -(h[q][0] * g0 + h[q][1] * g1 + h[q][2] * g2);
            if (V == (int)U) {  // S += F'F of the row itself (EBlockRowOuterProduct)
              const double* f = a.F + (size_t)a.srow[r0 + V] * 18;
              const double fj0 = f[j], fj1 = f[9 + j];
#pragma unroll
              for (int q = 0; q < 3; ++q) out[q] += f[ib + 3 * q] * fj0 + f[9 + ib + 3 * q] * fj1;
            }
            double* dst = a.S + (size_t)(cell & ~kFormDuplicate) * 81;
#pragma unroll
            for (int q = 0; q < 3; ++q) atomicAdd(dst + lane + 27 * q, out[q]);
            if (cell & kFormDuplicate) {  // two rows of one camera: the cell also takes the transpose
#pragma unroll
              for (int q = 0; q < 3; ++q) atomicAdd(dst + j * 9 + ib + 3 * q, out[q]);
            }
          }
        }
      }
    }
  }
}

// D_f^2 on the diagonal cells (in place) and sj = inverse of the diagonal cell
// (Cholesky of the upper triangle).
__global__ void __launch_bounds__(64)
explicit_diag_kernel(double* __restrict__ S, const int* __restrict__ row_ptr, const double* __restrict__ Df,
                     int C, double* __restrict__ sj) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double* cell = S + (size_t)row_ptr[c] * 81;
  double m[81];
#pragma unroll
  for (int k = 0; k < 81; ++k) m[k] = cell[k];
  if (Df != nullptr) {
#pragma unroll
    for (int i = 0; i < 9; ++i) {
      const double d = Df[9 * (size_t)c + i];
      m[i * 10] += d * d;
      cell[i * 10] = m[i * 10];
    }
  }
  LltInverseUpper<9>(m);
#pragma unroll
  for (int k = 0; k < 81; ++k) sj[(size_t)c * 81 + k] = m[k];
}

constexpr int kSpmvWarps = 8;
constexpr int kSpmvThreads = 32 * kSpmvWarps;

// y = S x with the upper block triangle stored: CTA i gathers the cells (i, j >= i) of its row
// and, through the column lists, the cells (r < i, i) transposed. Lane l < 27 holds the
// elements (a = l / 3, 3 (l % 3) .. + 2) of a cell: one coalesced 648-byte read per cell.
// No atomics, fixed summation order: every GPU computes bit-identical vectors.
__global__ void __launch_bounds__(kSpmvThreads)
explicit_spmv_kernel(const int* done, const double* __restrict__ S, const int* __restrict__ row_ptr,
                     const int* __restrict__ col, const int* __restrict__ t_ptr,
                     const int* __restrict__ t_cell, const int* __restrict__ t_row,
                     const double* __restrict__ x, double* __restrict__ y) {
  if (done != nullptr && *done) return;
  __shared__ double s_up[kSpmvWarps][27];
  __shared__ double s_lo[kSpmvWarps][27][3];
  const int i = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int a = lane / 3, g = lane - 3 * a;
  const bool active = lane < 27;
  const int u_begin = row_ptr[i], nU = row_ptr[i + 1] - u_begin;
  const int l_begin = t_ptr[i], nL = t_ptr[i + 1] - l_begin;
  double up = 0.0, lo0 = 0.0, lo1 = 0.0, lo2 = 0.0;
  // two cells per step: both cells' six loads are in flight before the first multiply
  for (int w = warp; w < nU; w += 2 * kSpmvWarps) {
    const int k0 = u_begin + w;
    const bool two = w + kSpmvWarps < nU;
    const int k1 = two ? k0 + kSpmvWarps : k0;
    const int j0 = col[k0], j1 = col[k1];
    if (active) {
      const double* c0 = S + (size_t)k0 * 81 + a * 9 + 3 * g;
      const double* c1 = S + (size_t)k1 * 81 + a * 9 + 3 * g;
      const double* x0 = x + (size_t)j0 * 9 + 3 * g;
      const double* x1 = x + (size_t)j1 * 9 + 3 * g;
      const double a0 = c0[0], a1 = c0[1], a2 = c0[2], b0 = c1[0], b1 = c1[1], b2 = c1[2];
      const double p0 = x0[0], p1 = x0[1], p2 = x0[2], q0 = x1[0], q1 = x1[1], q2 = x1[2];
      up += a0 * p0 + a1 * p1 + a2 * p2;
      if (two) up += b0 * q0 + b1 * q1 + b2 * q2;
    }
  }
  for (int w = warp; w < nL; w += 2 * kSpmvWarps) {
    const bool two = w + kSpmvWarps < nL;
    const int t0 = l_begin + w, t1 = two ? t0 + kSpmvWarps : t0;
    const int k0 = t_cell[t0], k1 = t_cell[t1];
    const int r0 = t_row[t0], r1 = t_row[t1];
    if (active) {
      const double* c0 = S + (size_t)k0 * 81 + a * 9 + 3 * g;
      const double* c1 = S + (size_t)k1 * 81 + a * 9 + 3 * g;
      const double a0 = c0[0], a1 = c0[1], a2 = c0[2], b0 = c1[0], b1 = c1[1], b2 = c1[2];
      const double xr0 = x[(size_t)r0 * 9 + a], xr1 = two ? x[(size_t)r1 * 9 + a] : 0.0;
      lo0 += a0 * xr0; lo1 += a1 * xr0; lo2 += a2 * xr0;
      lo0 += b0 * xr1; lo1 += b1 * xr1; lo2 += b2 * xr1;
    }
  }
  if (active) {
    s_up[warp][lane] = up;
    s_lo[warp][lane][0] = lo0; s_lo[warp][lane][1] = lo1; s_lo[warp][lane][2] = lo2;
  }
  __syncthreads();
  if (threadIdx.x < 9) {
    const int t = threadIdx.x;
    double sum = 0.0;
    for (int w = 0; w < kSpmvWarps; ++w) {
      for (int gg = 0; gg < 3; ++gg) sum += s_up[w][3 * t + gg];
      for (int aa = 0; aa < 9; ++aa) sum += s_lo[w][3 * aa + t / 3][t % 3];
    }
    y[(size_t)i * 9 + t] = sum;
  }
}

}  // namespace cb200
#endif  // CB200_EXPLICIT_KERNELS_CUH_
