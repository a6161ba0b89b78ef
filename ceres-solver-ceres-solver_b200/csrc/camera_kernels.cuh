// Camera-side kernels: the 9x9 block-diagonal preconditioners and the separate
// E / F block-SpMV products of the PartitionedMatrixView seam.
#ifndef CB200_CAMERA_KERNELS_CUH_
#define CB200_CAMERA_KERNELS_CUH_

#include "kernels.cuh"

namespace cb200 {

constexpr int kChunkObs = 2048;     // observations of one camera handled by one CTA
constexpr int kChunkThreads = 128;

// ----------------------------------------------------------------------------
// Block diagonal of the Schur complement (SCHUR_JACOBI) or of F'F (JACOBI, SPSE).
//
// Reference: SchurEliminator<2,3,9>::Eliminate into a BlockRandomAccessDiagonalMatrix
// (schur_eliminator_impl.h:176-304: ChunkDiagonalBlockAndGradient :441-505,
// ChunkOuterProduct :511-561, EBlockRowOuterProduct :664-714) walks the points and
// updates the 9x9 camera blocks under a per-block mutex. Here the sum is
// re-associated per camera: S_cc = D_c^2 + sum_{rows r of c} F_r' W_r F_r with
// W_r = I - E_r ete^-1 E_r' (from point_ete_kernel), so every CTA owns its output
// and the order of the sum is fixed. WEIGHTED = false gives F'F
// (UpdateBlockDiagonalFtF, partitioned_matrix_view_impl.h:530-660).
// One CTA per chunk (<= kChunkObs rows of one camera, listed camera-major in
// cam_perm); out: the 45 upper-triangle sums of the chunk.
// ----------------------------------------------------------------------------
constexpr int kChunkDepth = 3;     // 32-row groups in flight per warp (cp.async pipeline)
constexpr int kChunkSlot = 176;    // bytes per staged row: F (144) + weights (24), 16-byte aligned
constexpr int kChunkSmemBytes = kChunkDepth * kChunkThreads * kChunkSlot;

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(void* dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// The rows of a camera are scattered over the point-major Jacobian. Every warp
// streams groups of 32 rows through its own kChunkDepth-deep cp.async ring; a
// group's 32 x 144 bytes of F are fetched as 288 16-byte pieces, lane l taking
// pieces j*32 + l, so that one request touches the few cache lines of 3-4 rows
// instead of 32 (the L1 tag stage looks up one line per cycle). After the
// group has landed each lane owns one row and accumulates its 45 upper-triangle
// products. No CTA barrier in the loop.
template <bool WEIGHTED>
__global__ void __launch_bounds__(kChunkThreads)
camera_block_kernel(const double* __restrict__ F, const double* __restrict__ weights,
                    const int* __restrict__ cam_perm, const int* __restrict__ chunk_begin,
                    double* __restrict__ chunk_partial) {
  extern __shared__ __align__(16) unsigned char ring[];
  __shared__ double sred[kChunkThreads / 32][45];
  constexpr int kWarps = kChunkThreads / 32;
  constexpr int kGroupBytes = 32 * kChunkSlot;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int begin = chunk_begin[blockIdx.x], end = chunk_begin[blockIdx.x + 1];
  unsigned char* mine = ring + warp * (kChunkDepth * kGroupBytes);
  // group g of this warp = rows [begin + (g * kWarps + warp) * 32, +32)
  auto group_row0 = [&](int g) { return begin + (g * kWarps + warp) * 32; };
  auto load_index = [&](int g) {
    const int i = group_row0(g) + lane;
    return i < end ? __ldg(cam_perm + i) : -1;
  };
  auto issue = [&](int d, int r_own) {  // r_own: row of lane's slot in the group, -1 past the end
    unsigned char* dst = mine + d * kGroupBytes;
#pragma unroll
    for (int j = 0; j < 9; ++j) {
      const int idx = j * 32 + lane;
      const int o = idx / 9, pc = idx - 9 * o;
      const int r = __shfl_sync(0xffffffffu, r_own, o);
      if (r >= 0) cp_async16(dst + o * kChunkSlot + 16 * pc, F + (size_t)r * 18 + 2 * pc);
    }
    if (WEIGHTED) {
      if (kWeightStride == 4) {  // one aligned 32-byte sector per row: two 16-byte pieces
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int idx = j * 32 + lane;
          const int o = idx >> 1, pc = idx & 1;
          const int r = __shfl_sync(0xffffffffu, r_own, o);
          if (r >= 0) cp_async16(dst + o * kChunkSlot + 144 + 16 * pc, weights + (size_t)r * 4 + 2 * pc);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const int idx = j * 32 + lane;
          const int o = idx / 3, pc = idx - 3 * o;
          const int r = __shfl_sync(0xffffffffu, r_own, o);
          if (r >= 0) cp_async8(dst + o * kChunkSlot + 144 + 8 * pc, weights + (size_t)r * 3 + pc);
        }
      }
    }
    cp_async_commit();
  };
  double acc[45];
#pragma unroll
  for (int k = 0; k < 45; ++k) acc[k] = 0.0;
  const int rows = end - begin;
  const int my_groups = (rows - warp * 32 + kWarps * 32 - 1) / (kWarps * 32);  // groups with >= 1 row (<= 0: none)
#pragma unroll
  for (int d = 0; d < kChunkDepth; ++d) issue(d, d < my_groups ? load_index(d) : -1);
  // row indices of the next refill: loaded one iteration before the copies that need them
  int r_next = kChunkDepth < my_groups ? load_index(kChunkDepth) : -1;
  int d = 0;
  for (int g = 0; g < my_groups; ++g) {
    cp_async_wait<kChunkDepth - 1>();
    __syncwarp();  // every lane's pieces of group g have landed
    const bool live = group_row0(g) + lane < end;
    const unsigned char* row = mine + d * kGroupBytes + lane * kChunkSlot;
    const double2* f2 = reinterpret_cast<const double2*>(row);
    double f[18];
#pragma unroll
    for (int k = 0; k < 9; ++k) { const double2 v = f2[k]; f[2 * k] = v.x; f[2 * k + 1] = v.y; }
    double w00 = 1.0, w01 = 0.0, w11 = 1.0;
    if (WEIGHTED) {
      const double* w = reinterpret_cast<const double*>(row + 144);
      w00 = w[0]; w01 = w[1]; w11 = w[2];
    }
    __syncwarp();  // the slot has been read by all lanes: refill it
    issue(d, r_next);
    r_next = (g + kChunkDepth + 1 < my_groups) ? load_index(g + kChunkDepth + 1) : -1;
    d = (d + 1 == kChunkDepth) ? 0 : d + 1;
    if (!live) continue;  // (stale bytes of an earlier group)
    double g0[9], g1[9];  // W * F
    if (WEIGHTED) {
#pragma unroll
      for (int k = 0; k < 9; ++k) {
        g0[k] = w00 * f[k] + w01 * f[9 + k];
        g1[k] = w01 * f[k] + w11 * f[9 + k];
      }
    } else {
#pragma unroll
      for (int k = 0; k < 9; ++k) { g0[k] = f[k]; g1[k] = f[9 + k]; }
    }
    int q = 0;
#pragma unroll
    for (int i2 = 0; i2 < 9; ++i2)
#pragma unroll
      for (int j = i2; j < 9; ++j) { acc[q] += f[i2] * g0[j] + f[9 + i2] * g1[j]; ++q; }
  }
  cp_async_wait<0>();
#pragma unroll
  for (int k = 0; k < 45; ++k) {
    double v = acc[k];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0) sred[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 45) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kChunkThreads / 32; ++w) s += sred[w][threadIdx.x];
    chunk_partial[(size_t)blockIdx.x * 45 + threadIdx.x] = s;
  }
}

// Deterministic mode (cb200_options::deterministic): y_c = sum over the rows r of camera c of F_r' w_r, the
// camera scatter of a point pass (LeftMultiplyAndAccumulateF, partitioned_matrix_view_impl.h:266-375, whose
// multi-threaded form also walks the rows camera by camera) summed in a FIXED order instead of with FP64
// atomics: thread t of the chunk's CTA takes rows t, t + 128, ... in order, then a fixed shuffle / shared-memory
// tree. Same chunks as camera_block_kernel; out: 9 sums per chunk, folded per camera by
// camera_gather_finish_kernel. A repeatability option, not the fast path: F is gathered camera-major.
__global__ void __launch_bounds__(kChunkThreads)
camera_gather_kernel(const int* done, const double* __restrict__ F, const double* __restrict__ w_rows,
                     const int* __restrict__ cam_perm, const int* __restrict__ chunk_begin,
                     double* __restrict__ chunk_partial9) {
  if (done != nullptr && *done) return;
  __shared__ double sred[kChunkThreads / 32][9];
  const int begin = chunk_begin[blockIdx.x], end = chunk_begin[blockIdx.x + 1];
  double acc[9];
#pragma unroll
  for (int k = 0; k < 9; ++k) acc[k] = 0.0;
  for (int i = begin + threadIdx.x; i < end; i += kChunkThreads) {
    const int r = __ldg(cam_perm + i);
    const double2 w = reinterpret_cast<const double2*>(w_rows)[r];
    const double2* f2 = reinterpret_cast<const double2*>(F + (size_t)r * 18);
    double f[18];
#pragma unroll
    for (int k = 0; k < 9; ++k) { const double2 v = f2[k]; f[2 * k] = v.x; f[2 * k + 1] = v.y; }
#pragma unroll
    for (int k = 0; k < 9; ++k) acc[k] += f[k] * w.x + f[9 + k] * w.y;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    double v = acc[k];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    if (lane == 0) sred[warp][k] = v;
  }
  __syncthreads();
  if (threadIdx.x < 9) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kChunkThreads / 32; ++w) s += sred[w][threadIdx.x];
    chunk_partial9[(size_t)blockIdx.x * 9 + threadIdx.x] = s;
  }
}
__global__ void __launch_bounds__(256)
camera_gather_finish_kernel(const int* done, const double* __restrict__ chunk_partial9,
                            const int* __restrict__ cam_chunk_start, int C, double* __restrict__ out /* 9C */) {
  if (done != nullptr && *done) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 9 * C) return;
  const int c = i / 9, k = i - 9 * c;
  double s = 0.0;
  for (int ch = cam_chunk_start[c]; ch < cam_chunk_start[c + 1]; ++ch) s += chunk_partial9[(size_t)ch * 9 + k];
  out[i] = s;
}

// Cross terms of two rows (ri != rj) of the same point that see the same camera
// (the eliminator merges them in its per-camera buffer before the outer product,
// schur_eliminator_impl.h:140-148,491-501): S_cc -= F_i' E_i c E_j' F_j + transpose.
// Rare; one thread per unordered pair, FP64 atomics into dup_blocks [C x 81].
__global__ void duplicate_pair_kernel(const double* __restrict__ E, const double* __restrict__ F,
                                      const int* __restrict__ cam, const int* __restrict__ row_start,
                                      const double* __restrict__ De, const int* __restrict__ pairs,
                                      int num_pairs, double* __restrict__ dup_blocks) {
  // (grid-stride: deterministic mode runs the whole list on ONE thread so the additions keep their order)
  for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < num_pairs; t += gridDim.x * blockDim.x) {
  const int p = pairs[3 * t], ri = pairs[3 * t + 1], rj = pairs[3 * t + 2];
  double m[9];
  for (int k = 0; k < 9; ++k) m[k] = 0.0;
  for (int r = row_start[p]; r < row_start[p + 1]; ++r) {
    const double* e = E + (size_t)r * 6;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) m[i * 3 + j] += e[i] * e[j] + e[3 + i] * e[3 + j];
  }
  if (De != nullptr)
    for (int i = 0; i < 3; ++i) { const double d = De[3 * (size_t)p + i]; m[i * 4] += d * d; }
  double c[9];
  Inverse3x3(m, c);
  const double* ei = E + (size_t)ri * 6;
  const double* ej = E + (size_t)rj * 6;
  double X[4];  // X = E_i c E_j'
  for (int a = 0; a < 2; ++a) {
    double g[3];
    for (int j = 0; j < 3; ++j) g[j] = ei[3 * a] * c[j] + ei[3 * a + 1] * c[3 + j] + ei[3 * a + 2] * c[6 + j];
    for (int b = 0; b < 2; ++b) X[2 * a + b] = g[0] * ej[3 * b] + g[1] * ej[3 * b + 1] + g[2] * ej[3 * b + 2];
  }
  const double* fi = F + (size_t)ri * 18;
  const double* fj = F + (size_t)rj * 18;
  double* out = dup_blocks + (size_t)cam[ri] * 81;
  for (int a = 0; a < 9; ++a)
    for (int b = 0; b < 9; ++b) {
      // (F_i' X F_j)(a,b) + (F_j' X' F_i)(a,b)
      const double v1 = fi[a] * (X[0] * fj[b] + X[1] * fj[9 + b]) + fi[9 + a] * (X[2] * fj[b] + X[3] * fj[9 + b]);
      const double v2 = fj[a] * (X[0] * fi[b] + X[2] * fi[9 + b]) + fj[9 + a] * (X[1] * fi[b] + X[3] * fi[9 + b]);
      atomicAdd(out + a * 9 + b, -(v1 + v2));
    }
  }
}

// Sum the chunk partials of each camera in chunk order, add D_c^2 (and the
// duplicate-pair corrections), mirror to a full row-major 9x9 and optionally
// invert (BlockRandomAccessDiagonalMatrix::Invert,
// block_random_access_diagonal_matrix.cc:90-100; AddDiagonalAndInvert for F'F,
// implicit_schur_complement.cc:179-204). One thread per camera. `partial_in`
// non-null: blocks already hold all-reduced upper triangles [C x 45] instead.
__global__ void __launch_bounds__(256)
camera_block_finish_kernel(const double* __restrict__ chunk_partial,
                           const int* __restrict__ cam_chunk_start, int C,
                           double* __restrict__ upper /* C x 45 */) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // one thread per (camera, entry)
  if (i >= C * 45) return;
  const int c = i / 45, k = i - 45 * c;
  double s = 0.0;
  for (int ch = cam_chunk_start[c]; ch < cam_chunk_start[c + 1]; ++ch) s += chunk_partial[(size_t)ch * 45 + k];
  upper[i] = s;
}

__global__ void __launch_bounds__(64)
camera_block_invert_kernel(const double* __restrict__ upper, const double* __restrict__ Df,
                           const double* __restrict__ dup_blocks, int C, int invert,
                           double* __restrict__ blocks /* C x 81 */) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double m[81];
  int q = 0;
#pragma unroll
  for (int i = 0; i < 9; ++i)
#pragma unroll
    for (int j = i; j < 9; ++j) { const double v = upper[(size_t)c * 45 + q]; ++q; m[i * 9 + j] = v; m[j * 9 + i] = v; }
  if (Df != nullptr) {
#pragma unroll
    for (int i = 0; i < 9; ++i) { const double d = Df[9 * (size_t)c + i]; m[i * 10] += d * d; }
  }
  if (dup_blocks != nullptr) {
#pragma unroll
    for (int k = 0; k < 81; ++k) m[k] += dup_blocks[(size_t)c * 81 + k];
  }
  if (invert) LltInverseUpper<9>(m);
#pragma unroll
  for (int k = 0; k < 81; ++k) blocks[(size_t)c * 81 + k] = m[k];
}

// y (+)= B x for a block-diagonal matrix of C row-major 9x9 blocks: one thread per
// output element (BlockRandomAccessDiagonalMatrix::RightMultiplyAndAccumulate,
// block_random_access_diagonal_matrix.cc:102-116; BlockSparseMatrix::
// RightMultiplyAndAccumulate on the F'F inverse, block_sparse_matrix.cc:239-274).
__global__ void block9_multiply_kernel(const double* __restrict__ blocks, const double* __restrict__ x,
                                       double* __restrict__ y, int n, int accumulate) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int c = i / 9;
  const double* row = blocks + (size_t)i * 9;
  const double* xc = x + 9 * c;
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 9; ++k) s += row[k] * xc[k];
  y[i] = accumulate ? y[i] + s : s;
}

// ----------------------------------------------------------------------------
// PartitionedMatrixView<2,3,9> products as stand-alone kernels (unit seam; the
// solver itself uses the fused point kernels).
// ----------------------------------------------------------------------------
// y_rows += E x_e   (RightMultiplyAndAccumulateE, partitioned_matrix_view_impl.h:112-137)
// y_e    += E' x_rows (LeftMultiplyAndAccumulateE, :193-264); one thread per point.
template <bool LEFT>
__global__ void pmv_e_kernel(const double* __restrict__ E, const int* __restrict__ row_start, int P,
                             const double* __restrict__ x, double* __restrict__ y) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  double a0 = 0.0, a1 = 0.0, a2 = 0.0;
  double x0 = 0.0, x1 = 0.0, x2 = 0.0;
  if (!LEFT) { x0 = x[3 * (size_t)p]; x1 = x[3 * (size_t)p + 1]; x2 = x[3 * (size_t)p + 2]; }
  for (int r = row_start[p]; r < row_start[p + 1]; ++r) {
    const double* e = E + (size_t)r * 6;
    if (LEFT) {
      const double t0 = x[2 * (size_t)r], t1 = x[2 * (size_t)r + 1];
      a0 += e[0] * t0 + e[3] * t1; a1 += e[1] * t0 + e[4] * t1; a2 += e[2] * t0 + e[5] * t1;
    } else {
      y[2 * (size_t)r] += e[0] * x0 + e[1] * x1 + e[2] * x2;
      y[2 * (size_t)r + 1] += e[3] * x0 + e[4] * x1 + e[5] * x2;
    }
  }
  if (LEFT) { y[3 * (size_t)p] += a0; y[3 * (size_t)p + 1] += a1; y[3 * (size_t)p + 2] += a2; }
}

// y_rows += F x_f (RightMultiplyAndAccumulateF, :139-191); y_f += F' x_rows
// (LeftMultiplyAndAccumulateF, :266-375). One thread per row.
template <bool LEFT>
__global__ void pmv_f_kernel(const double* __restrict__ F, const int* __restrict__ cam, int N,
                             const double* __restrict__ x, double* __restrict__ y) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (LEFT) {
    // The nine sums of a row leave through shared memory so that consecutive lanes add to consecutive
    // doubles of a camera (3 sectors per row per reduction pass instead of 9 scattered ones).
    __shared__ double sT[4][32 * 9];  // blockDim.x == 128
    __shared__ int sC[4][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool on = r < N;
    if (on) {
      const double2* f2 = reinterpret_cast<const double2*>(F + (size_t)r * 18);
      double f[18];
#pragma unroll
      for (int k = 0; k < 9; ++k) { const double2 v = f2[k]; f[2 * k] = v.x; f[2 * k + 1] = v.y; }
      const double t0 = x[2 * (size_t)r], t1 = x[2 * (size_t)r + 1];
#pragma unroll
      for (int k = 0; k < 9; ++k) sT[warp][lane * 9 + k] = f[k] * t0 + f[9 + k] * t1;
      sC[warp][lane] = cam[r];
    }
    __syncwarp();
    const int rows = min(32, N - (r - lane));  // rows of this warp that exist
#pragma unroll
    for (int k = 0; k < 9; ++k) {
      const int idx = k * 32 + lane;
      const int o = idx / 9;
      if (o < rows) atomicAdd(y + 9 * (size_t)sC[warp][o] + (idx - 9 * o), sT[warp][idx]);
    }
    return;
  }
  if (r >= N) return;
  const double2* f2 = reinterpret_cast<const double2*>(F + (size_t)r * 18);
  double f[18];
#pragma unroll
  for (int k = 0; k < 9; ++k) { const double2 v = f2[k]; f[2 * k] = v.x; f[2 * k + 1] = v.y; }
  const int c = cam[r];
  double s0 = 0.0, s1 = 0.0;
#pragma unroll
  for (int k = 0; k < 9; ++k) { const double xk = x[9 * (size_t)c + k]; s0 += f[k] * xk; s1 += f[9 + k] * xk; }
  y[2 * (size_t)r] += s0;
  y[2 * (size_t)r + 1] += s1;
}

// Gather Jacobian blocks from an arbitrary value layout into E / F
// (cell.position need not be the evaluator's E-first layout:
// fake_bundle_adjustment_jacobian.cc:82-88, evaluation_benchmark.cc:123-136).
__global__ void gather_blocks_kernel(const double* __restrict__ values, const int* __restrict__ e_pos,
                                     const int* __restrict__ f_pos, int N, double* __restrict__ E,
                                     double* __restrict__ F) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 24LL * N) return;
  const int r = (int)(i / 24);
  const int k = (int)(i - 24LL * r);
  if (k < 6) E[(size_t)r * 6 + k] = values[(size_t)e_pos[r] + k];
  else F[(size_t)r * 18 + (k - 6)] = values[(size_t)f_pos[r] + (k - 6)];
}

}  // namespace cb200
#endif  // CB200_CAMERA_KERNELS_CUH_
