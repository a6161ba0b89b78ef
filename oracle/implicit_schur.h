// ORACLE — TEST INFRASTRUCTURE ONLY (see small_blas.h header).
//
// CPU restatement of ImplicitSchurComplement:
//   internal/ceres/implicit_schur_complement.cc:49-97    Init
//   internal/ceres/implicit_schur_complement.cc:106-144  RightMultiplyAndAccumulate (S x)
//   internal/ceres/implicit_schur_complement.cc:146-174  InversePowerSeriesOperatorRightMultiplyAccumulate
//   internal/ceres/implicit_schur_complement.cc:179-204  AddDiagonalAndInvert
//   internal/ceres/implicit_schur_complement.cc:208-243  BackSubstitute
//   internal/ceres/implicit_schur_complement.cc:251-276  UpdateRhs
#ifndef ORACLE_IMPLICIT_SCHUR_H_
#define ORACLE_IMPLICIT_SCHUR_H_

#include <memory>
#include <vector>

#include "partitioned_view.h"

namespace oracle {

enum PreconditionerType {
  IDENTITY = 0,
  JACOBI = 1,
  SCHUR_JACOBI = 2,
  SCHUR_POWER_SERIES_EXPANSION = 3,  // values of include/ceres/types.h:93-141
};

struct SolverOptions {
  int num_eliminate_blocks = 0;
  int num_threads = 1;
  int preconditioner_type = SCHUR_JACOBI;
  int min_num_iterations = 0;
  int max_num_iterations = 500;
  int residual_reset_period = 10;
  bool use_spse_initialization = false;
  int max_num_spse_iterations = 5;
  double spse_tolerance = 0.1;
};

// ParallelSetZero / ParallelAssign (parallel_vector_ops.h:49-87): parallel only
// for vectors of at least 2 * kMinBlockSizeParallelVectorOps elements.
constexpr int kMinBlockSizeParallelVectorOps = 1 << 16;
inline void ParallelSetZero(int nt, double* v, int n) {
  ParallelForRange(0, n, nt, [&](int, int s, int e) {
    std::memset(v + s, 0, sizeof(double) * (e - s));
  }, kMinBlockSizeParallelVectorOps);
}
inline void ParallelSetZero(int nt, std::vector<double>& v) {
  ParallelSetZero(nt, v.data(), (int)v.size());
}
template <typename F>  // f(i) -> value assigned to element i
inline void ParallelAssign(int nt, double* dst, int n, F f) {
  ParallelForRange(0, n, nt, [&](int, int s, int e) {
    for (int i = s; i < e; ++i) dst[i] = f(i);
  }, kMinBlockSizeParallelVectorOps);
}

class ImplicitSchurComplement {
 public:
  explicit ImplicitSchurComplement(const SolverOptions& o) : options_(o) {}

  void Init(const BlockStructure* bs, const BlockStructure* tbs,
            const double* values, const double* D, const double* b) {
    if (!A_) {
      A_ = std::make_unique<PartitionedMatrixView>(
          bs, tbs, values, options_.num_eliminate_blocks, options_.num_threads);
    }
    A_->set_values(values);
    D_ = D;
    b_ = b;
    compute_ftf_inverse_ =
        options_.use_spse_initialization ||
        options_.preconditioner_type == JACOBI ||
        options_.preconditioner_type == SCHUR_POWER_SERIES_EXPANSION;
    if (!have_ete_) {
      ete_inv_ = A_->CreateBlockDiagonalEtE();
      if (compute_ftf_inverse_) ftf_inv_ = A_->CreateBlockDiagonalFtF();
      rhs_.assign(A_->num_cols_f(), 0.0);
      tmp_rows_.assign(A_->num_rows(), 0.0);
      tmp_e_cols_.assign(A_->num_cols_e(), 0.0);
      tmp_e_cols_2_.assign(A_->num_cols_e(), 0.0);
      tmp_f_cols_.assign(A_->num_cols_f(), 0.0);
      have_ete_ = true;
    } else {
      A_->UpdateBlockDiagonalEtE(&ete_inv_);
      if (compute_ftf_inverse_) A_->UpdateBlockDiagonalFtF(&ftf_inv_);
    }
    AddDiagonalAndInvert(D_, &ete_inv_);
    if (compute_ftf_inverse_)
      AddDiagonalAndInvert(D_ == nullptr ? nullptr : D_ + A_->num_cols_e(),
                           &ftf_inv_);
    UpdateRhs();
  }

  int num_rows() const { return A_->num_cols_f(); }
  int num_cols() const { return A_->num_cols_f(); }
  const std::vector<double>& rhs() const { return rhs_; }
  const BlockDiagonal& block_diagonal_EtE_inverse() const { return ete_inv_; }
  const BlockDiagonal& block_diagonal_FtF_inverse() const { return ftf_inv_; }
  const PartitionedMatrixView& view() const { return *A_; }

  // y = D_f^2 x + F'(F x - E (E'E)^-1 E' F x). y is OVERWRITTEN (:130-140).
  void RightMultiplyAndAccumulate(const double* x, double* y) const {
    const int nt = options_.num_threads;
    ParallelSetZero(nt, tmp_rows_);
    A_->RightMultiplyAndAccumulateF(x, tmp_rows_.data());
    ParallelSetZero(nt, tmp_e_cols_);
    A_->LeftMultiplyAndAccumulateE(tmp_rows_.data(), tmp_e_cols_.data());
    ParallelSetZero(nt, tmp_e_cols_2_);
    ete_inv_.RightMultiplyAndAccumulate(tmp_e_cols_.data(), tmp_e_cols_2_.data(), nt);
    { double* t = tmp_e_cols_2_.data();
      ParallelAssign(nt, t, (int)tmp_e_cols_2_.size(), [t](int i) { return -t[i]; }); }
    A_->RightMultiplyAndAccumulateE(tmp_e_cols_2_.data(), tmp_rows_.data());
    const int n = num_cols();
    if (D_ != nullptr) {
      const double* Df = D_ + A_->num_cols_e();
      ParallelAssign(nt, y, n, [Df, x](int i) { return (Df[i] * Df[i]) * x[i]; });
    } else {
      ParallelSetZero(nt, y, n);
    }
    A_->LeftMultiplyAndAccumulateF(tmp_rows_.data(), y);
  }

  // y += (F'F)^-1 F'E (E'E)^-1 E'F x
  void InversePowerSeriesOperatorRightMultiplyAccumulate(const double* x,
                                                         double* y) const {
    const int nt = options_.num_threads;
    ParallelSetZero(nt, tmp_rows_);
    A_->RightMultiplyAndAccumulateF(x, tmp_rows_.data());
    ParallelSetZero(nt, tmp_e_cols_);
    A_->LeftMultiplyAndAccumulateE(tmp_rows_.data(), tmp_e_cols_.data());
    ParallelSetZero(nt, tmp_e_cols_2_);
    ete_inv_.RightMultiplyAndAccumulate(tmp_e_cols_.data(), tmp_e_cols_2_.data(), nt);
    ParallelSetZero(nt, tmp_rows_);
    A_->RightMultiplyAndAccumulateE(tmp_e_cols_2_.data(), tmp_rows_.data());
    ParallelSetZero(nt, tmp_f_cols_);
    A_->LeftMultiplyAndAccumulateF(tmp_rows_.data(), tmp_f_cols_.data());
    ftf_inv_.RightMultiplyAndAccumulate(tmp_f_cols_.data(), y, nt);
  }

  // y = [ (E'E)^-1 E'(b - F x) ; x ]
  void BackSubstitute(const double* x, double* y) {
    const int nt = options_.num_threads;
    const int nce = A_->num_cols_e(), ncf = A_->num_cols_f();
    const int nr = A_->num_rows();
    ParallelSetZero(nt, tmp_rows_);
    if (ncf > 0) A_->RightMultiplyAndAccumulateF(x, tmp_rows_.data());
    { double* t = tmp_rows_.data(); const double* b = b_;
      ParallelAssign(nt, t, nr, [t, b](int i) { return b[i] - t[i]; }); }
    ParallelSetZero(nt, tmp_e_cols_);
    A_->LeftMultiplyAndAccumulateE(tmp_rows_.data(), tmp_e_cols_.data());
    ParallelSetZero(nt, y, nce + ncf);
    ete_inv_.RightMultiplyAndAccumulate(tmp_e_cols_.data(), y, nt);
    for (int i = 0; i < ncf; ++i) y[nce + i] = x[i];
  }

 private:
  void AddDiagonalAndInvert(const double* D, BlockDiagonal* bd) {
    ParallelForRange(0, (int)bd->size.size(), options_.num_threads,
                     [&](int, int s, int e) {
      for (int i = s; i < e; ++i) {
        const int n = bd->size[i];
        double* m = &bd->values[bd->cell_pos[i]];
        if (D != nullptr) {
          const double* d = D + bd->block_pos[i];
          for (int k = 0; k < n; ++k) m[k * n + k] += d[k] * d[k];
        }
        LltInverseUpper(m, n);
      }
    });
  }

  // rhs = F'(b - E (E'E)^-1 E'b)
  void UpdateRhs() {
    const int nt = options_.num_threads;
    ParallelSetZero(nt, tmp_e_cols_);
    A_->LeftMultiplyAndAccumulateE(b_, tmp_e_cols_.data());
    ParallelSetZero(nt, tmp_e_cols_2_);
    ete_inv_.RightMultiplyAndAccumulate(tmp_e_cols_.data(), tmp_e_cols_2_.data(), nt);
    ParallelSetZero(nt, tmp_rows_);
    A_->RightMultiplyAndAccumulateE(tmp_e_cols_2_.data(), tmp_rows_.data());
    const int nr = A_->num_rows();
    { double* t = tmp_rows_.data(); const double* b = b_;
      ParallelAssign(nt, t, nr, [t, b](int i) { return b[i] - t[i]; }); }
    ParallelSetZero(nt, rhs_);
    A_->LeftMultiplyAndAccumulateF(tmp_rows_.data(), rhs_.data());
  }

  SolverOptions options_;
  std::unique_ptr<PartitionedMatrixView> A_;
  const double* D_ = nullptr;
  const double* b_ = nullptr;
  bool compute_ftf_inverse_ = false;
  bool have_ete_ = false;
  BlockDiagonal ete_inv_, ftf_inv_;
  std::vector<double> rhs_;
  mutable std::vector<double> tmp_rows_, tmp_e_cols_, tmp_e_cols_2_, tmp_f_cols_;
};

}  // namespace oracle

#endif  // ORACLE_IMPLICIT_SCHUR_H_
