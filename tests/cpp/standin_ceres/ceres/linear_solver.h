// Stand-in (test only) for the parts of internal/ceres/linear_solver.h the adapter touches:
// termination types (:57-74), LinearSolver::Options (:150-230), PerSolveOptions (:237-300),
// Summary (:320-326), TypedLinearSolver / BlockSparseMatrixSolver (:363-395).
#ifndef CERES_INTERNAL_LINEAR_SOLVER_H_
#define CERES_INTERNAL_LINEAR_SOLVER_H_
#include <string>
#include <vector>
#include "ceres/block_sparse_matrix.h"
namespace ceres {
enum PreconditionerType { IDENTITY, JACOBI, SCHUR_JACOBI, SCHUR_POWER_SERIES_EXPANSION, CLUSTER_JACOBI,
                          CLUSTER_TRIDIAGONAL, SUBSET };
}
namespace ceres::internal {
enum class LinearSolverTerminationType { SUCCESS, NO_CONVERGENCE, FAILURE, FATAL_ERROR };
class LinearSolver {
 public:
  struct Options {
    PreconditionerType preconditioner_type = JACOBI;
    int min_num_iterations = 1;
    int max_num_iterations = 1;
    int residual_reset_period = 10;
    bool use_explicit_schur_complement = false;
    bool use_spse_initialization = false;
    int max_num_spse_iterations = 5;
    double spse_tolerance = 0.1;
    std::vector<int> elimination_groups;
  };
  struct PerSolveOptions {
    double* D = nullptr;
    double r_tolerance = 0.0;
    double q_tolerance = 0.0;
  };
  struct Summary {
    double residual_norm = -1.0;
    int num_iterations = -1;
    LinearSolverTerminationType termination_type = LinearSolverTerminationType::FAILURE;
    std::string message;
  };
  virtual ~LinearSolver() = default;
  virtual Summary Solve(LinearOperator* A, const double* b, const PerSolveOptions& o, double* x) = 0;
};
template <typename MatrixType>
class TypedLinearSolver : public LinearSolver {
 public:
  LinearSolver::Summary Solve(LinearOperator* A, const double* b, const LinearSolver::PerSolveOptions& o,
                              double* x) override {
    return SolveImpl(static_cast<MatrixType*>(A), b, o, x);
  }
 private:
  virtual LinearSolver::Summary SolveImpl(MatrixType* A, const double* b,
                                          const LinearSolver::PerSolveOptions& o, double* x) = 0;
};
using BlockSparseMatrixSolver = TypedLinearSolver<BlockSparseMatrix>;
}  // namespace ceres::internal
#endif  // CERES_INTERNAL_LINEAR_SOLVER_H_
