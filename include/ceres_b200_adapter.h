// ceres_b200_adapter.h — the reference-side binding of libceres_b200_schur.so.
//
// Drop this header into internal/ceres/ of a Ceres checkout, link
// libceres_b200_schur.so, and return the class below from
// LinearSolver::Create, case ITERATIVE_SCHUR (internal/ceres/linear_solver.cc:111-116).
// Solver::Options{linear_solver_type = ITERATIVE_SCHUR,
// preconditioner_type = SCHUR_JACOBI} then runs on the B200 unchanged.
//
// It replaces IterativeSchurComplementSolver
// (internal/ceres/iterative_schur_complement_solver.h:72-96, .cc:64-199) and has
// the same contract: one instance per Solve(), reused for every LM iteration
// (linear_solver.h:137-142); x is left untouched on FAILURE / FATAL_ERROR.
//
// The flattening helper is a template on the block-structure type so that it
// also compiles (and is tested, tests/test_cpu_host.py) against a stand-in for
// ceres/block_structure.h outside a Ceres tree.
#ifndef CERES_B200_ADAPTER_H_
#define CERES_B200_ADAPTER_H_

#include <string>
#include <vector>

#include "ceres_b200_schur.h"

namespace ceres_b200 {

// CompressedRowBlockStructure (block_structure.h:52-96) -> the flat int arrays
// of cb200_structure. `Bs` needs .rows[i].{block.{size,position}, cells[j].{block_id,position}}
// and .cols[j].{size,position}.
struct FlatBlockStructure {
  std::vector<int> row_block_size, row_block_pos, cell_ptr, cell_block_id, cell_position;
  std::vector<int> col_block_size, col_block_pos;
  cb200_structure view() const {
    cb200_structure s;
    s.num_row_blocks = static_cast<int>(row_block_size.size());
    s.row_block_size = row_block_size.data();
    s.row_block_pos = row_block_pos.data();
    s.cell_ptr = cell_ptr.data();
    s.cell_block_id = cell_block_id.data();
    s.cell_position = cell_position.data();
    s.num_col_blocks = static_cast<int>(col_block_size.size());
    s.col_block_size = col_block_size.data();
    s.col_block_pos = col_block_pos.data();
    return s;
  }
};

template <typename Bs>
FlatBlockStructure FlattenBlockStructure(const Bs& bs) {
  FlatBlockStructure f;
  f.cell_ptr.push_back(0);
  for (const auto& row : bs.rows) {
    f.row_block_size.push_back(row.block.size);
    f.row_block_pos.push_back(row.block.position);
    for (const auto& cell : row.cells) {
      f.cell_block_id.push_back(cell.block_id);
      f.cell_position.push_back(cell.position);
    }
    f.cell_ptr.push_back(static_cast<int>(f.cell_block_id.size()));
  }
  for (const auto& col : bs.cols) {
    f.col_block_size.push_back(col.size);
    f.col_block_pos.push_back(col.position);
  }
  return f;
}

}  // namespace ceres_b200

#ifdef CERES_INTERNAL_LINEAR_SOLVER_H_  // only inside a Ceres tree, after "ceres/linear_solver.h"

#include "ceres/block_sparse_matrix.h"
#include "ceres/block_structure.h"
#include "glog/logging.h"

namespace ceres::internal {

class B200IterativeSchurComplementSolver final : public BlockSparseMatrixSolver {
 public:
  explicit B200IterativeSchurComplementSolver(LinearSolver::Options options)
      : options_(std::move(options)) {}
  ~B200IterativeSchurComplementSolver() override {
    if (registered_values_ != nullptr) cb200_host_unregister(registered_values_);
    cb200_destroy(handle_);
  }

 private:
  static int PreconditionerCode(PreconditionerType t) {
    switch (t) {
      case IDENTITY: return CB200_IDENTITY;
      case JACOBI: return CB200_JACOBI;
      case SCHUR_JACOBI: return CB200_SCHUR_JACOBI;
      case SCHUR_POWER_SERIES_EXPANSION: return CB200_SCHUR_POWER_SERIES_EXPANSION;
      default: return -1;  // CLUSTER_*: rejected by cb200_create, no CPU fallback
    }
  }

  LinearSolver::Summary SolveImpl(BlockSparseMatrix* A, const double* b,
                                  const LinearSolver::PerSolveOptions& per_solve_options,
                                  double* x) final {
    CHECK(A->block_structure() != nullptr);
    LinearSolver::Summary summary;
    if (handle_ == nullptr) {  // same sparsity structure for every later call
      const ceres_b200::FlatBlockStructure flat = ceres_b200::FlattenBlockStructure(*A->block_structure());
      const cb200_structure view = flat.view();
      cb200_options o{};
      o.num_eliminate_blocks = options_.elimination_groups[0];
      o.preconditioner_type = PreconditionerCode(options_.preconditioner_type);
      o.min_num_iterations = options_.min_num_iterations;
      o.max_num_iterations = options_.max_num_iterations;
      o.residual_reset_period = options_.residual_reset_period;
      o.use_spse_initialization = options_.use_spse_initialization;
      o.max_num_spse_iterations = options_.max_num_spse_iterations;
      o.spse_tolerance = options_.spse_tolerance;
      // linear_solver.cc:111-113: the same adapter serves ITERATIVE_SCHUR with and without the
      // explicit reduced camera matrix (SparseSchurComplementSolver's CG branch)
      o.use_explicit_schur_complement = options_.use_explicit_schur_complement;
      o.device_id = 0;
      o.rank = 0;
      o.world_size = 1;
      if (cb200_create(&view, &o, &handle_) != CB200_OK) {
        summary.termination_type = LinearSolverTerminationType::FATAL_ERROR;
        summary.message = cb200_last_error();
        return summary;
      }
    }
    if (registered_values_ != A->values()) {
      // the Jacobian's value array keeps its address for the whole Solve(): page-lock it once, so that the
      // 24 N doubles cross PCIe at full rate every iteration (best effort: the solve works without it)
      if (registered_values_ != nullptr) cb200_host_unregister(registered_values_);
      registered_values_ = nullptr;
      if (cb200_host_register(A->values(), sizeof(double) * static_cast<size_t>(A->num_nonzeros())) == CB200_OK)
        registered_values_ = A->values();
    }
    cb200_summary s{};
    const int rc = cb200_solve(handle_, A->values(), b, per_solve_options.D,
                               per_solve_options.q_tolerance, per_solve_options.r_tolerance, x, &s);
    if (rc != CB200_OK) {
      summary.termination_type = LinearSolverTerminationType::FATAL_ERROR;
      summary.message = cb200_last_error();
      return summary;
    }
    summary.num_iterations = s.num_iterations;
    summary.termination_type = static_cast<LinearSolverTerminationType>(s.termination_type);
    summary.message = s.message;
    return summary;
  }

  LinearSolver::Options options_;
  cb200_solver* handle_ = nullptr;
  const double* registered_values_ = nullptr;
};

}  // namespace ceres::internal

#endif  // CERES_INTERNAL_LINEAR_SOLVER_H_
#endif  // CERES_B200_ADAPTER_H_
