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

#include <cstdlib>
#include <memory>

#include "absl/log/check.h"  // as linear_solver.h:43 (this tree logs through abseil, not glog)
#include "ceres/block_sparse_matrix.h"
#include "ceres/block_structure.h"

namespace ceres::internal {

class B200IterativeSchurComplementSolver final : public BlockSparseMatrixSolver {
 public:
  // `fallback`: the solver LinearSolver::Create returned for this configuration before the patch
  // (IterativeSchurComplementSolver, or SparseSchurComplementSolver with use_explicit_schur_complement;
  // linear_solver.cc:111-116). Problems OUTSIDE the device path's scope -- block structures that are not
  // <2,3,9> bundle adjustment (camera priors or other F-only rows, other block sizes, no F blocks at all:
  // iterative_schur_complement_solver.cc:85-93), CLUSTER_* preconditioners -- are handed to it unchanged, so
  // the patch never turns a Solve() that used to work into a failure. Device errors are never delegated:
  // they are FATAL_ERROR. Without a fallback, out-of-scope problems are FATAL_ERROR too.
  explicit B200IterativeSchurComplementSolver(LinearSolver::Options options,
                                              std::unique_ptr<LinearSolver> fallback = nullptr)
      : options_(std::move(options)), fallback_(std::move(fallback)) {}
  ~B200IterativeSchurComplementSolver() override {
    if (registered_values_ != nullptr) cb200_host_unregister(registered_values_);
    cb200_destroy(handle_);
  }

  // GPUs one handle drives (cb200_options::num_gpus / device_ids): LinearSolver::Options has no field for
  // device placement, so it comes from the environment of the process that calls ceres::Solve:
  //   CERES_B200_NUM_GPUS=4            GPUs 0..3
  //   CERES_B200_DEVICES=2,3,6,7       these CUDA ordinals (implies their count)
  //   CERES_B200_DETERMINISTIC=1       bitwise-repeatable solves (fixed-order camera sums)
  int num_gpus() const { return num_gpus_; }

 private:
  static int PreconditionerCode(PreconditionerType t) {
    switch (t) {
      case IDENTITY: return CB200_IDENTITY;
      case JACOBI: return CB200_JACOBI;
      case SCHUR_JACOBI: return CB200_SCHUR_JACOBI;
      case SCHUR_POWER_SERIES_EXPANSION: return CB200_SCHUR_POWER_SERIES_EXPANSION;
      default: return -1;  // CLUSTER_*: outside the device path
    }
  }

  static void PlacementFromEnvironment(cb200_options* o) {
    if (const char* e = std::getenv("CERES_B200_DEVICES")) {
      int n = 0;
      for (const char* p = e; *p != '\0' && n < 16;) {
        char* end = nullptr;
        const long v = std::strtol(p, &end, 10);
        if (end == p) break;
        o->device_ids[n++] = static_cast<int>(v);
        p = (*end == ',') ? end + 1 : end;
      }
      o->num_gpus = n;
      if (n == 1) { o->device_id = o->device_ids[0]; o->num_gpus = 0; }
    } else if (const char* e = std::getenv("CERES_B200_NUM_GPUS")) {
      o->num_gpus = std::atoi(e);
    }
    if (const char* e = std::getenv("CERES_B200_DETERMINISTIC")) o->deterministic = std::atoi(e) != 0;
  }

  LinearSolver::Summary SolveImpl(BlockSparseMatrix* A, const double* b,
                                  const LinearSolver::PerSolveOptions& per_solve_options,
                                  double* x) final {
    CHECK(A->block_structure() != nullptr);
    LinearSolver::Summary summary;
    if (delegate_) return fallback_->Solve(A, b, per_solve_options, x);
    if (handle_ == nullptr) {  // same sparsity structure for every later call (linear_solver.h:137-142)
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
      PlacementFromEnvironment(&o);
      const int rc = cb200_create(&view, &o, &handle_);
      if ((rc == CB200_ERR_UNSUPPORTED_STRUCTURE || rc == CB200_ERR_INVALID_ARGUMENT) && fallback_ != nullptr) {
        delegate_ = true;  // out of the device path's scope: the reference's own solver, as before the patch
        return fallback_->Solve(A, b, per_solve_options, x);
      }
      if (rc != CB200_OK) {
        summary.termination_type = LinearSolverTerminationType::FATAL_ERROR;
        summary.message = cb200_last_error();
        return summary;
      }
      num_gpus_ = o.num_gpus > 1 ? o.num_gpus : 1;
    }
    if (registered_values_ != A->values()) {
      // the Jacobian's value array keeps its address for the whole Solve(): page-lock it once, so that the
      // 24 N doubles cross PCIe at full rate every iteration (best effort: the solve works without it; the
      // one-off cost is that of cudaHostRegister). A BlockSparseMatrix created with
      // use_page_locked_memory (block_sparse_matrix.cc:178-180) is already pinned: registering is then a no-op.
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
  std::unique_ptr<LinearSolver> fallback_;
  bool delegate_ = false;
  cb200_solver* handle_ = nullptr;
  const double* registered_values_ = nullptr;
  int num_gpus_ = 1;
};

}  // namespace ceres::internal

#endif  // CERES_INTERNAL_LINEAR_SOLVER_H_
#endif  // CERES_B200_ADAPTER_H_
