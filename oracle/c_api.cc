// ORACLE — TEST INFRASTRUCTURE ONLY (see small_blas.h header).
//
// extern "C" surface of the CPU oracle, bound with ctypes by tests/, smoke() and
// bench.py's cpu_baseline / --impl reference legs. Takes the reference's
// CompressedRowBlockStructure (internal/ceres/block_structure.h:52-96) flattened
// into plain int arrays, so that the KAT problems with scalar / mixed blocks
// (internal/ceres/linear_least_squares_problems.cc:301-940) and <2,3,9> BAL
// Jacobians go through the same code.
#include <chrono>
#include <cstring>

#include "explicit_schur.h"
#include "solver.h"

using namespace oracle;

struct OracleHandle {
  BlockStructure bs;
  BlockStructure tbs;
  SolverOptions options;
  std::unique_ptr<IterativeSchurComplementSolver> solver;
  std::unique_ptr<ImplicitSchurComplement> isc;  // stand-alone operator tests
  std::unique_ptr<SchurJacobiPreconditioner> sj;
  std::unique_ptr<ExplicitSchurComplementSolver> explicit_solver;  // use_explicit_schur_complement
  std::string message;
};

extern "C" {

struct OracleSummary {
  double residual_norm;
  int num_iterations;
  int termination_type;
  char message[512];
};

OracleHandle* oracle_create(int num_row_blocks, const int* row_size,
                            const int* row_pos, const int* cell_ptr,
                            const int* cell_block_id, const int* cell_position,
                            int num_col_blocks, const int* col_size,
                            const int* col_pos, int num_eliminate_blocks,
                            int num_threads, int preconditioner_type,
                            int min_num_iterations, int max_num_iterations,
                            int residual_reset_period,
                            int use_spse_initialization,
                            int max_num_spse_iterations,
                            double spse_tolerance) {
  auto* h = new OracleHandle;
  h->bs.cols.resize(num_col_blocks);
  for (int i = 0; i < num_col_blocks; ++i)
    h->bs.cols[i] = Block{col_size[i], col_pos[i]};
  h->bs.rows.resize(num_row_blocks);
  int cum = 0;
  for (int r = 0; r < num_row_blocks; ++r) {
    CompressedRow& row = h->bs.rows[r];
    row.block = Block{row_size[r], row_pos[r]};
    row.nnz = 0;
    for (int c = cell_ptr[r]; c < cell_ptr[r + 1]; ++c) {
      row.cells.push_back(Cell{cell_block_id[c], cell_position[c]});
      row.nnz += row_size[r] * col_size[cell_block_id[c]];
    }
    cum += row.nnz;
    row.cumulative_nnz = cum;
  }
  h->tbs = CreateTranspose(h->bs);
  SolverOptions& o = h->options;
  o.num_eliminate_blocks = num_eliminate_blocks;
  o.num_threads = num_threads;
  o.preconditioner_type = preconditioner_type;
  o.min_num_iterations = min_num_iterations;
  o.max_num_iterations = max_num_iterations;
  o.residual_reset_period = residual_reset_period;
  o.use_spse_initialization = use_spse_initialization != 0;
  o.max_num_spse_iterations = max_num_spse_iterations;
  o.spse_tolerance = spse_tolerance;
  h->solver = std::make_unique<IterativeSchurComplementSolver>(&h->bs, o);
  return h;
}

void oracle_destroy(OracleHandle* h) { delete h; }

// op: 0 RightE (y_rows += E x_e), 1 RightF, 2 LeftE (y_e += E' x_rows), 3 LeftF.
void oracle_pmv(OracleHandle* h, const double* values, int op, const double* x,
                double* y) {
  PartitionedMatrixView v(&h->bs, &h->tbs, values,
                          h->options.num_eliminate_blocks,
                          h->options.num_threads);
  switch (op) {
    case 0: v.RightMultiplyAndAccumulateE(x, y); break;
    case 1: v.RightMultiplyAndAccumulateF(x, y); break;
    case 2: v.LeftMultiplyAndAccumulateE(x, y); break;
    case 3: v.LeftMultiplyAndAccumulateF(x, y); break;
  }
}

// which: 0 -> block diagonal of E'E, 1 -> of F'F. out holds the contiguous
// row-major blocks. Returns the number of doubles written.
int oracle_block_diagonal(OracleHandle* h, const double* values, int which,
                          double* out) {
  PartitionedMatrixView v(&h->bs, &h->tbs, values,
                          h->options.num_eliminate_blocks,
                          h->options.num_threads);
  BlockDiagonal d = which == 0 ? v.CreateBlockDiagonalEtE()
                               : v.CreateBlockDiagonalFtF();
  if (out) std::memcpy(out, d.values.data(), sizeof(double) * d.values.size());
  return (int)d.values.size();
}

// ImplicitSchurComplement::Init. values/D/b must outlive later isc calls.
void oracle_isc_init(OracleHandle* h, const double* values, const double* D,
                     const double* b) {
  if (!h->isc) h->isc = std::make_unique<ImplicitSchurComplement>(h->options);
  h->isc->Init(&h->bs, &h->tbs, values, D, b);
}
int oracle_isc_num_cols(OracleHandle* h) { return h->isc->num_cols(); }
void oracle_isc_rhs(OracleHandle* h, double* out) {
  std::memcpy(out, h->isc->rhs().data(), sizeof(double) * h->isc->rhs().size());
}
void oracle_isc_multiply(OracleHandle* h, const double* x, double* y) {
  h->isc->RightMultiplyAndAccumulate(x, y);
}
void oracle_isc_pse_operator(OracleHandle* h, const double* x, double* y) {
  h->isc->InversePowerSeriesOperatorRightMultiplyAccumulate(x, y);
}
void oracle_isc_back_substitute(OracleHandle* h, const double* z, double* y) {
  h->isc->BackSubstitute(z, y);
}
int oracle_isc_ete_inverse(OracleHandle* h, double* out) {
  const auto& v = h->isc->block_diagonal_EtE_inverse().values;
  if (out) std::memcpy(out, v.data(), sizeof(double) * v.size());
  return (int)v.size();
}
int oracle_isc_ftf_inverse(OracleHandle* h, double* out) {
  const auto& v = h->isc->block_diagonal_FtF_inverse().values;
  if (out) std::memcpy(out, v.data(), sizeof(double) * v.size());
  return (int)v.size();
}
void oracle_pse_apply(OracleHandle* h, int max_num_spse_iterations,
                      double spse_tolerance, const double* x, double* y) {
  PowerSeriesApply(*h->isc, max_num_spse_iterations, spse_tolerance,
                   h->options.num_threads, x, y);
}

// SchurJacobiPreconditioner. invert != 0: Update (eliminate + invert);
// invert == 0: block-diag(S) before inversion. Returns doubles written.
int oracle_schur_jacobi(OracleHandle* h, const double* values, const double* D,
                        int invert, double* out) {
  if (!h->sj)
    h->sj = std::make_unique<SchurJacobiPreconditioner>(
        &h->bs, h->options.num_eliminate_blocks, h->options.num_threads);
  if (invert) h->sj->Update(values, D); else h->sj->Eliminate(values, D);
  const auto& v = h->sj->values();
  if (out) std::memcpy(out, v.data(), sizeof(double) * v.size());
  return (int)v.size();
}

// IterativeSchurComplementSolver::SolveImpl. D may be NULL. If trace != NULL it
// receives the reduced solution after each of the first max_trace PCG
// iterations (max_trace x num_cols_f doubles, row-major).
void oracle_solve(OracleHandle* h, const double* values, const double* b,
                  const double* D, double q_tolerance, double r_tolerance,
                  double* x, OracleSummary* out, double* trace, int max_trace) {
  std::function<void(int, const Vec&)> tr = nullptr;
  if (trace != nullptr && max_trace > 0) {
    tr = [trace, max_trace](int it, const Vec& s) {
      if (it <= max_trace)
        std::memcpy(trace + (size_t)(it - 1) * s.size(), s.data(),
                    sizeof(double) * s.size());
    };
  }
  Summary s = h->solver->Solve(values, b, D, q_tolerance, r_tolerance, x, tr);
  if (out) {
    out->residual_norm = s.residual_norm;
    out->num_iterations = s.num_iterations;
    out->termination_type = s.termination_type;
    std::snprintf(out->message, sizeof(out->message), "%s", s.message.c_str());
  }
}

// Reduced (camera) solution of the last oracle_solve.
int oracle_reduced_solution(OracleHandle* h, double* out) {
  const Vec& v = h->solver->reduced_solution();
  if (out) std::memcpy(out, v.data(), sizeof(double) * v.size());
  return (int)v.size();
}

// ---- use_explicit_schur_complement = true (explicit_schur.h) ----------------
static ExplicitSchurComplementSolver* ExplicitSolver(OracleHandle* h) {
  if (!h->explicit_solver)
    h->explicit_solver = std::make_unique<ExplicitSchurComplementSolver>(&h->bs, h->options);
  return h->explicit_solver.get();
}

// InitStorage + SchurEliminator::Eliminate. Sizes come back through the three
// counters; a second call with the arrays non-NULL fills them: block CRS of the
// upper triangle (row_ptr: num_blocks + 1, cols / positions: num_cells), the
// values and the reduced right hand side.
void oracle_explicit_eliminate(OracleHandle* h, const double* values, const double* b,
                               const double* D, int* num_blocks, int* num_cells,
                               int* num_values, int* row_ptr, int* cols, int* positions,
                               double* lhs_values, double* rhs) {
  ExplicitSchurComplementSolver* s = ExplicitSolver(h);
  s->Eliminate(values, b, D);
  const BlockRandomAccessSparseMatrix* m = s->lhs();
  if (num_blocks) *num_blocks = (int)m->blocks().size();
  if (num_cells) *num_cells = (int)m->cols().size();
  if (num_values) *num_values = (int)m->values().size();
  if (row_ptr) std::memcpy(row_ptr, m->row_ptr().data(), sizeof(int) * m->row_ptr().size());
  if (cols) std::memcpy(cols, m->cols().data(), sizeof(int) * m->cols().size());
  if (positions) std::memcpy(positions, m->positions().data(), sizeof(int) * m->positions().size());
  if (lhs_values) std::memcpy(lhs_values, m->values().data(), sizeof(double) * m->values().size());
  if (rhs) std::memcpy(rhs, s->rhs().data(), sizeof(double) * s->rhs().size());
}

// y += S x on the matrix of the last oracle_explicit_eliminate / oracle_explicit_solve.
void oracle_explicit_multiply(OracleHandle* h, const double* x, double* y) {
  ExplicitSolver(h)->lhs()->SymmetricRightMultiplyAndAccumulate(x, y);
}

// SchurComplementSolver::SolveImpl with SolveReducedLinearSystemUsingConjugateGradients.
void oracle_explicit_solve(OracleHandle* h, const double* values, const double* b,
                           const double* D, double q_tolerance, double r_tolerance,
                           double* x, OracleSummary* out, double* trace, int max_trace) {
  std::function<void(int, const Vec&)> tr = nullptr;
  if (trace != nullptr && max_trace > 0) {
    tr = [trace, max_trace](int it, const Vec& s) {
      if (it <= max_trace)
        std::memcpy(trace + (size_t)(it - 1) * s.size(), s.data(), sizeof(double) * s.size());
    };
  }
  Summary s = ExplicitSolver(h)->Solve(values, b, D, q_tolerance, r_tolerance, x, tr);
  if (out) {
    out->residual_norm = s.residual_norm;
    out->num_iterations = s.num_iterations;
    out->termination_type = s.termination_type;
    std::snprintf(out->message, sizeof(out->message), "%s", s.message.c_str());
  }
}

// ConjugateGradientsSolver on a dense row-major SPD matrix with the identity
// preconditioner (conjugate_gradients_solver_test.cc:47-160).
void oracle_cg_dense(int n, const double* A, const double* b, double* x,
                     int min_it, int max_it, int reset_period, double q_tol,
                     double r_tol, OracleSummary* out) {
  CGOptions o;
  o.min_num_iterations = min_it;
  o.max_num_iterations = max_it;
  o.residual_reset_period = reset_period;
  o.q_tolerance = q_tol;
  o.r_tolerance = r_tol;
  Vec rhs(b, b + n), sol(x, x + n);
  Operator lhs = [&](const Vec& a, Vec& y) { MatVec(A, n, n, a.data(), y.data(), 1); };
  Operator id = [](const Vec& a, Vec& y) {
    for (size_t i = 0; i < a.size(); ++i) y[i] += a[i];
  };
  Summary s = ConjugateGradientsSolver(o, lhs, rhs, id, sol);
  std::memcpy(x, sol.data(), sizeof(double) * n);
  if (out) {
    out->residual_norm = s.residual_norm;
    out->num_iterations = s.num_iterations;
    out->termination_type = s.termination_type;
    std::snprintf(out->message, sizeof(out->message), "%s", s.message.c_str());
  }
}

// Small dense helpers exposed for unit tests of the restated Eigen primitives.
int oracle_llt_inverse_upper(double* m, int n) { return LltInverseUpper(m, n) ? 1 : 0; }
void oracle_inverse_small(const double* m, int n, double* out) { InverseSmall(m, n, out); }

}  // extern "C"
