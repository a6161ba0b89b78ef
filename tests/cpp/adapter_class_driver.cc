// Test driver: the Ceres-side class of include/ceres_b200_adapter.h (B200IterativeSchurComplementSolver)
// compiled against stand-ins for the Ceres headers it needs (tests/cpp/standin_ceres) and driven through
// LinearSolver::Solve the way levenberg_marquardt_strategy.cc:97-132 does. Prints key=value pairs.
// argv[1]: i = implicit operator, e = use_explicit_schur_complement, u = a structure outside the device
// path's scope (a camera-prior row) with a fallback solver to delegate to, U = the same without fallback;
// argv[2] (optional): file the solution x is written to (raw doubles).
#include <cmath>
#include <cstdio>
#include <memory>
#include <random>

#include "ceres/linear_solver.h"
#include "../../include/ceres_b200_adapter.h"

using namespace ceres::internal;

// what LinearSolver::Create returned before the patch, reduced to a marker
class MarkerSolver final : public LinearSolver {
 public:
  Summary Solve(LinearOperator*, const double*, const PerSolveOptions&, double* x) override {
    Summary s;
    s.termination_type = LinearSolverTerminationType::SUCCESS;
    s.num_iterations = 1;
    s.message = "delegated-to-reference-solver";
    x[0] = 42.0;
    ++calls;
    return s;
  }
  int calls = 0;
};

int main(int argc, char** argv) {
  const char mode = argc > 1 ? argv[1][0] : 'i';
  const bool explicit_schur = mode == 'e';
  const int P = 60, C = 5;
  CompressedRowBlockStructure bs;
  for (int p = 0; p < P; ++p) bs.cols.push_back({3, 3 * p});
  for (int c = 0; c < C; ++c) bs.cols.push_back({9, 3 * P + 9 * c});
  int N = 0;
  for (int p = 0; p < P; ++p) N += 2 + p % 3;
  int r = 0;
  for (int p = 0; p < P; ++p)
    for (int k = 0; k < 2 + p % 3; ++k, ++r) {
      CompressedList row;
      row.block = {2, 2 * r};
      row.cells.push_back({p, 6 * r});
      row.cells.push_back({P + (p + k) % C, 6 * N + 18 * r});
      bs.rows.push_back(row);
    }
  if (mode == 'u' || mode == 'U') {  // a prior on camera 0: a row with an F cell only (not <2,3,9> bundle adjustment rows)
    CompressedList row;
    row.block = {2, 2 * r};
    row.cells.push_back({P, 24 * N});
    bs.rows.push_back(row);
  }
  std::mt19937_64 rng(5);
  std::normal_distribution<double> g(0.0, 1.0);
  const int extra = (mode == 'u' || mode == 'U') ? 1 : 0;
  std::vector<double> values(24 * (size_t)N + 18 * extra), b(2 * (size_t)(N + extra)), D(3 * P + 9 * C), x(3 * P + 9 * C, NAN);
  for (double& v : values) v = g(rng);
  for (double& v : b) v = g(rng);
  for (double& v : D) v = 1.0 + std::fabs(g(rng));
  BlockSparseMatrix A(bs, values);
  LinearSolver::Options options;
  options.preconditioner_type = ceres::SCHUR_JACOBI;
  options.min_num_iterations = 0;
  options.max_num_iterations = 100;
  options.elimination_groups = {P, C};
  options.use_explicit_schur_complement = explicit_schur;
  MarkerSolver* marker = nullptr;
  std::unique_ptr<LinearSolver> fallback;
  if (mode == 'u') { marker = new MarkerSolver; fallback.reset(marker); }
  auto b200 = std::make_unique<B200IterativeSchurComplementSolver>(options, std::move(fallback));
  B200IterativeSchurComplementSolver* adapter = b200.get();
  std::unique_ptr<LinearSolver> solver = std::move(b200);
  LinearSolver::PerSolveOptions per_solve;
  per_solve.D = D.data();
  per_solve.q_tolerance = -1.0;
  per_solve.r_tolerance = 1e-10;
  LinearSolver::Summary s = solver->Solve(&A, b.data(), per_solve, x.data());
  if (mode == 'u') {  // the decision is taken once: the second call goes straight to the fallback
    s = solver->Solve(&A, b.data(), per_solve, x.data());
    std::printf("delegated_calls=%d x0=%g ", marker->calls, x[0]);
  }
  if (argc > 2) {
    if (FILE* f = std::fopen(argv[2], "wb")) { std::fwrite(x.data(), sizeof(double), x.size(), f); std::fclose(f); }
  }
  // residual of the normal equations (J'J + D^2) x = J'b, computed here with plain loops
  double worst = 0.0, scale = 0.0;
  if (s.termination_type == LinearSolverTerminationType::SUCCESS && extra == 0) {
    const int ncols = 3 * P + 9 * C;
    std::vector<double> Jx(2 * (size_t)N, 0.0), lhs(ncols, 0.0), rhs(ncols, 0.0);
    for (int row = 0; row < N; ++row)
      for (const Cell& cell : bs.rows[row].cells) {
        const Block& col = bs.cols[cell.block_id];
        for (int i = 0; i < 2; ++i)
          for (int j = 0; j < col.size; ++j) Jx[2 * row + i] += values[cell.position + i * col.size + j] * x[col.position + j];
      }
    for (int row = 0; row < N; ++row)
      for (const Cell& cell : bs.rows[row].cells) {
        const Block& col = bs.cols[cell.block_id];
        for (int i = 0; i < 2; ++i)
          for (int j = 0; j < col.size; ++j) {
            lhs[col.position + j] += values[cell.position + i * col.size + j] * Jx[2 * row + i];
            rhs[col.position + j] += values[cell.position + i * col.size + j] * b[2 * row + i];
          }
      }
    for (int j = 0; j < ncols; ++j) {
      worst = std::fmax(worst, std::fabs(lhs[j] + D[j] * D[j] * x[j] - rhs[j]));
      scale = std::fmax(scale, std::fabs(rhs[j]));
    }
  }
  std::printf("termination=%d iterations=%d gpus=%d residual=%.3e message=%s\n", (int)s.termination_type, s.num_iterations,
              adapter->num_gpus(), scale > 0 ? worst / scale : -1.0, s.message.c_str());
  return 0;
}
