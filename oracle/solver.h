// ORACLE — TEST INFRASTRUCTURE ONLY (see small_blas.h header).
//
// CPU restatement of the solver layer:
//   internal/ceres/conjugate_gradients_solver.h:107-305  ConjugateGradientsSolver
//   internal/ceres/eigen_vector_ops.h:46-101             Norm/Dot/Axpby/SetZero/Copy
//   internal/ceres/power_series_expansion_preconditioner.cc:57-82
//   internal/ceres/iterative_schur_complement_solver.cc:64-157  SolveImpl
//   internal/ceres/iterative_schur_complement_solver.cc:159-199 CreatePreconditioner
//   internal/ceres/linear_solver.h:57-74,320-326         termination types, Summary
#ifndef ORACLE_SOLVER_H_
#define ORACLE_SOLVER_H_

#include <cmath>
#include <cstdio>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "implicit_schur.h"
#include "schur_jacobi.h"

namespace oracle {

enum TerminationType {  // linear_solver.h:57-74
  LINEAR_SOLVER_SUCCESS = 0,
  LINEAR_SOLVER_NO_CONVERGENCE = 1,
  LINEAR_SOLVER_FAILURE = 2,
  LINEAR_SOLVER_FATAL_ERROR = 3,
};

struct Summary {
  double residual_norm = -1.0;
  int num_iterations = -1;
  int termination_type = LINEAR_SOLVER_FAILURE;
  std::string message;
};

using Vec = std::vector<double>;
using Operator = std::function<void(const Vec&, Vec&)>;  // y (+)= A x

inline double Dot(const Vec& a, const Vec& b) {
  double s = 0.0;
  for (size_t i = 0; i < a.size(); ++i) s += a[i] * b[i];
  return s;
}
inline double Norm(const Vec& a) { return std::sqrt(Dot(a, a)); }
inline void Axpby(double a, const Vec& x, double b, const Vec& y, Vec& z) {
  for (size_t i = 0; i < x.size(); ++i) z[i] = a * x[i] + b * y[i];
}

struct CGOptions {
  int min_num_iterations = 1;
  int max_num_iterations = 1;
  int residual_reset_period = 10;
  double r_tolerance = 0.0;
  double q_tolerance = 0.0;
};

inline std::string Fmt(const char* f, double a = 0, double b = 0, double c = 0) {
  char buf[512];
  std::snprintf(buf, sizeof(buf), f, a, b, c);
  return buf;
}

// trace(iteration, solution) is called after each solution update when set.
inline Summary ConjugateGradientsSolver(
    const CGOptions& options, const Operator& lhs, const Vec& rhs,
    const Operator& preconditioner, Vec& solution,
    const std::function<void(int, const Vec&)>& trace = nullptr) {
  auto IsZeroOrInfinity = [](double x) { return x == 0.0 || std::isinf(x); };
  const size_t n = rhs.size();
  Vec p(n), r(n), z(n), tmp(n);
  Summary summary;
  summary.termination_type = LINEAR_SOLVER_NO_CONVERGENCE;
  summary.message = "Maximum number of iterations reached.";
  summary.num_iterations = 0;

  const double norm_rhs = Norm(rhs);
  if (norm_rhs == 0.0) {
    std::fill(solution.begin(), solution.end(), 0.0);
    summary.termination_type = LINEAR_SOLVER_SUCCESS;
    summary.message = "Convergence. |b| = 0.";
    return summary;
  }
  const double tol_r = options.r_tolerance * norm_rhs;

  std::fill(tmp.begin(), tmp.end(), 0.0);
  lhs(solution, tmp);
  Axpby(1.0, rhs, -1.0, tmp, r);
  double norm_r = Norm(r);
  if (options.min_num_iterations == 0 && norm_r <= tol_r) {
    summary.termination_type = LINEAR_SOLVER_SUCCESS;
    summary.message = Fmt("Convergence. |r| = %e <= %e.", norm_r, tol_r);
    return summary;
  }

  double rho = 1.0;
  Axpby(1.0, rhs, 1.0, r, tmp);
  double Q0 = -Dot(solution, tmp);

  for (summary.num_iterations = 1;; ++summary.num_iterations) {
    std::fill(z.begin(), z.end(), 0.0);
    preconditioner(r, z);

    const double last_rho = rho;
    rho = Dot(r, z);
    if (IsZeroOrInfinity(rho)) {
      summary.termination_type = LINEAR_SOLVER_FAILURE;
      summary.message = Fmt("Numerical failure. rho = r'z = %e.", rho);
      break;
    }
    if (summary.num_iterations == 1) {
      p = z;
    } else {
      const double beta = rho / last_rho;
      if (IsZeroOrInfinity(beta)) {
        summary.termination_type = LINEAR_SOLVER_FAILURE;
        summary.message = Fmt(
            "Numerical failure. beta = rho_n / rho_{n-1} = %e, "
            "rho_n = %e, rho_{n-1} = %e", beta, rho, last_rho);
        break;
      }
      Axpby(1.0, z, beta, p, p);
    }

    Vec& q = z;
    std::fill(q.begin(), q.end(), 0.0);
    lhs(p, q);
    const double pq = Dot(p, q);
    if ((pq <= 0) || std::isinf(pq)) {
      summary.termination_type = LINEAR_SOLVER_NO_CONVERGENCE;
      summary.message = Fmt(
          "Matrix is indefinite, no more progress can be made. "
          "p'q = %e. |p| = %e, |q| = %e", pq, Norm(p), Norm(q));
      break;
    }
    const double alpha = rho / pq;
    if (std::isinf(alpha)) {
      summary.termination_type = LINEAR_SOLVER_FAILURE;
      summary.message = Fmt(
          "Numerical failure. alpha = rho / pq = %e, rho = %e, pq = %e.", alpha,
          rho, pq);
      break;
    }
    Axpby(1.0, solution, alpha, p, solution);
    if (trace) trace(summary.num_iterations, solution);

    if (summary.num_iterations % options.residual_reset_period == 0) {
      std::fill(tmp.begin(), tmp.end(), 0.0);
      lhs(solution, tmp);
      Axpby(1.0, rhs, -1.0, tmp, r);
    } else {
      Axpby(1.0, r, -alpha, q, r);
    }

    Axpby(1.0, rhs, 1.0, r, tmp);
    const double Q1 = -Dot(solution, tmp);
    const double zeta = summary.num_iterations * (Q1 - Q0) / Q1;
    if (zeta < options.q_tolerance &&
        summary.num_iterations >= options.min_num_iterations) {
      summary.termination_type = LINEAR_SOLVER_SUCCESS;
      char buf[512];
      std::snprintf(buf, sizeof(buf),
                    "Iteration: %d Convergence: zeta = %e < %e. |r| = %e",
                    summary.num_iterations, zeta, options.q_tolerance, Norm(r));
      summary.message = buf;
      break;
    }
    Q0 = Q1;

    norm_r = Norm(r);
    if (norm_r <= tol_r &&
        summary.num_iterations >= options.min_num_iterations) {
      summary.termination_type = LINEAR_SOLVER_SUCCESS;
      char buf[512];
      std::snprintf(buf, sizeof(buf),
                    "Iteration: %d Convergence. |r| = %e <= %e.",
                    summary.num_iterations, norm_r, tol_r);
      summary.message = buf;
      break;
    }
    if (summary.num_iterations >= options.max_num_iterations) break;
  }
  return summary;
}

// PowerSeriesExpansionPreconditioner::RightMultiplyAndAccumulate:
// y = sum_{i=0..k} Z^i (F'F)^-1 x ; y is overwritten.
inline void PowerSeriesApply(const ImplicitSchurComplement& isc,
                             int max_num_spse_iterations, double spse_tolerance,
                             int num_threads, const double* x, double* y) {
  const int n = isc.num_rows();
  Vec series_term(n), previous_series_term(n);
  for (int i = 0; i < n; ++i) y[i] = 0.0;
  isc.block_diagonal_FtF_inverse().RightMultiplyAndAccumulate(x, y, num_threads);
  for (int i = 0; i < n; ++i) previous_series_term[i] = y[i];
  double ynorm = 0.0;
  for (int i = 0; i < n; ++i) ynorm += y[i] * y[i];
  const double norm_threshold = spse_tolerance * std::sqrt(ynorm);
  for (int i = 1;; i++) {
    std::fill(series_term.begin(), series_term.end(), 0.0);
    isc.InversePowerSeriesOperatorRightMultiplyAccumulate(
        previous_series_term.data(), series_term.data());
    for (int k = 0; k < n; ++k) y[k] = y[k] + series_term[k];
    if (i >= max_num_spse_iterations || Norm(series_term) < norm_threshold) break;
    std::swap(previous_series_term, series_term);
  }
}

class IterativeSchurComplementSolver {
 public:
  IterativeSchurComplementSolver(const BlockStructure* bs,
                                 const SolverOptions& options)
      : bs_(bs), options_(options) {
    tbs_ = CreateTranspose(*bs);
  }

  const ImplicitSchurComplement* schur_complement() const { return isc_.get(); }
  const SchurJacobiPreconditioner* schur_jacobi() const { return sj_.get(); }
  const Vec& reduced_solution() const { return reduced_solution_; }

  Summary Solve(const double* values, const double* b, const double* D,
                double q_tolerance, double r_tolerance, double* x,
                const std::function<void(int, const Vec&)>& trace = nullptr) {
    const int neb = options_.num_eliminate_blocks;
    if (!isc_) isc_ = std::make_unique<ImplicitSchurComplement>(options_);
    isc_->Init(bs_, &tbs_, values, D, b);

    const int num_sc_blocks = (int)bs_->cols.size() - neb;
    if (num_sc_blocks == 0) {
      Summary s;
      s.num_iterations = 0;
      s.termination_type = LINEAR_SOLVER_SUCCESS;
      isc_->BackSubstitute(nullptr, x);
      return s;
    }
    const int n = isc_->num_rows();
    reduced_solution_.assign(n, 0.0);
    if (options_.use_spse_initialization) {
      PowerSeriesApply(*isc_, options_.max_num_spse_iterations,
                       options_.spse_tolerance, options_.num_threads,
                       isc_->rhs().data(), reduced_solution_.data());
    }

    Operator precond;
    const int nt = options_.num_threads;
    switch (options_.preconditioner_type) {
      case IDENTITY:
        precond = [](const Vec& a, Vec& y) {
          for (size_t i = 0; i < a.size(); ++i) y[i] += a[i];
        };
        break;
      case JACOBI:
        precond = [this, nt](const Vec& a, Vec& y) {
          isc_->block_diagonal_FtF_inverse().RightMultiplyAndAccumulate(
              a.data(), y.data(), nt);
        };
        break;
      case SCHUR_POWER_SERIES_EXPANSION:
        // spse_tolerance forced to 0 so the preconditioner stays fixed (:178-186)
        precond = [this, nt](const Vec& a, Vec& y) {
          PowerSeriesApply(*isc_, options_.max_num_spse_iterations, 0.0, nt,
                           a.data(), y.data());
        };
        break;
      case SCHUR_JACOBI:
        if (!sj_)
          sj_ = std::make_unique<SchurJacobiPreconditioner>(bs_, neb, nt);
        if (!sj_->Update(values, D)) {
          Summary s;
          s.num_iterations = 0;
          s.termination_type = LINEAR_SOLVER_FAILURE;
          s.message = "Preconditioner update failed.";
          return s;
        }
        precond = [this](const Vec& a, Vec& y) {
          sj_->RightMultiplyAndAccumulate(a.data(), y.data());
        };
        break;
      default: {
        Summary s;
        s.termination_type = LINEAR_SOLVER_FATAL_ERROR;
        s.message = "Unknown Preconditioner Type";
        return s;
      }
    }

    CGOptions cg;
    cg.min_num_iterations = options_.min_num_iterations;
    cg.max_num_iterations = options_.max_num_iterations;
    cg.residual_reset_period = options_.residual_reset_period;
    cg.q_tolerance = q_tolerance;
    cg.r_tolerance = r_tolerance;
    Operator lhs = [this](const Vec& a, Vec& y) {
      isc_->RightMultiplyAndAccumulate(a.data(), y.data());
    };
    Summary summary = ConjugateGradientsSolver(cg, lhs, isc_->rhs(), precond,
                                               reduced_solution_, trace);
    if (summary.termination_type != LINEAR_SOLVER_FAILURE &&
        summary.termination_type != LINEAR_SOLVER_FATAL_ERROR) {
      isc_->BackSubstitute(reduced_solution_.data(), x);
    }
    return summary;
  }

 private:
  const BlockStructure* bs_;
  BlockStructure tbs_;
  SolverOptions options_;
  std::unique_ptr<ImplicitSchurComplement> isc_;
  std::unique_ptr<SchurJacobiPreconditioner> sj_;
  Vec reduced_solution_;
};

}  // namespace oracle

#endif  // ORACLE_SOLVER_H_
