"""ORACLE — TEST INFRASTRUCTURE ONLY.

CPU restatement (numpy + the oracle linear solver) of the trust-region loop that surrounds the
ITERATIVE_SCHUR solve, for the final-cost parity statement of BASELINE.json's north_star
("final cost within 1e-8 relative at the same LM iteration count"):

  TrustRegionMinimizer::Minimize / IterationZero / EvaluateGradientAndJacobian /
  ComputeTrustRegionStep / HandleInvalidStep / ParameterToleranceReached /
  FunctionToleranceReached / IsStepSuccessful / HandleSuccessfulStep
                                      internal/ceres/trust_region_minimizer.cc:68-138,196-310,381-500,720-840
  LevenbergMarquardtStrategy          internal/ceres/levenberg_marquardt_strategy.cc:69-171
  TrustRegionStepEvaluator            internal/ceres/trust_region_step_evaluator.cc:38-110
  BlockSparseMatrix::SquaredColumnNorm / ScaleColumns   internal/ceres/block_sparse_matrix.cc:351-450

Monotonic steps, no bounds, no inner iterations (what examples/bundle_adjuster.cc runs by default).
The residuals, the Jacobian, the column norms / scaling and J*step come from oracle/snavely_reference.py
(ReferenceProblem: numpy restatement of SnavelyReprojectionError + Jet autodiff, no product code);
only the problem's DATA (structure, observations, parameters) is taken from the object passed in.
independent_evaluator=False keeps the object's own evaluate / squared_column_norm / scale_columns /
jacobian_times instead (bench.py's CPU timing leg, where the product's OpenMP host evaluator is only a
faster clock, not a checker). linear_solve(values, residuals, D, eta) returns (y, summary) for J y = r.
"""
import numpy as np

DEFAULTS = dict(max_num_iterations=50, initial_trust_region_radius=1e4, max_trust_region_radius=1e16,
                min_trust_region_radius=1e-32, min_relative_decrease=1e-3, min_lm_diagonal=1e-6,
                max_lm_diagonal=1e32, eta=1e-1, function_tolerance=1e-6, gradient_tolerance=1e-10,
                parameter_tolerance=1e-8, jacobi_scaling=True, max_num_consecutive_invalid_steps=5)


def minimize(problem, cameras, points, linear_solve, independent_evaluator=True, **options):
    o = dict(DEFAULTS)
    o.update(options)
    if independent_evaluator:
        from oracle.snavely_reference import ReferenceProblem
        problem = ReferenceProblem.of(problem)
    P = problem.P
    x = np.concatenate([points, cameras]).astype(np.float64)  # column order of the Jacobian: points, cameras
    split = lambda v: (v[3 * P:], v[:3 * P])  # -> cameras, points
    iterations = []
    scale = None

    def evaluate_gradient_and_jacobian(xv, first):
        nonlocal scale
        cam, pts = split(xv)
        cost, residuals, values = problem.evaluate(cameras=cam, points=pts)
        gradient = jacobian_transpose_times(problem, values, residuals)
        norm2 = problem.squared_column_norm(values)
        if o["jacobi_scaling"]:
            if first:
                scale = 1.0 / (1.0 + np.sqrt(norm2))
            problem.scale_columns(values, scale)
        return cost, residuals, values, np.abs(gradient).max(), np.linalg.norm(gradient)

    x_cost, residuals, values, gmax, gnorm = evaluate_gradient_and_jacobian(x, True)
    summary = dict(initial_cost=x_cost, num_successful_steps=0, num_unsuccessful_steps=0,
                   num_linear_solver_iterations=0)
    it = dict(iteration=0, step_is_valid=True, step_is_successful=True, cost=x_cost, cost_change=0.0,
              gradient_max_norm=gmax, gradient_norm=gnorm, step_norm=0.0, relative_decrease=0.0,
              linear_solver_iterations=0)
    radius, decrease_factor = o["initial_trust_region_radius"], 2.0
    reuse_diagonal, diagonal = False, None
    current_cost = reference_cost = minimum_cost = candidate_best = x_cost
    acc_reference = acc_candidate = 0.0
    num_invalid, at_least_one_successful = 0, False

    def finish(termination, message):
        cam, pts = split(x)
        summary.update(termination_type=termination, message=message, final_cost=x_cost, iterations=iterations,
                       cameras=cam.copy(), points=pts.copy())
        return summary

    while True:
        summary["num_successful_steps" if it["step_is_successful"] else "num_unsuccessful_steps"] += 1
        it["trust_region_radius"] = radius
        iterations.append(it)
        if it["iteration"] >= o["max_num_iterations"]:
            return finish("NO_CONVERGENCE", "Maximum number of iterations reached. Number of iterations: %d." % it["iteration"])
        if it["step_is_successful"] and it["gradient_max_norm"] <= o["gradient_tolerance"]:
            return finish("CONVERGENCE", "Gradient tolerance reached.")
        if radius <= o["min_trust_region_radius"]:
            return finish("CONVERGENCE", "Minimum trust region radius reached.")
        prev = it
        it = dict(iteration=prev["iteration"] + 1, step_is_valid=False, step_is_successful=False, cost_change=0.0,
                  step_norm=0.0, relative_decrease=0.0)
        # LevenbergMarquardtStrategy::ComputeStep
        if not reuse_diagonal:
            diagonal = np.clip(problem.squared_column_norm(values), o["min_lm_diagonal"], o["max_lm_diagonal"])
        D = np.sqrt(diagonal / radius)
        y, ls = linear_solve(values, residuals, D, o["eta"])
        reuse_diagonal = True
        it["linear_solver_iterations"] = ls["num_iterations"]
        summary["num_linear_solver_iterations"] += ls["num_iterations"]
        if ls["termination_type"] == 3:
            return finish("FAILURE", "Linear solver failed due to unrecoverable non-numeric causes.")
        step_is_valid = False
        if ls["termination_type"] != 2 and np.all(np.isfinite(y)):
            step = -y
            model_residuals = problem.jacobian_times(values, step)
            model_cost_change = -model_residuals @ (residuals + model_residuals / 2.0)
            step_is_valid = model_cost_change > 0.0
        it["step_is_valid"] = bool(step_is_valid)
        if not step_is_valid:
            num_invalid += 1
            if num_invalid >= o["max_num_consecutive_invalid_steps"]:
                return finish("FAILURE", "Number of consecutive invalid steps more than "
                              "Solver::Options::max_num_consecutive_invalid_steps: %d" % o["max_num_consecutive_invalid_steps"])
            radius /= decrease_factor
            decrease_factor *= 2.0
            it.update(cost=x_cost, gradient_max_norm=prev["gradient_max_norm"], gradient_norm=prev["gradient_norm"])
            continue
        num_invalid = 0
        delta = step * scale if o["jacobi_scaling"] else step
        candidate = x + delta
        cam, pts = split(candidate)
        candidate_cost, _, _ = problem.evaluate(cameras=cam, points=pts, jacobian=False)
        if not np.isfinite(candidate_cost):
            candidate_cost = np.finfo(np.float64).max
        x_norm = np.linalg.norm(x)
        it["step_norm"] = np.linalg.norm(x - candidate)
        if at_least_one_successful and it["step_norm"] <= o["parameter_tolerance"] * (x_norm + o["parameter_tolerance"]):
            return finish("CONVERGENCE", "Parameter tolerance reached.")
        it["cost_change"] = x_cost - candidate_cost
        if abs(it["cost_change"]) <= o["function_tolerance"] * x_cost:
            return finish("CONVERGENCE", "Function tolerance reached.")
        if candidate_cost >= np.finfo(np.float64).max:
            it["relative_decrease"] = np.finfo(np.float64).min
        else:
            it["relative_decrease"] = max((current_cost - candidate_cost) / model_cost_change,
                                          (reference_cost - candidate_cost) / (acc_reference + model_cost_change))
        if it["relative_decrease"] > o["min_relative_decrease"]:
            at_least_one_successful = True
            x = candidate
            x_cost, residuals, values, gmax, gnorm = evaluate_gradient_and_jacobian(x, False)
            it.update(cost=x_cost, gradient_max_norm=gmax, gradient_norm=gnorm, step_is_successful=True)
            radius = min(o["max_trust_region_radius"],
                         radius / max(1.0 / 3.0, 1.0 - (2.0 * it["relative_decrease"] - 1.0) ** 3))
            decrease_factor = 2.0
            reuse_diagonal = False
            current_cost = candidate_cost
            acc_candidate += model_cost_change
            acc_reference += model_cost_change
            if current_cost < minimum_cost:
                minimum_cost = candidate_best = current_cost
                acc_candidate = 0.0
            elif current_cost > candidate_best:
                candidate_best = current_cost
                acc_candidate = 0.0
            reference_cost, acc_reference = candidate_best, acc_candidate
        else:
            it.update(cost=candidate_cost, gradient_max_norm=prev["gradient_max_norm"],
                      gradient_norm=prev["gradient_norm"])
            radius /= decrease_factor
            decrease_factor *= 2.0
            reuse_diagonal = True


def jacobian_transpose_times(problem, values, r):
    """J' r (the gradient the evaluator returns, program_evaluator.h:189-260)."""
    N, P, C = problem.N, problem.P, problem.C
    row_point = np.repeat(np.arange(P), np.diff(problem.point_row_start))
    E = values[:6 * N].reshape(N, 2, 3)
    F = values[6 * N:].reshape(N, 2, 9)
    rr = r.reshape(N, 2)
    ge = np.einsum("nij,ni->nj", E, rr)
    gf = np.einsum("nij,ni->nj", F, rr)
    out = np.zeros(3 * P + 9 * C)
    for k in range(3):
        out[k:3 * P:3] = np.bincount(row_point, weights=ge[:, k], minlength=P)
    for k in range(9):
        out[3 * P + k::9] = np.bincount(problem.cam_idx, weights=gf[:, k], minlength=C)
    return out
