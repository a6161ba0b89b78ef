// bundle_adjuster: the reference's examples/bundle_adjuster.cc flow on the B200 path.
//
//   bundle_adjuster --input=problem.txt [--num_iterations=5] [--eta=1e-2]
//       [--preconditioner=schur_jacobi] [--max_linear_solver_iterations=500] ...
//
// Same flags (the subset this path implements; anything else the reference offers is
// refused, not silently ignored), same order of operations as SolveProblem
// (bundle_adjuster.cc:349-377): read the BAL file, optional initial PLY, Normalize,
// Perturb, build the Schur-ordered program, minimise with TRUST_REGION /
// LEVENBERG_MARQUARDT / ITERATIVE_SCHUR and tolerances 1e-16, print the report,
// optional final PLY. Everything after the file parsing runs on the GPU through the
// C ABI (include/ceres_b200_schur.h): device Snavely evaluator, device LM loop.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../ceres-solver-ceres-solver_b200/csrc/bal_io.h"
#include "../include/ceres_b200_schur.h"

namespace {

struct Flags {
  std::map<std::string, std::string> v = {
      {"input", ""},
      {"trust_region_strategy", "levenberg_marquardt"},
      {"linear_solver", "iterative_schur"},
      {"explicit_schur_complement", "false"},
      {"preconditioner", "schur_jacobi"},
      {"use_spse_initialization", "false"},
      {"spse_tolerance", "0.1"},
      {"max_num_spse_iterations", "5"},
      {"eta", "1e-2"},
      {"num_iterations", "5"},
      {"max_linear_solver_iterations", "500"},
      {"rotation_sigma", "0.0"},
      {"translation_sigma", "0.0"},
      {"point_sigma", "0.0"},
      {"random_seed", "38401"},
      {"initial_ply", ""},
      {"final_ply", ""},
      {"final_bal", ""},        // extra: the refined problem as a BAL file
      {"device", "0"},          // extra: CUDA device
      {"inner_iterations", "false"},
      {"use_quaternions", "false"},
      {"use_manifolds", "false"},
      {"robustify", "false"},
      {"line_search", "false"},
      {"nonmonotonic_steps", "false"},
      {"mixed_precision_solves", "false"},
  };
  bool Bool(const std::string& k) const { const std::string& s = v.at(k); return s == "true" || s == "1"; }
  double Double(const std::string& k) const { return std::atof(v.at(k).c_str()); }
  int Int(const std::string& k) const { return std::atoi(v.at(k).c_str()); }
  const std::string& Str(const std::string& k) const { return v.at(k); }
};

// absl-style: --name=value, --name value, --flag, --noflag
bool Parse(int argc, char** argv, Flags* f, std::string* error) {
  for (int i = 1; i < argc; ++i) {
    std::string a = argv[i];
    if (a.rfind("--", 0) != 0) { *error = "unexpected argument " + a; return false; }
    a = a.substr(2);
    std::string name = a, value;
    bool has_value = false;
    const size_t eq = a.find('=');
    if (eq != std::string::npos) { name = a.substr(0, eq); value = a.substr(eq + 1); has_value = true; }
    if (!has_value && f->v.count(name) == 0 && name.rfind("no", 0) == 0 && f->v.count(name.substr(2))) {
      name = name.substr(2); value = "false"; has_value = true;
    }
    if (f->v.count(name) == 0) { *error = "unknown flag --" + name; return false; }
    if (!has_value) {
      const std::string& cur = f->v[name];
      if (cur == "true" || cur == "false") value = "true";
      else if (i + 1 < argc) value = argv[++i];
      else { *error = "flag --" + name + " needs a value"; return false; }
    }
    f->v[name] = value;
  }
  return true;
}

int PreconditionerCode(const std::string& s) {
  if (s == "identity") return CB200_IDENTITY;
  if (s == "jacobi") return CB200_JACOBI;
  if (s == "schur_jacobi") return CB200_SCHUR_JACOBI;
  if (s == "schur_power_series_expansion") return CB200_SCHUR_POWER_SERIES_EXPANSION;
  return -1;
}

int Die(const std::string& what) {
  std::fprintf(stderr, "bundle_adjuster: %s\n", what.c_str());
  return 1;
}

}  // namespace

int main(int argc, char** argv) {
  Flags flags;
  std::string error;
  if (!Parse(argc, argv, &flags, &error)) return Die(error);
  if (flags.Str("input").empty()) return Die("Usage: bundle_adjuster --input=bal_problem");
  // what this path is: say so instead of quietly doing something else
  if (flags.Str("linear_solver") != "iterative_schur")
    return Die("only --linear_solver=iterative_schur runs on this path (got " + flags.Str("linear_solver") + ")");
  if (flags.Str("trust_region_strategy") != "levenberg_marquardt")
    return Die("only --trust_region_strategy=levenberg_marquardt is implemented");
  for (const char* k : {"inner_iterations", "use_quaternions", "use_manifolds", "robustify",
                        "line_search", "nonmonotonic_steps", "mixed_precision_solves"})
    if (flags.Bool(k)) return Die(std::string("--") + k + " is not implemented on this path");
  const int precond = PreconditionerCode(flags.Str("preconditioner"));
  if (precond < 0)
    return Die("--preconditioner=" + flags.Str("preconditioner") +
               " is not implemented (identity, jacobi, schur_jacobi, schur_power_series_expansion are)");

  cb200_bal::BalProblem bal;
  error = cb200_bal::Read(flags.Str("input"), &bal);
  if (!error.empty()) return Die(error);
  if (!flags.Str("initial_ply").empty()) {
    error = cb200_bal::WritePly(flags.Str("initial_ply"), bal);
    if (!error.empty()) return Die(error);
  }
  std::srand(flags.Int("random_seed"));  // as the reference does; Perturb itself uses a default-seeded mt19937
  cb200_bal::Normalize(&bal);
  cb200_bal::Perturb(&bal, flags.Double("rotation_sigma"), flags.Double("translation_sigma"),
                     flags.Double("point_sigma"));

  // Schur-ordered program: points first (eliminated), rows grouped by point; values in the
  // BlockJacobianWriter layout (all E cells, then all F cells: block_jacobian_writer.cc:102-165)
  const int C = bal.num_cameras, P = bal.num_points, N = bal.num_observations;
  std::vector<int> order, row_start;
  cb200_bal::GroupByPoint(bal, &order, &row_start);
  std::vector<int> row_block_size(N, 2), row_block_pos(N), cell_ptr(N + 1), cell_block_id(2 * (size_t)N),
      cell_position(2 * (size_t)N), col_block_size(P + C), col_block_pos(P + C);
  std::vector<double> observations(2 * (size_t)N);
  for (int p = 0; p < P; ++p) {
    for (int r = row_start[p]; r < row_start[p + 1]; ++r) {
      const int o = order[r];
      row_block_pos[r] = 2 * r;
      cell_ptr[r] = 2 * r;
      cell_block_id[2 * (size_t)r] = p;
      cell_block_id[2 * (size_t)r + 1] = P + bal.camera_index[o];
      cell_position[2 * (size_t)r] = 6 * r;
      cell_position[2 * (size_t)r + 1] = 6 * N + 18 * r;
      observations[2 * (size_t)r] = bal.observations[2 * (size_t)o];
      observations[2 * (size_t)r + 1] = bal.observations[2 * (size_t)o + 1];
    }
  }
  cell_ptr[N] = 2 * N;
  for (int p = 0; p < P; ++p) { col_block_size[p] = 3; col_block_pos[p] = 3 * p; }
  for (int c = 0; c < C; ++c) { col_block_size[P + c] = 9; col_block_pos[P + c] = 3 * P + 9 * c; }
  cb200_structure st{};
  st.num_row_blocks = N;
  st.row_block_size = row_block_size.data();
  st.row_block_pos = row_block_pos.data();
  st.cell_ptr = cell_ptr.data();
  st.cell_block_id = cell_block_id.data();
  st.cell_position = cell_position.data();
  st.num_col_blocks = P + C;
  st.col_block_size = col_block_size.data();
  st.col_block_pos = col_block_pos.data();

  cb200_options o{};
  o.num_eliminate_blocks = P;
  o.preconditioner_type = precond;
  o.min_num_iterations = 0;
  o.max_num_iterations = flags.Int("max_linear_solver_iterations");
  o.residual_reset_period = 10;
  o.use_spse_initialization = flags.Bool("use_spse_initialization");
  o.max_num_spse_iterations = flags.Int("max_num_spse_iterations");
  o.spse_tolerance = flags.Double("spse_tolerance");
  o.use_explicit_schur_complement = flags.Bool("explicit_schur_complement");  // bundle_adjuster.cc:91,168
  o.device_id = flags.Int("device");
  o.rank = 0;
  o.world_size = 1;
  cb200_solver* solver = nullptr;
  if (cb200_create(&st, &o, &solver) != CB200_OK) return Die(cb200_last_error());
  auto fail = [&](const char* what) {
    std::fprintf(stderr, "bundle_adjuster: %s: %s\n", what, cb200_last_error());
    cb200_destroy(solver);
    return 1;
  };
  if (cb200_bal_attach(solver, observations.data()) != CB200_OK) return fail("attach");
  if (cb200_bal_set_parameters(solver, bal.cameras.data(), bal.points.data()) != CB200_OK) return fail("parameters");

  cb200_lm_options lm{};
  lm.max_num_iterations = flags.Int("num_iterations");
  lm.initial_trust_region_radius = 1e4;
  lm.max_trust_region_radius = 1e16;
  lm.min_trust_region_radius = 1e-32;
  lm.min_relative_decrease = 1e-3;
  lm.min_lm_diagonal = 1e-6;
  lm.max_lm_diagonal = 1e32;
  lm.eta = flags.Double("eta");
  lm.function_tolerance = 1e-16;   // bundle_adjuster.cc:366-368
  lm.gradient_tolerance = 1e-16;
  lm.parameter_tolerance = 1e-16;
  lm.jacobi_scaling = 1;
  lm.max_num_consecutive_invalid_steps = 5;
  std::vector<cb200_lm_iteration> its(lm.max_num_iterations + 2);
  cb200_lm_summary sum{};
  const auto t0 = std::chrono::steady_clock::now();
  if (cb200_lm_minimize(solver, &lm, &sum, its.data(), (int)its.size()) != CB200_OK) return fail("minimize");
  const double wall = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (cb200_bal_get_parameters(solver, bal.cameras.data(), bal.points.data()) != CB200_OK) return fail("parameters");

  // ---- report, in the shape of Solver::Summary::FullReport + the minimizer's progress table ----
  std::printf("\nSolver Summary (B200 ITERATIVE_SCHUR path)\n\n");
  std::printf("Parameter blocks      %12d   (cameras %d, points %d)\n", C + P, C, P);
  std::printf("Parameters            %12d\n", 9 * C + 3 * P);
  std::printf("Residual blocks       %12d\n", N);
  std::printf("Residuals             %12d\n\n", 2 * N);
  std::printf("Minimizer                        TRUST_REGION\nTrust region strategy     LEVENBERG_MARQUARDT\n");
  std::printf("Linear solver                 ITERATIVE_SCHUR\nPreconditioner        %s\n\n",
              flags.Str("preconditioner").c_str());
  std::printf("iter      cost      cost_change  |gradient|   |step|    tr_ratio  tr_radius  ls_iter  iter_time\n");
  for (int i = 0; i < sum.num_iterations && i < (int)its.size(); ++i) {
    const cb200_lm_iteration& it = its[i];
    std::printf("%4d % .6e  % .2e   %.2e   %.2e  % .2e   %.2e   %6d   %.2e\n", it.iteration, it.cost, it.cost_change,
                it.gradient_max_norm, it.step_norm, it.relative_decrease, it.trust_region_radius,
                it.linear_solver_iterations, it.iteration_ms * 1e-3);
  }
  std::printf("\nCost:\nInitial        %22.6e\nFinal          %22.6e\nChange         %22.6e\n\n", sum.initial_cost,
              sum.final_cost, sum.initial_cost - sum.final_cost);
  std::printf("Minimizer iterations  %12d\nSuccessful steps      %12d\nUnsuccessful steps    %12d\n",
              sum.num_iterations, sum.num_successful_steps, sum.num_unsuccessful_steps);
  std::printf("Linear solver iterations %9d\n\n", sum.num_linear_solver_iterations);
  std::printf("Time (in seconds):\n  Residual + Jacobian evaluation %10.6f\n  Linear solver                  %10.6f\n"
              "  Minimizer                      %10.6f\n  Wall (incl. parameter copies)  %10.6f\n\n",
              sum.evaluate_ms * 1e-3, sum.linear_solver_ms * 1e-3, sum.total_ms * 1e-3, wall);
  const char* term = sum.termination_type == CB200_CONVERGENCE      ? "CONVERGENCE"
                     : sum.termination_type == CB200_NO_CONVERGENCE ? "NO_CONVERGENCE"
                                                                    : "FAILURE";
  std::printf("Termination:  %s (%s)\n", term, sum.message);

  int rc = 0;
  if (!flags.Str("final_ply").empty()) {
    error = cb200_bal::WritePly(flags.Str("final_ply"), bal);
    if (!error.empty()) rc = Die(error);
  }
  if (!flags.Str("final_bal").empty()) {
    error = cb200_bal::Write(flags.Str("final_bal"), bal);
    if (!error.empty()) rc = Die(error);
  }
  cb200_destroy(solver);
  return rc;
}
