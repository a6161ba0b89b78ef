/* ceres_b200_schur.h — C ABI of libceres_b200_schur.so
 *
 * B200-native (sm_100a, FP64 CUDA) ITERATIVE_SCHUR linear solve for Ceres
 * Solver's Levenberg-Marquardt loop. This header is the drop-in boundary: every
 * entry point names the reference interface (file:line under the Ceres tree) it
 * replaces. Plain pointers and sizes only; all pointers are HOST pointers unless
 * a function says otherwise. INTEGRATION.md shows the C++ adapter
 * (B200IterativeSchurComplementSolver : BlockSparseMatrixSolver) a Ceres
 * maintainer adds on top of this header.
 *
 * There is no CPU fallback: every compute entry point fails with CB200_ERR_CUDA
 * when no CUDA device is usable.
 */
#ifndef CERES_B200_SCHUR_H_
#define CERES_B200_SCHUR_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes returned by every int function */
enum {
  CB200_OK = 0,
  CB200_ERR_INVALID_ARGUMENT = -1,
  CB200_ERR_UNSUPPORTED_STRUCTURE = -2, /* not a <2,3,9> Schur structure        */
  CB200_ERR_CUDA = -3,
  CB200_ERR_NCCL = -4,
  CB200_ERR_STATE = -5                   /* call order (e.g. solve before upload) */
};

/* ceres::PreconditionerType values this solver accepts
 * (include/ceres/types.h; iterative_schur_complement_solver.cc:159-199).
 * CLUSTER_JACOBI / CLUSTER_TRIDIAGONAL are rejected at create time. */
enum {
  CB200_IDENTITY = 0,
  CB200_JACOBI = 1,
  CB200_SCHUR_JACOBI = 2,
  CB200_SCHUR_POWER_SERIES_EXPANSION = 3
};

/* ceres::internal::LinearSolverTerminationType (linear_solver.h:57-74) */
enum {
  CB200_LINEAR_SOLVER_SUCCESS = 0,
  CB200_LINEAR_SOLVER_NO_CONVERGENCE = 1,
  CB200_LINEAR_SOLVER_FAILURE = 2,
  CB200_LINEAR_SOLVER_FATAL_ERROR = 3
};

/* CompressedRowBlockStructure (block_structure.h:52-96) flattened into int
 * arrays: row block r has size row_block_size[r], first scalar row
 * row_block_pos[r], and cells cell_ptr[r] .. cell_ptr[r+1]-1, each with a column
 * block id and the position of its row-major values in the values array. */
typedef struct cb200_structure {
  int num_row_blocks;
  const int* row_block_size;
  const int* row_block_pos;
  const int* cell_ptr;       /* num_row_blocks + 1 */
  const int* cell_block_id;
  const int* cell_position;
  int num_col_blocks;
  const int* col_block_size;
  const int* col_block_pos;
} cb200_structure;

/* Mirror of the LinearSolver::Options fields this solver consumes
 * (linear_solver.h:150-230; iterative_schur_complement_solver.cc:74,100-129,178-186)
 * plus the device placement. */
typedef struct cb200_options {
  int num_eliminate_blocks;   /* options.elimination_groups[0]                   */
  int preconditioner_type;    /* CB200_* above                                   */
  int min_num_iterations;     /* 0 in Solver::Options defaults (solver.h:594)    */
  int max_num_iterations;
  int residual_reset_period;  /* 10                                              */
  int use_spse_initialization;
  int max_num_spse_iterations;
  double spse_tolerance;
  /* device placement, two ways (linear_solver.h:150-230 has num_threads / context for the same job):
   * (a) ONE handle drives several GPUs of the box from the caller's thread (what a single-process
   *     ceres::Solve needs): leave rank = 0, world_size = 1 and set num_gpus (and optionally
   *     device_ids; default 0 .. num_gpus-1). The library shards the points over the GPUs, runs one
   *     internal worker thread per GPU and sums the camera-sized partials through NVLink peer
   *     memory (cudaDeviceEnablePeerAccess; no NCCL on the implicit path).
   * (b) one process per GPU: rank/world_size shard the points (and their residual rows) into
   *     contiguous ranges of balanced observation count; world_size > 1 needs nccl_unique_id
   *     (128 bytes from cb200_nccl_unique_id on rank 0, broadcast by the caller). num_gpus <= 1. */
  int device_id;
  int rank;
  int world_size;
  const char* nccl_unique_id;
  int y_replicas;             /* 0 = default; privatised copies of the camera
                                 accumulator that spread FP64 atomics           */
  int status_check_period;    /* 0 = default; PCG iterations launched between
                                 two reads of the device convergence flag       */
  int use_explicit_schur_complement; /* linear_solver.h:161, linear_solver.cc:111-116:
                                 form the reduced camera matrix S as a block-sparse
                                 matrix once per solve and run PCG on it
                                 (SparseSchurComplementSolver::
                                 SolveReducedLinearSystemUsingConjugateGradients,
                                 schur_complement_solver.cc:337-408). SCHUR_JACOBI only,
                                 no SPSE initialisation (solver.cc:277-290).          */
  int explicit_switch_iterations; /* 0 = never (default). n > 0 (one GPU, IDENTITY / JACOBI /
                                 SCHUR_JACOBI): a solve on the implicit operator that is
                                 still running after n PCG iterations forms the explicit
                                 matrix once and continues THE SAME iteration on it (same
                                 operator, same preconditioner, same vectors; products
                                 agree to 1e-12). Not a reference option: an internal
                                 speed-up for long solves.              */
  int num_gpus;               /* 0 / 1 = one GPU (device_id). n > 1 = mode (a) above      */
  int device_ids[16];         /* mode (a): CUDA ordinals; all zero = 0 .. num_gpus-1      */
  int deterministic;          /* 1 = bitwise-repeatable solves: the camera scatter F'w of every
                                 product is summed per camera in a fixed order (a second,
                                 camera-major pass) instead of FP64 atomics, so borderline
                                 convergence tests cannot flip between runs. 1.8x the time per
                                 product (measured, Venice shape). Implicit operator only.                        */
  int cuda_graphs;            /* 0 = default (on): the launch sequence of a solve is captured once
                                 and replayed as CUDA graphs; -1 = off (plain stream launches)  */
  int h2d_chunks;             /* 0 = default (16): cb200_solve copies the Jacobian in this many
                                 point-range chunks and runs the E'E / rhs pass on each chunk
                                 while the next one is still crossing PCIe; 1 = one copy      */
} cb200_options;

/* LinearSolver::Summary (linear_solver.h:320-326) + per-phase device timings. */
typedef struct cb200_summary {
  double residual_norm;       /* left at -1 like the reference                   */
  int num_iterations;
  int termination_type;
  char message[512];
  /* milliseconds: h2d / d2h / total from CUDA events on the solver's stream, the phases in between
   * from %globaltimer stamps the kernels write at the phase boundaries */
  float h2d_ms, setup_ms, pcg_ms, back_substitute_ms, d2h_ms;
  int kernel_launches;        /* kernels launched by the last solve              */
  float total_ms;             /* whole call on the device: cb200_solve_resident = setup + PCG + back
                                 substitution; cb200_solve = first H2D byte to last D2H byte (with the
                                 overlapped upload the phases above overlap, so they need not add up) */
} cb200_summary;

typedef struct cb200_solver cb200_solver;

const char* cb200_last_error(void);
int cb200_version(void);

/* 128-byte ncclUniqueId for a multi-process communicator (rank 0 calls it). */
int cb200_nccl_unique_id(char* out128);

/* LinearSolver::Create, case ITERATIVE_SCHUR (linear_solver.cc:111-116) +
 * the first-call work of IterativeSchurComplementSolver::SolveImpl
 * (iterative_schur_complement_solver.cc:76-83: DetectStructure,
 * detect_structure.cc:39-120; PartitionedMatrixView ctor,
 * partitioned_matrix_view_impl.h:47-105). Validates <2,3,9>, flattens the
 * structure to device arrays, shards by point, allocates every device buffer. */
int cb200_create(const cb200_structure* structure, const cb200_options* options,
                 cb200_solver** out);
void cb200_destroy(cb200_solver* s);

/* The point sharding cb200_create applies, without a device (host only): rank g
 * owns points [point_bounds[g], point_bounds[g+1]) and row blocks
 * [row_bounds[g], row_bounds[g+1]); contiguous ranges of balanced observation
 * count (the idea of PartitionRangeForParallelFor over cumulative_nnz,
 * block_sparse_matrix.cc:58-67). Both outputs have world_size + 1 entries. */
int cb200_shard_plan(const cb200_structure* structure, int num_eliminate_blocks, int world_size,
                     int* point_bounds, int* row_bounds);

/* The block structure cb200_create gives the explicit Schur complement, without a
 * device (host only): SparseSchurComplementSolver::InitStorage
 * (schur_complement_solver.cc:224-290) + the BlockRandomAccessSparseMatrix layout
 * (block_random_access_sparse_matrix.cc:47-103). Upper block triangle in block CRS,
 * columns ascending, cell k holds 81 row-major doubles at 81 k. row_ptr (C + 1) and
 * cols (*num_cells) may be NULL to query the size. */
int cb200_explicit_plan(const cb200_structure* structure, int num_eliminate_blocks,
                        int* num_cells, int* row_ptr, int* cols);

/* Page-lock a caller-owned host buffer (cudaHostRegister) so that cb200_solve /
 * cb200_upload copy from it at full PCIe rate instead of through the driver's
 * staging buffer. Meant for BlockSparseMatrix::values(), whose address is stable
 * for the life of a Solve(); the reference's own switch for this is
 * BlockSparseMatrix(..., use_page_locked_memory) (block_sparse_matrix.cc:178-180,
 * 810-826). Registering a buffer twice is not an error. */
int cb200_host_register(const void* ptr, size_t bytes);
int cb200_host_unregister(const void* ptr);

/* sizes: num_cols = 3P + 9C (x), num_cols_f = 9C, num_rows = 2N. */
int cb200_num_rows(const cb200_solver* s);
int cb200_num_cols(const cb200_solver* s);
int cb200_num_cols_e(const cb200_solver* s);
int cb200_num_cols_f(const cb200_solver* s);
/* this rank's point range [begin,end) and row-block range */
int cb200_shard(const cb200_solver* s, int* point_begin, int* point_end,
                int* row_block_begin, int* row_block_end);

/* TypedLinearSolver<BlockSparseMatrix>::Solve -> SolveImpl
 * (linear_solver.h:366-387; iterative_schur_complement_solver.cc:64-157).
 * values = A->values() (full array, positions as in the structure), b has
 * num_rows entries, D has num_cols entries or is NULL, x has num_cols entries;
 * x is left untouched on FAILURE / FATAL_ERROR (:150-154). With world_size > 1
 * each rank writes its own points' entries of x and all camera entries. */
int cb200_solve(cb200_solver* s, const double* values, const double* b,
                const double* D, double q_tolerance, double r_tolerance,
                double* x, cb200_summary* summary);

/* The same solve split at the host<->device copies (measurement with inputs
 * resident in HBM; also what an on-device evaluator would call). Any of
 * values / b / D may be NULL in cb200_upload to keep what is resident
 * (D == NULL on the first upload means no regularisation). */
int cb200_upload(cb200_solver* s, const double* values, const double* b,
                 const double* D);
int cb200_solve_resident(cb200_solver* s, double q_tolerance, double r_tolerance,
                         cb200_summary* summary);
int cb200_download_solution(cb200_solver* s, double* x);
/* per_solve_options.D == nullptr for the following resident solves
 * (linear_solver.h:237-257); the next cb200_upload with D != NULL re-enables it. */
int cb200_clear_diagonal(cb200_solver* s);
/* change min/max_num_iterations of an existing handle (fixed-work timing runs) */
int cb200_set_iterations(cb200_solver* s, int min_num_iterations, int max_num_iterations);

/* Iterate trace for parity tests: the reduced (camera) solution after each of
 * the first max_trace PCG iterations of the next solves is kept on the device
 * and copied into trace (max_trace x num_cols_f, row-major) at the end. */
int cb200_set_trace(cb200_solver* s, double* trace, int max_trace);

/* ---- finer-grained seams (unit parity, ncu); operate on uploaded values ---- */

/* PartitionedMatrixView<2,3,9> products (partitioned_matrix_view_impl.h):
 * RightMultiplyAndAccumulateE :112-137, RightMultiplyAndAccumulateF :139-191,
 * LeftMultiplyAndAccumulateE :193-264, LeftMultiplyAndAccumulateF :266-375.
 * y is read, accumulated into and written back (full-length host vectors;
 * with world_size > 1 only this rank's rows / points contribute). */
int cb200_pmv_right_e(cb200_solver* s, const double* x_e, double* y_rows);
int cb200_pmv_right_f(cb200_solver* s, const double* x_f, double* y_rows);
int cb200_pmv_left_e(cb200_solver* s, const double* x_rows, double* y_e);
int cb200_pmv_left_f(cb200_solver* s, const double* x_rows, double* y_f);

/* ImplicitSchurComplement::Init (implicit_schur_complement.cc:49-97):
 * UpdateBlockDiagonalEtE (+FtF for JACOBI / SPSE), AddDiagonalAndInvert
 * (:179-204), UpdateRhs (:251-276). */
int cb200_isc_init(cb200_solver* s);
int cb200_isc_ete_inverse(cb200_solver* s, double* out /* 9 * local points */);
int cb200_isc_ftf_inverse(cb200_solver* s, double* out /* 81 * C */);
int cb200_isc_rhs(cb200_solver* s, double* out /* 9C */);
/* ImplicitSchurComplement::RightMultiplyAndAccumulate (:106-144): y = S x */
int cb200_isc_multiply(cb200_solver* s, const double* x, double* y);
/* InversePowerSeriesOperatorRightMultiplyAccumulate (:146-174): y += ... */
int cb200_isc_pse_operator(cb200_solver* s, const double* x, double* y);
/* BackSubstitute (:208-243): y = [ (E'E+D^2)^-1 E'(b - F z) ; z ] */
int cb200_isc_back_substitute(cb200_solver* s, const double* z, double* y);

/* SchurJacobiPreconditioner::UpdateImpl (schur_jacobi_preconditioner.cc:87-97):
 * SchurEliminator<2,3,9>::Eliminate into a BlockRandomAccessDiagonalMatrix
 * (schur_eliminator_impl.h:176-304) then Invert
 * (block_random_access_diagonal_matrix.cc:90-100). invert == 0 stops before
 * the inversion. blocks (81*C, row-major 9x9) may be NULL. D is the one last
 * uploaded. */
int cb200_schur_jacobi_update(cb200_solver* s, int invert, double* blocks);
/* SchurJacobiPreconditioner::RightMultiplyAndAccumulate (:99-104): y += M^-1 x */
int cb200_schur_jacobi_apply(cb200_solver* s, const double* x, double* y);
/* PowerSeriesExpansionPreconditioner::RightMultiplyAndAccumulate
 * (power_series_expansion_preconditioner.cc:57-82): y = sum Z^i (F'F)^-1 x */
int cb200_pse_apply(cb200_solver* s, int max_num_spse_iterations,
                    double spse_tolerance, const double* x, double* y);

/* y += B x for a block-diagonal matrix of num_blocks row-major 9x9 blocks (host pointers: blocks 81 n, x 9 n,
 * y 9 n) -- what cb200_schur_jacobi_apply computes, WITHOUT a solver handle and WITHOUT any allocation: no
 * cudaMalloc / cudaFree / allocating library call, so it can be used where those are forbidden. It has no
 * counterpart in the reference. The operands are staged in
 * one file-scope static device array of 99 * 16384 doubles (12.98 MB, held by every GPU that loads the library
 * until the process exits): num_blocks > 16384 is rejected with CB200_ERR_INVALID_ARGUMENT. Runs on the current
 * device, legacy default stream, synchronously. All calls share the array, which keeps the last call's contents
 * (every call overwrites what it reads; nothing needs to start at zero): calls that overlap on the GPU race. */
int cb200_block_diagonal_apply_static(const double* blocks, const double* x, double* y, int num_blocks);

/* Device-resident timing of one op (CUDA events on the solver's stream, `iters`
 * back-to-back launches after `warmup`); ms_out = average per launch.
 * op: 0 S*x (isc multiply), 1 EtE inverse build, 2 rhs, 3 Schur-Jacobi build,
 * 4 back substitute, 6 the fused S*x kernel alone; the separate products that the
 * reference's evaluation_benchmark.cc times (:333-491): 7 / 8 Right / Left E,
 * 9 / 10 Right / Left F, 11 UpdateBlockDiagonalFtF, 12 PSE preconditioner apply;
 * explicit-Schur solvers: 13 SchurEliminator::Eliminate (rhs pass + outer products +
 * diagonal), 14 the outer-product kernel alone, 15 S*x on the explicit matrix. */
int cb200_time_op(cb200_solver* s, int op, int warmup, int iters, float* ms_out,
                  int* launches_out);

/* ---- use_explicit_schur_complement seams (solvers created with the option) ----
 * cb200_explicit_eliminate: SchurEliminator::Eliminate on the resident values / b / D
 * (schur_eliminator_impl.h:176-304): lhs_values (num_cells x 81, the structure of
 * cb200_explicit_plan; NULL to skip) and the reduced rhs (9C; NULL to skip).
 * cb200_explicit_multiply: y = S x,
 * BlockRandomAccessSparseMatrix::SymmetricRightMultiplyAndAccumulate
 * (block_random_access_sparse_matrix.cc:128-170) on the matrix of the last
 * eliminate / solve. */
int cb200_explicit_num_cells(const cb200_solver* s);
int cb200_explicit_eliminate(cb200_solver* s, double* lhs_values, double* rhs);
int cb200_explicit_multiply(cb200_solver* s, const double* x, double* y);

/* ---- widened boundary: the Snavely bundle-adjustment
 * evaluator and the trust-region loop run on the device-resident problem, so the
 * Jacobian never crosses PCIe. The structure given to cb200_create must be the
 * evaluator's canonical layout. -------------------------------------------------- */

/* ceres::TerminationType (include/ceres/types.h) */
enum { CB200_CONVERGENCE = 0, CB200_NO_CONVERGENCE = 1, CB200_FAILURE = 2 };

/* The Solver::Options fields TrustRegionMinimizer and LevenbergMarquardtStrategy
 * consume (include/ceres/solver.h:263-321, trust_region_minimizer.cc,
 * levenberg_marquardt_strategy.cc:46-65). */
typedef struct cb200_lm_options {
  int max_num_iterations;               /* 50                                   */
  double initial_trust_region_radius;   /* 1e4                                  */
  double max_trust_region_radius;       /* 1e16                                 */
  double min_trust_region_radius;       /* 1e-32                                */
  double min_relative_decrease;         /* 1e-3                                 */
  double min_lm_diagonal;               /* 1e-6                                 */
  double max_lm_diagonal;               /* 1e32                                 */
  double eta;                           /* 1e-1 (bundle_adjuster.cc:114: 1e-2)  */
  double function_tolerance;            /* 1e-6                                 */
  double gradient_tolerance;            /* 1e-10                                */
  double parameter_tolerance;           /* 1e-8                                 */
  int jacobi_scaling;                   /* 1                                    */
  int max_num_consecutive_invalid_steps;/* 5                                    */
} cb200_lm_options;

/* ceres::IterationSummary (include/ceres/iteration_callback.h), the fields this loop fills */
typedef struct cb200_lm_iteration {
  int iteration;
  int step_is_valid, step_is_successful;
  double cost, cost_change, gradient_max_norm, gradient_norm, step_norm;
  double relative_decrease, trust_region_radius, eta;
  int linear_solver_iterations;
  float iteration_ms;                   /* wall clock of the iteration           */
} cb200_lm_iteration;

typedef struct cb200_lm_summary {
  double initial_cost, final_cost;
  int num_iterations;                   /* entries written to `iterations` (incl. iteration 0) */
  int num_successful_steps, num_unsuccessful_steps;
  int termination_type;                 /* CB200_CONVERGENCE / NO_CONVERGENCE / FAILURE */
  char message[256];
  float total_ms, evaluate_ms, linear_solver_ms;
  int num_linear_solver_iterations;
} cb200_lm_summary;

/* Program + ProgramEvaluator for SnavelyReprojectionError residuals
 * (examples/snavely_reprojection_error.h:57-92, program_evaluator.h:189-260):
 * observations has 2N entries in row order; cameras 9C, points 3P. */
int cb200_bal_attach(cb200_solver* s, const double* observations);
int cb200_bal_set_parameters(cb200_solver* s, const double* cameras, const double* points);
int cb200_bal_get_parameters(cb200_solver* s, double* cameras, double* points);
/* Evaluator::Evaluate at the resident parameters: cost, and with jacobian != 0 the
 * residuals and the UNSCALED Jacobian, written straight into the solver's device
 * arrays (as cb200_upload would). gradient (3P + 9C, J'r) may be NULL. */
int cb200_bal_evaluate(cb200_solver* s, int jacobian, double* cost, double* gradient);
/* the device-resident Jacobian / residuals back in the evaluator's layout (tests) */
int cb200_download_jacobian(cb200_solver* s, double* values, double* residuals);
/* TrustRegionMinimizer::Minimize with LevenbergMarquardtStrategy
 * (trust_region_minimizer.cc:68-138, levenberg_marquardt_strategy.cc:69-171),
 * starting from the resident parameters; the result replaces them. iterations may
 * be NULL; at most max_iterations_out entries are written. */
int cb200_lm_minimize(cb200_solver* s, const cb200_lm_options* options, cb200_lm_summary* summary,
                      cb200_lm_iteration* iterations, int max_iterations_out);

#ifdef __cplusplus
}
#endif
#endif /* CERES_B200_SCHUR_H_ */
