"""GPU parity tests of ITERATIVE_SCHUR with use_explicit_schur_complement = true
(SURVEY 8(f) row 3): the explicit block-sparse reduced camera matrix, its symmetric product and
the PCG solve on it against the CPU oracle (oracle/explicit_schur.h, the restatement of
SparseSchurComplementSolver + SchurEliminator + BlockRandomAccessSparseMatrix) on the same seeded
inputs. Tolerances: cells of S, rhs and S x <= 1e-12 relative; PCG iterates and the solution
<= 1e-9 relative, equal iteration counts, termination types and message texts.
Modelled on schur_eliminator_test.cc:127-226, block_random_access_sparse_matrix_test.cc:126-141
and solver_test.cc:1107-1230.
"""
import re

import numpy as np
import pytest

from oracle import binding as ob
from tests import util

pytestmark = pytest.mark.gpu

OP_TOL = 1e-12
ITERATE_TOL = 1e-9


def problem(C=9, P=200, N=900, seed=11, dup_pairs=0):
    m = util.pkg()
    p = m.SyntheticBAL(C, P, N, seed=seed, dup_pairs=dup_pairs, geometry=False)
    values, b, D = p.random_values(seed)
    return m, p.structure, values, b, D


def long_point_problem():
    """Points of 3 .. 300 rows: several kFormRows chunks per point, chunk pairs (A, B > A)."""
    m = util.pkg()
    C, P = 300, 40
    lens = np.full(P, 3)
    for i, n in [(7, 300), (20, 257), (12, 100), (30, 33), (31, 64), (39, 16), (5, 17), (6, 1)]:
        lens[i] = n
    prs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    rng = np.random.default_rng(0)
    # cameras deliberately NOT sorted within a point: the plan sorts them
    cam = np.concatenate([rng.choice(C, size=L, replace=False) for L in lens]).astype(np.int32)
    st = m.BlockStructure.bal(prs, cam, C)
    N = len(cam)
    values, b, D = rng.standard_normal(24 * N), rng.standard_normal(2 * N), rng.standard_normal(3 * P + 9 * C) + 3
    return m, st, values, b, D


def compare_eliminate(m, st, values, b, D):
    o = ob.Oracle(util.oracle_structure(st), num_threads=4, preconditioner_type=ob.SCHUR_JACOBI)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_JACOBI, use_explicit_schur_complement=True)
    if D is None:
        g.upload(values, b)
    else:
        g.upload(values, b, D)
    row_ptr, cols, pos, want, rhs_want = o.explicit_eliminate(values, b, D)
    got, rhs = g.explicit_eliminate()
    assert got.shape == (len(cols), 81)
    assert util.rel_err(rhs, rhs_want) < OP_TOL
    want = want.reshape(-1, 81)
    # every cell on its own scale (small off-diagonal cells must be right too)
    scale = np.maximum(np.abs(want).max(axis=1, keepdims=True), 1e-300)
    assert (np.abs(got - want) / scale).max() < 1e-11
    assert util.rel_err(got, want) < OP_TOL
    x = np.random.default_rng(5).standard_normal(st.num_cols_f)
    assert util.rel_err(g.explicit_multiply(x), o.explicit_multiply(x)) < OP_TOL
    g.close()


@pytest.mark.parametrize("dup_pairs", [0, 40])
@pytest.mark.parametrize("use_D", [True, False])
def test_explicit_eliminate_matches_oracle(dup_pairs, use_D):
    m, st, values, b, D = problem(dup_pairs=dup_pairs)
    compare_eliminate(m, st, values, b, D if use_D else None)


def test_explicit_eliminate_long_and_unsorted_points():
    compare_eliminate(*long_point_problem())


def test_explicit_schur_complement_is_the_dense_one():
    m, st, values, b, D = problem(C=4, P=30, N=100, seed=3)
    A = util.oracle_structure(st).dense(values)
    H = A.T @ A + np.diag(D * D)
    ne = st.num_cols_e
    S = H[ne:, ne:] - H[ne:, :ne] @ np.linalg.solve(H[:ne, :ne], H[:ne, ne:])
    g = m.IterativeSchurComplementSolver(st, use_explicit_schur_complement=True)
    g.upload(values, b, D)
    cells, rhs = g.explicit_eliminate()
    row_ptr, cols = m.explicit_plan(st)
    for i in range(4):
        for k in range(row_ptr[i], row_ptr[i + 1]):
            j = cols[k]
            assert np.abs(cells[k].reshape(9, 9) - S[9 * i:9 * i + 9, 9 * j:9 * j + 9]).max() < 1e-10 * np.abs(S).max()
    got = np.stack([g.explicit_multiply(e) for e in np.eye(st.num_cols_f)], axis=1)
    assert np.abs(got - S).max() < 1e-10 * np.abs(S).max()


def compare_solve(m, st, values, b, D, q_tol=-1.0, r_tol=1e-10, max_it=200, min_it=0, check_period=0):
    o = ob.Oracle(util.oracle_structure(st), num_threads=4, preconditioner_type=ob.SCHUR_JACOBI,
                  min_num_iterations=min_it, max_num_iterations=max_it)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_JACOBI, min_num_iterations=min_it,
                                         max_num_iterations=max_it, status_check_period=check_period,
                                         use_explicit_schur_complement=True)
    ntrace = min(max_it, 64)
    g.set_trace(ntrace)
    xo, so = o.explicit_solve(values, b, D, q_tol, r_tol, max_trace=ntrace)
    xg, sg = g.solve(values, b, D, q_tol, r_tol, x=np.full(st.num_cols, np.nan))
    assert sg["termination_type"] == so["termination_type"], (sg, so)
    assert sg["num_iterations"] == so["num_iterations"], (sg["message"], so["message"])
    for k in range(len(so.get("trace", []))):
        assert util.rel_err(sg["trace"][k], so["trace"][k]) < ITERATE_TOL, k
    assert util.rel_err(xg, xo) < ITERATE_TOL
    strip = lambda t: re.sub(r"-?\d\.\d+e[-+]\d+", "#", t)
    assert strip(sg["message"]) == strip(so["message"])
    g.close()
    return xg, sg, so


@pytest.mark.parametrize("use_D", [True, False])
def test_explicit_solve_matches_oracle_and_dense(use_D):
    m, st, values, b, D = problem()
    x, sg, so = compare_solve(m, st, values, b, D if use_D else None)
    assert sg["termination_type"] == 0
    A = util.oracle_structure(st).dense(values)
    H = A.T @ A + (np.diag(D * D) if use_D else 0)
    assert util.rel_err(x, np.linalg.solve(H, A.T @ b)) < 1e-7


@pytest.mark.parametrize("period", [1, 3, 8])
def test_explicit_solve_lm_tolerances_and_check_period(period):
    m, st, values, b, D = problem(C=21, P=1500, N=7000, seed=5, dup_pairs=6)
    compare_solve(m, st, values, b, D, q_tol=1e-2, r_tol=-1.0, check_period=period)


def test_explicit_solve_residual_reset_and_long_points():
    m, st, values, b, D = long_point_problem()
    _, sg, _ = compare_solve(m, st, values, b, D, r_tol=1e-9, max_it=300)
    assert sg["num_iterations"] > 10  # crosses a residual reset (period 10)


def test_explicit_solve_without_success_leaves_points_zero():
    """schur_complement_solver.cc:135,155-160: x is zero-filled, the CG solution is always written,
    back substitution runs only after SUCCESS."""
    m, st, values, b, D = problem()
    x, sg, so = compare_solve(m, st, values, b, D, r_tol=1e-30, max_it=2)
    assert ob.TERMINATION[sg["termination_type"]] == "NO_CONVERGENCE"
    assert np.all(x[:st.num_cols_e] == 0.0) and np.all(np.isfinite(x[st.num_cols_e:]))


def test_explicit_and_implicit_solves_agree_at_lm_shape():
    """Ladybug-like LM iteration: the explicit and the implicit solver are two orders of arithmetic on
    one reduced system; same iteration count, solutions equal to 1e-9."""
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug", scale=0.05)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    xe, se, _ = compare_solve(m, p.structure, values, residuals, D, q_tol=1e-3, r_tol=-1.0, max_it=300)
    gi = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=300)
    xi, si = gi.solve(values, residuals, D, 1e-3, -1.0)
    assert si["num_iterations"] == se["num_iterations"]
    assert util.rel_err(xe, xi) < ITERATE_TOL


def test_lm_loop_with_explicit_schur_matches_reference_loop():
    """The device trust-region loop (cb200_lm_minimize) on a handle created with
    use_explicit_schur_complement, against the CPU restatement of the loop driving the oracle's explicit
    solver: same accepted steps, same PCG iteration counts, costs within 1e-8 relative."""
    from oracle import lm_reference as lmr
    m = util.pkg()
    p = m.SyntheticBAL(31, 3000, 14000, seed=7, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    st = p.structure
    o = ob.Oracle(util.oracle_structure(st), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=500)
    want = lmr.minimize(p, p.cameras, p.points, lambda v, r, D, e: o.explicit_solve(v, r, D, e, -1.0), eta=1e-2,
                        max_num_iterations=8)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=500,
                                         use_explicit_schur_complement=True)
    g.bal_attach(p.observations)
    g.bal_set_parameters(p.cameras, p.points)
    got = g.lm_minimize(eta=1e-2, max_num_iterations=8)
    assert got["termination_type"] == want["termination_type"]
    assert len(got["iterations"]) == len(want["iterations"])
    for a, b in zip(got["iterations"], want["iterations"]):
        assert a["iteration"] == b["iteration"]
        assert bool(a["step_is_successful"]) == bool(b["step_is_successful"])
        assert a["linear_solver_iterations"] == b["linear_solver_iterations"], (a, b)
        assert abs(a["cost"] - b["cost"]) <= 1e-8 * b["cost"], (a, b)
    assert abs(got["final_cost"] - want["final_cost"]) <= 1e-8 * want["final_cost"]
    assert want["final_cost"] < 0.2 * want["initial_cost"]


def test_explicit_seams_report_call_order_errors():
    """Programmer errors come back as negative status codes with a message, never as a crash
    (SURVEY 8(b) "Errors")."""
    m, st, values, b, D = problem()
    gi = m.IterativeSchurComplementSolver(st)
    gi.upload(values, b, D)
    with pytest.raises(m.B200Error, match="error -5"):
        gi.explicit_eliminate()          # not created with use_explicit_schur_complement
    with pytest.raises(m.B200Error, match="error -5"):
        gi.explicit_multiply(np.zeros(st.num_cols_f))
    ge = m.IterativeSchurComplementSolver(st, use_explicit_schur_complement=True)
    with pytest.raises(m.B200Error, match="error -5"):
        ge.explicit_multiply(np.zeros(st.num_cols_f))   # nothing eliminated yet
    with pytest.raises(m.B200Error, match="error -5"):
        ge.explicit_eliminate()                          # nothing uploaded
    ge.upload(values, b, D)
    ge.explicit_eliminate()
    ge.upload(values)                                    # new values invalidate the matrix
    with pytest.raises(m.B200Error, match="error -5"):
        ge.explicit_multiply(np.zeros(st.num_cols_f))


@pytest.mark.parametrize("switch_at,period", [(4, 4), (1, 1), (8, 3)])
def test_switch_to_explicit_matrix_mid_solve_keeps_the_iterates(switch_at, period):
    """explicit_switch_iterations: the implicit solve forms the explicit matrix after `switch_at` iterations
    and continues the same PCG on it. Same operator, so every iterate, the iteration count, the
    termination and the solution still match the oracle's IMPLICIT solver within the usual bars."""
    from tests.test_gpu_parity import compare_solve as compare_implicit
    m, st, values, b, D = long_point_problem()
    _, sg, _ = compare_implicit(m, st, values, b, D, "SCHUR_JACOBI", r_tol=1e-9, max_it=300, check_period=period,
                                explicit_switch_iterations=switch_at)
    assert sg["num_iterations"] > 12  # ran on both operators, across a residual reset on the explicit one
    m2 = util.pkg()
    p = m2.SyntheticBAL.named("ladybug", scale=0.05)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    for ptype in ("SCHUR_JACOBI", "JACOBI"):
        compare_implicit(m2, p.structure, values, residuals, D, ptype, q_tol=1e-3, r_tol=-1.0, max_it=300,
                         check_period=period, explicit_switch_iterations=switch_at)


def test_switch_option_validation():
    m, st, values, b, D = problem()
    with pytest.raises(m.B200Error, match="explicit_switch_iterations"):
        m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_POWER_SERIES_EXPANSION,
                                         explicit_switch_iterations=5)
