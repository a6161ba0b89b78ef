"""GPU parity tests added in round 2: the headline configurations at FULL size against the oracle, the
control-flow exits of ConjugateGradientsSolver the reference's callers rely on, and the round-2 machinery
(CUDA graphs, chunked upload, back substitution from the iteration history, deterministic mode) each
checked against the plain path and the oracle. Multi-GPU handles: tests/test_gpu_multi_inprocess.py.
"""
import os
import re

import numpy as np
import pytest

from oracle import binding as ob
from tests import util
from tests.test_gpu_parity import compare_solve, small_problem

pytestmark = pytest.mark.gpu

OP_TOL = 1e-12
ITERATE_TOL = 1e-9


# ---- BASELINE.json configs at full size against the oracle (implicit_schur_complement_test.cc:155-213,
# ---- iterative_schur_complement_solver_test.cc:76-115 are the reference tests these follow) ---------------

def _lm_problem(name):
    m = util.pkg()
    p = m.SyntheticBAL.named(name, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    return m, p, values, residuals, D


@pytest.mark.parametrize("ptype", ["SCHUR_JACOBI", "SCHUR_POWER_SERIES_EXPANSION"])
def test_venice_full_size_matches_oracle(ptype):
    """configs[2] / configs[4]: the bench problem itself (Venice shape, Snavely Jacobian, Jacobi scaling, LM
    diagonal): unit operators <= 1e-12 and the whole solve (eta = 1e-2, iterate by iterate) <= 1e-9."""
    m, p, values, residuals, D = _lm_problem("venice")
    st = p.structure
    code = ["IDENTITY", "JACOBI", "SCHUR_JACOBI", "SCHUR_POWER_SERIES_EXPANSION"].index(ptype)
    threads = os.cpu_count()
    o = ob.Oracle(util.oracle_structure(st), num_threads=threads, preconditioner_type=code, max_num_iterations=500)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=code, max_num_iterations=500)
    g.upload(values, residuals, D)
    g.isc_init()
    o.isc_init(values, D, residuals)
    x = np.random.default_rng(3).standard_normal(st.num_cols_f)
    assert util.rel_err(g.isc_rhs(), o.isc_rhs()) < OP_TOL
    assert util.rel_err(g.isc_multiply(x), o.isc_multiply(x)) < OP_TOL
    assert util.rel_err(g.ete_inverse(), o.ete_inverse()) < OP_TOL
    if ptype == "SCHUR_JACOBI":
        assert util.rel_err(g.schur_jacobi(invert=False), o.schur_jacobi(values, D, invert=False)) < OP_TOL
    else:
        assert util.rel_err(g.pse_apply(x, 5, 0.0), o.pse_apply(x, 5, 0.0)) < 1e-11
    g.close()
    del o
    compare_solve(m, st, values, residuals, D, ptype, q_tol=1e-2, r_tol=-1.0, max_it=500, threads=threads)


def test_final13682_full_size_matches_oracle():
    """configs[3] (13 682 cameras, 4.46 M points, 29.0 M observations): the operators of one solve against
    the oracle at full size (the oracle needs about six times its Venice time here, so no iterate trace)."""
    m = util.pkg()
    p = m.SyntheticBAL.named("final-13682", geometry=False)
    values, b, D = p.random_values(3)
    D = np.abs(D) + 1.0
    st = p.structure
    threads = os.cpu_count()
    o = ob.Oracle(util.oracle_structure(st), num_threads=threads, preconditioner_type=ob.SCHUR_JACOBI)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_JACOBI)
    g.upload(values, b, D)
    g.isc_init()
    o.isc_init(values, D, b)
    x = np.random.default_rng(5).standard_normal(st.num_cols_f)
    assert util.rel_err(g.isc_rhs(), o.isc_rhs()) < OP_TOL
    assert util.rel_err(g.isc_multiply(x), o.isc_multiply(x)) < OP_TOL
    assert util.rel_err(g.isc_back_substitute(x), o.isc_back_substitute(x)) < OP_TOL
    assert util.rel_err(g.schur_jacobi(invert=False), o.schur_jacobi(values, D, invert=False)) < OP_TOL
    g.close()


def test_problem16_and_ladybug_shapes_match_oracle():
    """configs[0] (problem-16-22106 shape; the data file itself is absent from the reference tree) and
    configs[1] (Ladybug shape): LM-shaped solves, iterate by iterate."""
    for name in ("problem-16", "ladybug"):
        m, p, values, residuals, D = _lm_problem(name)
        compare_solve(m, p.structure, values, residuals, D, "SCHUR_JACOBI", q_tol=1e-2, r_tol=-1.0, max_it=500,
                      threads=os.cpu_count())


# ---- ConjugateGradientsSolver's failure exits and the "x untouched on FAILURE" contract ---------------------
# (conjugate_gradients_solver.h:167-216; iterative_schur_complement_solver.cc:150-154)

def _solve_both(m, st, values, b, D, code, sentinel=None, **kw):
    o = ob.Oracle(util.oracle_structure(st), num_threads=2, preconditioner_type=code, max_num_iterations=50)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=code, max_num_iterations=50, **kw)
    # (both tolerances off: with |rhs| = inf the r test would "converge" at once, |r| = inf <= inf)
    _, so = o.solve(values, b, D, -1.0, -1.0)
    x = np.full(st.num_cols, sentinel) if sentinel is not None else None
    xg, sg = g.solve(values, b, D, -1.0, -1.0, x=x)
    g.close()
    strip = lambda t: re.sub(r"-?\d\.\d+e[-+]\d+", "#", t)
    assert sg["termination_type"] == so["termination_type"], (sg, so)
    assert sg["num_iterations"] == so["num_iterations"], (sg, so)
    assert strip(sg["message"]) == strip(so["message"]), (sg["message"], so["message"])
    return xg, sg


@pytest.mark.parametrize("fused", [True, False])
def test_cg_rho_failure_leaves_x_untouched(fused, monkeypatch):
    """rho = r'z overflows on the first iteration -> FAILURE "Numerical failure. rho = r'z = inf."; the
    caller's x (pre-filled with a marker, levenberg_marquardt_strategy.cc:110) must come back bit-identical."""
    if not fused:
        monkeypatch.setenv("CB200_NO_FUSED_STEP", "1")
    m, p, values, b, D = small_problem(C=9, P=200, N=900, seed=12)
    x, s = _solve_both(m, p.structure, values, b * 1e200, D, ob.IDENTITY, sentinel=-7.25)
    assert s["termination_type"] == 2 and "rho = r'z = inf" in s["message"]
    assert np.all(x == -7.25)


@pytest.mark.parametrize("fused", [True, False])
def test_cg_indefinite_and_alpha_exits(fused, monkeypatch):
    if not fused:
        monkeypatch.setenv("CB200_NO_FUSED_STEP", "1")
    m, p, values, b, D = small_problem(C=9, P=200, N=900, seed=13)
    N = p.N
    D = np.abs(D) + 0.5
    D[3 * p.P:] = 0.0  # no D_f: S is F'(I - E (E'E + D_e^2)^-1 E')F alone
    # S p underflows to exactly zero: p'q = 0 -> NO_CONVERGENCE "Matrix is indefinite, no more progress can
    # be made. p'q = 0.000000e+00 ..." (conjugate_gradients_solver.h:196-204)
    v = values.copy()
    v[6 * N:] *= 1e-170
    x, s = _solve_both(m, p.structure, v, b * 1e170, D, ob.IDENTITY)
    assert s["termination_type"] == 1 and "Matrix is indefinite" in s["message"], s
    # alpha = rho / pq overflows (S in the denormal range): FAILURE "Numerical failure. alpha = rho / pq = inf"
    v = values.copy()
    v[6 * N:] *= 1e-157
    x, s = _solve_both(m, p.structure, v, b * 1e156, D, ob.IDENTITY, sentinel=3.5)
    assert s["termination_type"] == 2 and "alpha = rho / pq" in s["message"], s
    assert np.all(x == 3.5)


def test_unfused_iteration_path_matches_oracle(monkeypatch):
    """Above kMaxStepBlocks * 14 = 57 344 cameras (or without cooperative launch) the PCG iteration runs as
    separate vector kernels instead of the fused step kernel: same iterates."""
    monkeypatch.setenv("CB200_NO_FUSED_STEP", "1")
    m, p, values, b, D = small_problem(C=9, P=200, N=900, seed=11)
    for ptype in ("IDENTITY", "JACOBI", "SCHUR_JACOBI"):
        compare_solve(m, p.structure, values, b, D, ptype)
    compare_solve(m, p.structure, values, b, D, "SCHUR_JACOBI", q_tol=-1.0, r_tol=-1.0, min_it=25, max_it=25)


def test_many_cameras_fused_path():
    """20 000 cameras (1429 CTAs of the cooperative step kernel, past the old 1024-CTA limit)."""
    m = util.pkg()
    p = m.SyntheticBAL(20000, 30000, 130000, seed=4, geometry=False)
    values, b, D = p.random_values(2)
    D = np.abs(D) + 1.0
    compare_solve(m, p.structure, values, b, D, "SCHUR_JACOBI", r_tol=1e-8, max_it=100, threads=8)


# ---- round-2 machinery against the plain path ----------------------------------------------------------------

def test_graph_replay_and_plain_launches_agree():
    """The launch sequence of a solve replayed from CUDA graphs (third and later solves of a handle) gives
    the iterates of plain stream launches (cuda_graphs = -1), for short and long solves."""
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug", scale=0.1)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    # (fixed iteration counts: a Q-test that sits on a borderline iteration may flip between two runs of the
    #  atomics-based default path, graphs or not -- that is what cb200_options::deterministic is for)
    for its, ptype in ((3, m.SCHUR_JACOBI), (27, m.SCHUR_JACOBI), (6, m.SCHUR_POWER_SERIES_EXPANSION)):
        plain = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=ptype, min_num_iterations=its,
                                                 max_num_iterations=its, cuda_graphs=-1)
        xp, sp = plain.solve(values, residuals, D, -1.0, -1.0)
        plain.close()
        g = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=ptype, min_num_iterations=its,
                                             max_num_iterations=its)
        for rep in range(4):
            xg, sg = g.solve(values, residuals, D, -1.0, -1.0)
            assert sg["num_iterations"] == sp["num_iterations"] == its and sg["message"] == sp["message"], rep
            # (two runs of the atomics-based scatter differ in the last bits and PCG amplifies that: ~1e-9 after 27
            #  iterations, run to run, with or without graphs)
            assert util.rel_err(xg, xp) < (1e-8 if its > 10 else 1e-11), rep
        # a different diagonal (graphs keyed on launch parameters must be rebuilt, not replayed stale)
        D2 = 2.0 * D
        xg, sg = g.solve(values, residuals, D2, -1.0, -1.0)
        g.clear_diagonal()
        g.upload(values, residuals, None)
        g.solve_resident(-1.0, -1.0)
        for rep in range(3):
            xg3, sg3 = g.solve(values, residuals, D2, -1.0, -1.0)
            assert util.rel_err(xg3, xg) < (1e-8 if its > 10 else 1e-11)
        assert util.rel_err(xg, xp) > 1e-3  # (and it IS a different problem)
        g.close()


def test_history_back_substitution_matches_the_pass_over_j(monkeypatch):
    """y_e = v_rhs - sum alpha_k v_k (solves of <= 8 iterations) against the kModeBack pass, and against the
    oracle, at 0, 1, few and more-than-eight iterations."""
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug", scale=0.1)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    st = p.structure
    o = ob.Oracle(util.oracle_structure(st), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=300)
    for q_tol, max_it in ((1e-1, 300), (1e-2, 300), (1e-6, 3), (1e-6, 8), (1e-6, 9), (1e-6, 0), (1e-6, 20)):
        o2 = ob.Oracle(util.oracle_structure(st), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI,
                       max_num_iterations=max_it)
        xo, so = o2.solve(values, residuals, D, q_tol, -1.0)
        g = m.IterativeSchurComplementSolver(st, max_num_iterations=max_it)
        for rep in range(3):  # plain, captured, replayed
            xh, sh = g.solve(values, residuals, D, q_tol, -1.0)
            assert sh["num_iterations"] == so["num_iterations"], (q_tol, max_it, sh, so)
            assert util.rel_err(xh, xo) < ITERATE_TOL, (q_tol, max_it, rep)
        g.close()
    monkeypatch.setenv("CB200_NO_BACKSUB_HISTORY", "1")
    g = m.IterativeSchurComplementSolver(st, max_num_iterations=300)
    xf, sf = g.solve(values, residuals, D, 1e-2, -1.0)
    g.close()
    monkeypatch.delenv("CB200_NO_BACKSUB_HISTORY")
    g = m.IterativeSchurComplementSolver(st, max_num_iterations=300)
    xh, sh = g.solve(values, residuals, D, 1e-2, -1.0)
    g.close()
    assert sh["num_iterations"] == sf["num_iterations"] <= 8
    assert util.rel_err(xh, xf) < 1e-12
    assert sh["kernel_launches"] >= 1


def test_chunked_upload_matches_one_copy():
    """cb200_solve with the Jacobian copied in 16 point-range chunks (E'E / rhs pass per chunk while the next
    chunk is in flight) against one monolithic copy, from pinned and from pageable memory; ragged structure
    (long and empty points) so that the chunk boundaries meet tile groups and the tail kernels."""
    m = util.pkg()
    rng = np.random.default_rng(31)
    C, P = 60, 6000
    lens = rng.choice([0, 1, 2, 3, 5, 9, 17, 40, 300], size=P, p=[.04, .2, .25, .2, .15, .1, .04, .015, .005])
    prs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    cam = np.concatenate([np.sort(rng.choice(C, size=min(L, C), replace=False)) if L <= C else
                          np.sort(rng.integers(0, C, size=L)) for L in lens if L > 0]).astype(np.int32)
    # (points of 300 rows see cameras repeatedly: duplicate pairs, fine for this test)
    st = m.BlockStructure.bal(prs, cam, C)
    N = len(cam)
    values, b = rng.standard_normal(24 * N), rng.standard_normal(2 * N)
    D = np.abs(rng.standard_normal(3 * P + 9 * C)) + 0.5
    one = m.IterativeSchurComplementSolver(st, max_num_iterations=100, h2d_chunks=1)
    x1, s1 = one.solve(values, b, D, -1.0, 1e-9)
    one.close()
    o = ob.Oracle(util.oracle_structure(st), num_threads=4, preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=100)
    xo, so = o.solve(values, b, D, -1.0, 1e-9)
    assert s1["num_iterations"] == so["num_iterations"] and util.rel_err(x1, xo) < ITERATE_TOL
    for chunks in (0, 7):
        g = m.IterativeSchurComplementSolver(st, max_num_iterations=100, h2d_chunks=chunks)
        for rep in range(3):
            xc, sc = g.solve(values, b, D, -1.0, 1e-9)
            assert sc["num_iterations"] == s1["num_iterations"], (chunks, rep)
            assert util.rel_err(xc, x1) < 1e-11, (chunks, rep)
        m.host_register(values)
        try:
            xc, sc = g.solve(values, b, D, -1.0, 1e-9)
        finally:
            m.host_unregister(values)
        assert util.rel_err(xc, x1) < 1e-11
        xn, sn = g.solve(values, b, 3.0 * D, -1.0, 1e-9)  # another diagonal through the same handle
        g.close()
        one = m.IterativeSchurComplementSolver(st, max_num_iterations=100, h2d_chunks=1)
        xn1, sn1 = one.solve(values, b, 3.0 * D, -1.0, 1e-9)
        one.close()
        assert sn["num_iterations"] == sn1["num_iterations"] and util.rel_err(xn, xn1) < 1e-10


@pytest.mark.parametrize("ptype", ["SCHUR_JACOBI", "SCHUR_POWER_SERIES_EXPANSION", "JACOBI"])
def test_deterministic_mode_is_bitwise_repeatable(ptype):
    """cb200_options::deterministic: the camera scatter summed in a fixed order -> two solves, and two
    handles, agree bit for bit (the default path's FP64 atomics do not), and both match the oracle."""
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug", scale=0.1, dup_pairs=6)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    code = ["IDENTITY", "JACOBI", "SCHUR_JACOBI", "SCHUR_POWER_SERIES_EXPANSION"].index(ptype)
    out = []
    for handle in range(2):
        g = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=code, max_num_iterations=300,
                                             deterministic=True)
        g.set_trace(16)
        for rep in range(3):
            x, s = g.solve(values, residuals, D, 1e-3, -1.0)
            out.append((x.copy(), s["num_iterations"], s["trace"].copy(), s["message"]))
        g.upload(values, residuals, D)
        g.isc_init()
        xv = np.random.default_rng(1).standard_normal(9 * p.C)
        out[-1] += (g.isc_multiply(xv), g.isc_rhs())
        g.close()
    for x, its, trace, msg in [o[:4] for o in out[1:]]:
        assert its == out[0][1] and msg == out[0][3]
        assert np.array_equal(x, out[0][0]) and np.array_equal(trace, out[0][2])
    assert np.array_equal(out[2][4], out[5][4]) and np.array_equal(out[2][5], out[5][5])
    compare_solve(m, p.structure, values, residuals, D, ptype, q_tol=1e-3, r_tol=-1.0, max_it=300, threads=8,
                  deterministic=True)
    o = ob.Oracle(util.oracle_structure(p.structure), num_threads=8, preconditioner_type=code)
    o.isc_init(values, D, residuals)
    assert util.rel_err(out[2][4], o.isc_multiply(np.random.default_rng(1).standard_normal(9 * p.C))) < OP_TOL
    assert util.rel_err(out[2][5], o.isc_rhs()) < OP_TOL


def test_device_evaluator_matches_independent_snavely_oracle():
    """(f)2: bal_eval_kernel against oracle/snavely_reference.py (numpy Jets written from
    snavely_reprojection_error.h:57-92 and rotation.h:792-850; shares no code with csrc/), including
    cameras with exactly-zero and tiny rotations (both branches of AngleAxisRotatePoint)."""
    from oracle import snavely_reference as sr
    m = util.pkg()
    p = m.SyntheticBAL(40, 3000, 14000, seed=3)
    cams = p.cameras.copy().reshape(-1, 9)
    cams[0, 0:3] = 0.0
    cams[1, 0:3] = [3e-9, -2e-9, 1e-9]
    cams = cams.reshape(-1)
    g = m.IterativeSchurComplementSolver(p.structure)
    g.bal_attach(p.observations)
    g.bal_set_parameters(cams, p.points)
    cost, residuals, values = sr.evaluate(p.point_row_start, p.cam_idx, p.observations, cams, p.points)
    gc, grad = g.bal_evaluate(jacobian=True)
    gv, gr = g.download_jacobian()
    assert abs(gc - cost) <= 1e-13 * cost
    assert util.rel_err(gr, residuals) < 1e-13
    tiny = np.isin(p.cam_idx, [1])
    E, F = gv[:6 * p.N].reshape(-1, 6), gv[6 * p.N:].reshape(-1, 18)
    Er, Fr = values[:6 * p.N].reshape(-1, 6), values[6 * p.N:].reshape(-1, 18)
    assert util.rel_err(E[~tiny], Er[~tiny]) < 1e-12 and util.rel_err(F[~tiny], Fr[~tiny]) < 1e-12
    assert util.rel_err(E[tiny], Er[tiny]) < 1e-7 and util.rel_err(F[tiny], Fr[tiny]) < 1e-7  # 1/theta cancellation
    rp = sr.ReferenceProblem.of(p)
    from oracle import lm_reference as lmr
    want = lmr.jacobian_transpose_times(rp, values, residuals)
    assert util.rel_err(grad, want) < 1e-9
    c2, _ = g.bal_evaluate(jacobian=False)
    assert abs(c2 - cost) <= 1e-13 * cost


def test_explicit_schur_lm_step_without_convergence():
    """with use_explicit_schur_complement a PCG solve that stops at max_num_iterations returns
    the step [0; cg_solution] (schur_complement_solver.cc:135-160) and the trust-region loop still
    evaluates it (trust_region_minimizer.cc:430-447): same accepted steps and costs as the CPU loop."""
    from oracle import lm_reference as lmr
    m = util.pkg()
    p = m.SyntheticBAL(31, 3000, 14000, seed=7, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    st = p.structure
    o = ob.Oracle(util.oracle_structure(st), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=3)
    want = lmr.minimize(p, p.cameras, p.points, lambda v, r, D, e: o.explicit_solve(v, r, D, e, -1.0), eta=1e-6,
                        max_num_iterations=6)
    g = m.IterativeSchurComplementSolver(st, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=3,
                                         use_explicit_schur_complement=True)
    g.bal_attach(p.observations)
    g.bal_set_parameters(p.cameras, p.points)
    got = g.lm_minimize(eta=1e-6, max_num_iterations=6)
    assert got["termination_type"] == want["termination_type"], (got["message"], want["message"])
    assert len(got["iterations"]) == len(want["iterations"])
    assert any(not b["step_is_successful"] or b["linear_solver_iterations"] == 3 for b in want["iterations"][1:])
    for a, b in zip(got["iterations"], want["iterations"]):
        assert bool(a["step_is_valid"]) == bool(b["step_is_valid"]), (a, b)
        assert bool(a["step_is_successful"]) == bool(b["step_is_successful"]), (a, b)
        assert a["linear_solver_iterations"] == b["linear_solver_iterations"]
        assert abs(a["cost"] - b["cost"]) <= 1e-8 * b["cost"], (a, b)
    assert abs(got["final_cost"] - want["final_cost"]) <= 1e-8 * want["final_cost"]
