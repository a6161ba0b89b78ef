"""CPU tests: pin the oracle (oracle/*.h, the CPU restatement of the reference's
ITERATIVE_SCHUR path) against the reference's own known-answer fixtures
(tests/golden/kat_problems.py, transcribed from linear_least_squares_problems.cc)
and against the identities the reference's tests assert:
  partitioned_matrix_view_test.cc:103-230, implicit_schur_complement_test.cc:55-237,
  iterative_schur_complement_solver_test.cc:60-149, schur_eliminator_test.cc,
  conjugate_gradients_solver_test.cc:47-160,
  power_series_expansion_preconditioner_test.cc:85-173.
"""
import numpy as np
import pytest

from oracle import binding as ob
from tests.golden import kat_problems as kp


def make(pid, **kw):
    st_kw, values, b, D, extra = kp.flat(pid)
    st = ob.Structure(**st_kw)
    return st, values, b, D, extra, ob.Oracle(st, **kw)


def dense_solution(A, b, D):
    H = A.T @ A + (np.diag(D * D) if D is not None else 0)
    return np.linalg.solve(H, A.T @ b)


def dense_schur(A, D, ne):
    H = A.T @ A + (np.diag(D * D) if D is not None else 0)
    Hee, Hef, Hff = H[:ne, :ne], H[:ne, ne:], H[ne:, ne:]
    return Hff - Hef.T @ np.linalg.solve(Hee, Hef)


@pytest.mark.parametrize("pid", [2, 4, 5, 6])
@pytest.mark.parametrize("threads", [1, 2, 4, 8])
def test_partitioned_view_products(pid, threads):
    st, values, b, D, _, o = make(pid, num_threads=threads)
    A = st.dense(values)
    ne = st.num_cols_e
    rng = np.random.default_rng(pid)
    xe, xf = rng.standard_normal(ne), rng.standard_normal(st.num_cols_f)
    xr = rng.standard_normal(st.num_rows)
    y = np.zeros(st.num_rows); o.pmv(values, 0, xe, y)
    np.testing.assert_allclose(y, A[:, :ne] @ xe, atol=1e-14)
    y = np.zeros(st.num_rows); o.pmv(values, 1, xf, y)
    np.testing.assert_allclose(y, A[:, ne:] @ xf, atol=1e-14)
    y = np.zeros(ne); o.pmv(values, 2, xr, y)
    np.testing.assert_allclose(y, A[:, :ne].T @ xr, atol=1e-14)
    y = np.zeros(st.num_cols_f); o.pmv(values, 3, xr, y)
    np.testing.assert_allclose(y, A[:, ne:].T @ xr, atol=1e-14)


@pytest.mark.parametrize("pid", [2, 4, 5, 6])
@pytest.mark.parametrize("threads", [1, 4])
def test_block_diagonals(pid, threads):
    st, values, b, D, _, o = make(pid, num_threads=threads)
    A = st.dense(values)
    for which, lo, hi in [(0, 0, st.num_eliminate_blocks),
                          (1, st.num_eliminate_blocks, len(st.col_size))]:
        got = o.block_diagonal(values, which)
        H = A.T @ A
        k = 0
        for cbi in range(lo, hi):
            s, p = st.col_size[cbi], st.col_pos[cbi]
            np.testing.assert_allclose(got[k:k + s * s].reshape(s, s),
                                       H[p:p + s, p:p + s], atol=1e-14)
            k += s * s
        assert k == len(got)


@pytest.mark.parametrize("pid", [2, 4, 5, 6])
@pytest.mark.parametrize("use_D", [False, True])
@pytest.mark.parametrize("threads", [1, 3])
def test_implicit_schur_complement(pid, use_D, threads):
    if pid in (4, 6) and not use_D:
        pytest.skip("rank deficient without regularisation (reference note :543)")
    st, values, b, D, _, o = make(pid, num_threads=threads,
                                  preconditioner_type=ob.JACOBI)
    Dv = D if use_D else None
    A = st.dense(values)
    ne, nf = st.num_cols_e, st.num_cols_f
    S = dense_schur(A, Dv, ne)
    H = A.T @ A + (np.diag(D * D) if use_D else 0)
    g = A.T @ b
    rhs = g[ne:] - H[:ne, ne:].T @ np.linalg.solve(H[:ne, :ne], g[:ne])
    o.isc_init(values, Dv, b)
    scale = max(1.0, np.abs(S).max())
    for i in range(nf):
        e = np.zeros(nf); e[i] = 1.0
        np.testing.assert_allclose(o.isc_multiply(e), S[:, i], atol=1e-12 * scale)
    np.testing.assert_allclose(o.isc_rhs(), rhs, atol=1e-12 * max(1, np.abs(rhs).max()))
    z = np.linalg.solve(S, rhs)
    sol = dense_solution(A, b, Dv)
    np.testing.assert_allclose(o.isc_back_substitute(z), sol, atol=1e-12)
    # InversePowerSeriesOperator == (F'F)^-1 F'E (E'E)^-1 E'F
    Z = np.linalg.solve(H[ne:, ne:] * 0 + _blockdiag(H[ne:, ne:], st), H[ne:, ne:] * 0 +
                        H[:ne, ne:].T @ np.linalg.solve(_blockdiag_e(H[:ne, :ne], st), H[:ne, ne:]))
    for i in range(nf):
        e = np.zeros(nf); e[i] = 1.0
        np.testing.assert_allclose(o.isc_pse_operator(e, np.zeros(nf)), Z[:, i],
                                   atol=1e-12)


def _blockdiag(Hff, st):
    out = np.zeros_like(Hff)
    base = st.num_cols_e
    for cbi in range(st.num_eliminate_blocks, len(st.col_size)):
        s, p = st.col_size[cbi], st.col_pos[cbi] - base
        out[p:p + s, p:p + s] = Hff[p:p + s, p:p + s]
    return out


def _blockdiag_e(Hee, st):
    out = np.zeros_like(Hee)
    for cbi in range(st.num_eliminate_blocks):
        s, p = st.col_size[cbi], st.col_pos[cbi]
        out[p:p + s, p:p + s] = Hee[p:p + s, p:p + s]
    return out


def test_problem2_published_solution():
    """linear_least_squares_problems.cc:311-316 gives x = A\\b to 4 decimals
    (the unregularised solution, D = nullptr)."""
    st, values, b, D, extra, o = make(2, max_num_iterations=5)
    x, s = o.solve(values, b, None, r_tolerance=1e-12)
    assert ob.TERMINATION[s["termination_type"]] == "SUCCESS"
    np.testing.assert_allclose(x, extra["x"], atol=5e-5)


def test_problem5_published_schur_complement():
    """:628-676 gives S, r and S\\r (no regularisation) to 4 decimals."""
    st, values, b, D, extra, o = make(5, preconditioner_type=ob.JACOBI)
    o.isc_init(values, None, b)
    nf = st.num_cols_f
    S = np.stack([o.isc_multiply(np.eye(nf)[i]) for i in range(nf)], axis=1)
    np.testing.assert_allclose(S, extra["S"], atol=5e-5)
    np.testing.assert_allclose(o.isc_rhs(), extra["r"], atol=5e-5)
    np.testing.assert_allclose(np.linalg.solve(S, o.isc_rhs()), extra["S_solve_r"],
                               atol=5e-5)


# iterative_schur_complement_solver_test.cc:125-149
@pytest.mark.parametrize("pid,ptype,spse_init", [
    (2, ob.SCHUR_JACOBI, False), (2, ob.SCHUR_JACOBI, True),
    (5, ob.SCHUR_POWER_SERIES_EXPANSION, False), (3, ob.SCHUR_JACOBI, False),
    (2, ob.JACOBI, False), (2, ob.IDENTITY, False)])
@pytest.mark.parametrize("use_D", [False, True])
@pytest.mark.parametrize("threads", [1, 4])
def test_iterative_schur_solver_vs_dense(pid, ptype, spse_init, use_D, threads):
    st, values, b, D, _, o = make(
        pid, num_threads=threads, preconditioner_type=ptype,
        max_num_iterations=st_cols(pid), max_num_spse_iterations=1,
        use_spse_initialization=spse_init)
    Dv = D if use_D else None
    x, s = o.solve(values, b, Dv, q_tolerance=0.0, r_tolerance=1e-12)
    ref = dense_solution(st.dense(values), b, Dv)
    assert np.linalg.norm(x - ref) < 1e-13 * max(1.0, np.linalg.norm(ref)), s


def st_cols(pid):
    return sum(c[0] for c in kp.PROBLEMS[pid]["cols"])


@pytest.mark.parametrize("pid", [2, 4, 5, 6])
@pytest.mark.parametrize("use_D", [False, True])
@pytest.mark.parametrize("threads", [1, 4])
def test_schur_jacobi_blocks(pid, use_D, threads):
    if pid in (4, 6) and not use_D:
        pytest.skip("rank deficient without regularisation")
    st, values, b, D, _, o = make(pid, num_threads=threads)
    Dv = D if use_D else None
    S = dense_schur(st.dense(values), Dv, st.num_cols_e)
    got = o.schur_jacobi(values, Dv, invert=False)
    inv = o.schur_jacobi(values, Dv, invert=True)
    k, base = 0, st.num_cols_e
    for cbi in range(st.num_eliminate_blocks, len(st.col_size)):
        s, p = st.col_size[cbi], st.col_pos[cbi] - base
        blk = S[p:p + s, p:p + s]
        np.testing.assert_allclose(got[k:k + s * s].reshape(s, s), blk,
                                   rtol=1e-12, atol=1e-12 * np.abs(blk).max())
        np.testing.assert_allclose(inv[k:k + s * s].reshape(s, s), np.linalg.inv(blk),
                                   rtol=1e-10, atol=1e-12)
        k += s * s


def test_power_series_converges_to_inverse():
    """power_series_expansion_preconditioner_test.cc:85-128: on problem 5 (F'F is
    block diagonal) enough series terms reproduce S^-1."""
    st, values, b, D, _, o = make(5, preconditioner_type=ob.SCHUR_POWER_SERIES_EXPANSION)
    o.isc_init(values, D, b)
    S = dense_schur(st.dense(values), D, st.num_cols_e)
    nf = st.num_cols_f
    for i in range(nf):
        y = o.pse_apply(np.eye(nf)[i], 100, 1e-30)
        np.testing.assert_allclose(y, np.linalg.inv(S)[:, i], atol=1e-12)
    # one term only is NOT the inverse (:130-173)
    y = o.pse_apply(np.eye(nf)[0], 1, 1e-30)
    assert np.abs(y - np.linalg.inv(S)[:, 0]).max() > 1e-4


def test_cg_identity_and_tridiagonal():
    """conjugate_gradients_solver_test.cc:47-160."""
    x, s = ob.cg_dense(np.eye(3), [1, 2, 3], [1, 1, 1], 1, 10, 20, 0.0, 1e-9)
    assert ob.TERMINATION[s["termination_type"]] == "SUCCESS"
    assert s["num_iterations"] == 1
    np.testing.assert_allclose(x, [1, 2, 3], rtol=4e-16)
    A = [[2, -1, 0], [-1, 2, -1], [0, -1, 2]]
    x, s = ob.cg_dense(A, [-1, 0, 3], [1, 1, 1], 1, 10, 20, 0.0, 1e-9)
    assert ob.TERMINATION[s["termination_type"]] == "SUCCESS"
    np.testing.assert_allclose(x, [0, 1, 2], atol=1e-14)


def test_llt_inverse_uses_upper_triangle_only():
    rng = np.random.default_rng(0)
    for n in (1, 2, 3, 9):
        M = rng.standard_normal((n + 3, n))
        A = M.T @ M + np.eye(n)
        m = np.triu(A) + np.tril(rng.standard_normal((n, n)), -1)  # garbage below
        m = np.ascontiguousarray(m)
        ob.lib().oracle_llt_inverse_upper(ob._dp(m), n)
        np.testing.assert_allclose(m, np.linalg.inv(A), rtol=1e-11, atol=1e-13)
        out = np.zeros((n, n))
        ob.lib().oracle_inverse_small(ob._dp(np.ascontiguousarray(A)), n, ob._dp(out))
        np.testing.assert_allclose(out, np.linalg.inv(A), rtol=1e-11, atol=1e-13)


# ---- use_explicit_schur_complement = true (oracle/explicit_schur.h) ----------
def _dense_from_upper(row_ptr, cols, pos, vals, sizes):
    """Dense symmetric matrix of a BlockRandomAccessSparseMatrix (upper block triangle)."""
    off = np.concatenate([[0], np.cumsum(sizes)])
    S = np.zeros((off[-1], off[-1]))
    for i in range(len(sizes)):
        for k in range(row_ptr[i], row_ptr[i + 1]):
            j = cols[k]
            blk = vals[pos[k]:pos[k] + sizes[i] * sizes[j]].reshape(sizes[i], sizes[j])
            S[off[i]:off[i + 1], off[j]:off[j + 1]] = blk
            if i != j:
                S[off[j]:off[j + 1], off[i]:off[i + 1]] = blk.T
    return S


# schur_eliminator_test.cc:127-226 (EliminateSolveAndCompare): lhs, rhs and the back
# substituted solution against the dense Schur complement, with and without D.
@pytest.mark.parametrize("pid", [2, 4, 5, 6])
@pytest.mark.parametrize("use_D", [False, True])
@pytest.mark.parametrize("threads", [1, 4])
def test_explicit_schur_eliminate_vs_dense(pid, use_D, threads):
    if pid in (4, 6) and not use_D:
        pytest.skip("rank deficient without regularisation")
    st, values, b, D, _, o = make(pid, num_threads=threads)
    Dv = D if use_D else None
    A = st.dense(values)
    ne = st.num_cols_e
    S = dense_schur(A, Dv, ne)
    H = A.T @ A + (np.diag(D * D) if use_D else 0)
    g = A.T @ b
    rhs_ref = g[ne:] - H[:ne, ne:].T @ np.linalg.solve(H[:ne, :ne], g[:ne])
    row_ptr, cols, pos, vals, rhs = o.explicit_eliminate(values, b, Dv)
    sizes = st.col_size[st.num_eliminate_blocks:]
    got = _dense_from_upper(row_ptr, cols, pos, vals, sizes)
    # only structurally non-zero pairs are stored; the dense S is zero elsewhere
    np.testing.assert_allclose(got, S, atol=1e-12 * max(1.0, np.abs(S).max()))
    np.testing.assert_allclose(rhs, rhs_ref, atol=1e-12 * max(1.0, np.abs(rhs_ref).max()))
    # block_random_access_sparse_matrix_test.cc:126-141: symmetric multiply == dense
    x = np.random.default_rng(pid).standard_normal(st.num_cols_f)
    np.testing.assert_allclose(o.explicit_multiply(x), S @ x, atol=1e-12 * max(1.0, np.abs(S).max()))
    assert all(cols[k] >= i for i in range(len(sizes)) for k in range(row_ptr[i], row_ptr[i + 1]))


# solver_test.cc:1107-1140 allows ITERATIVE_SCHUR + use_explicit_schur_complement only with
# SCHUR_JACOBI; the solve must agree with the dense normal equations like the implicit one.
@pytest.mark.parametrize("pid", [2, 3, 4, 5, 6])
@pytest.mark.parametrize("use_D", [False, True])
@pytest.mark.parametrize("threads", [1, 4])
def test_explicit_schur_solver_vs_dense_and_implicit(pid, use_D, threads):
    if pid in (4, 6) and not use_D:
        pytest.skip("rank deficient without regularisation")
    st, values, b, D, _, o = make(pid, num_threads=threads, preconditioner_type=ob.SCHUR_JACOBI,
                                  max_num_iterations=st_cols(pid))
    Dv = D if use_D else None
    x, s = o.explicit_solve(values, b, Dv, q_tolerance=0.0, r_tolerance=1e-12)
    ref = dense_solution(st.dense(values), b, Dv)
    assert ob.TERMINATION[s["termination_type"]] == "SUCCESS", s
    assert np.linalg.norm(x - ref) < 1e-12 * max(1.0, np.linalg.norm(ref)), s
    xi, si = o.solve(values, b, Dv, q_tolerance=0.0, r_tolerance=1e-12)
    assert si["num_iterations"] == s["num_iterations"]
    np.testing.assert_allclose(x, xi, atol=1e-12 * max(1.0, np.abs(ref).max()))


def test_explicit_schur_no_back_substitution_without_success():
    """schur_complement_solver.cc:155-160: x keeps its zero point part unless CG succeeded."""
    st, values, b, D, _, o = make(2, preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=1)
    x, s = o.explicit_solve(values, b, D, q_tolerance=0.0, r_tolerance=1e-30)
    assert ob.TERMINATION[s["termination_type"]] == "NO_CONVERGENCE"
    assert np.all(x[:st.num_cols_e] == 0.0) and np.any(x[st.num_cols_e:] != 0.0)


def _random_schur_problem(seed, e_size, f_sizes, row_size, num_e=7, trailing_rows=3):
    """A random block structure in the shape schur_eliminator_test.cc:228-330 builds by hand
    (static or mixed block sizes, rows with one e cell and 1..3 f cells, and rows WITHOUT an e
    cell at the bottom: NoEBlockRowsUpdate, schur_eliminator_impl.h:566-593)."""
    rng = np.random.default_rng(seed)
    nf = len(f_sizes)
    col_size = [e_size] * num_e + list(f_sizes)
    col_pos = np.concatenate([[0], np.cumsum(col_size)[:-1]])
    row_sz, row_pos, cell_ptr, cid, cpos = [], [], [0], [], []
    pos, rp = 0, 0
    for e in range(num_e):
        for _ in range(int(rng.integers(1, 5))):
            rs = row_size if row_size > 0 else int(rng.integers(1, 4))
            fs = sorted(rng.choice(nf, size=int(rng.integers(1, min(3, nf) + 1)), replace=False))
            row_sz.append(rs); row_pos.append(rp); rp += rs
            for b in [e] + [num_e + int(f) for f in fs]:
                cid.append(b); cpos.append(pos); pos += rs * col_size[b]
            cell_ptr.append(len(cid))
    for _ in range(trailing_rows):
        rs = row_size if row_size > 0 else int(rng.integers(1, 4))
        fs = sorted(rng.choice(nf, size=int(rng.integers(1, min(3, nf) + 1)), replace=False))
        row_sz.append(rs); row_pos.append(rp); rp += rs
        for f in fs:
            cid.append(num_e + int(f)); cpos.append(pos); pos += rs * col_size[num_e + int(f)]
        cell_ptr.append(len(cid))
    st = ob.Structure(row_sz, row_pos, cell_ptr, cid, cpos, col_size, col_pos, num_e)
    values = rng.standard_normal(pos)
    b = rng.standard_normal(rp)
    D = np.abs(rng.standard_normal(int(sum(col_size)))) + 0.5
    return st, values, b, D


@pytest.mark.parametrize("shape", [
    (3, [9] * 6, 2),            # the bundle-adjustment structure <2,3,9>
    (3, [3, 4, 9, 2, 6], 2),    # varying f-block size (schur_eliminator_test.cc:262-330)
    (2, [1, 5, 3, 3], 0),       # every size dynamic
    (5, [2, 2, 2], 3),          # e blocks of 5: InvertPSDMatrix takes the Cholesky route (invert_psd_matrix.h:60-72)
])
@pytest.mark.parametrize("threads", [1, 4])
def test_explicit_and_implicit_oracle_on_random_block_structures(shape, threads):
    e_size, f_sizes, row_size = shape
    for seed in range(4):
        st, values, b, D = _random_schur_problem(seed, e_size, f_sizes, row_size)
        A = st.dense(values)
        ne = st.num_cols_e
        S = dense_schur(A, D, ne)
        ref = dense_solution(A, b, D)
        o = ob.Oracle(st, num_threads=threads, preconditioner_type=ob.SCHUR_JACOBI,
                      max_num_iterations=4 * st.num_cols_f)
        row_ptr, cols, pos, vals, rhs = o.explicit_eliminate(values, b, D)
        got = _dense_from_upper(row_ptr, cols, pos, vals, st.col_size[st.num_eliminate_blocks:])
        # pairs that no row produces are not stored; the dense S is exactly zero there
        np.testing.assert_allclose(got, S, atol=1e-12 * np.abs(S).max())
        x, s = o.explicit_solve(values, b, D, q_tolerance=-1.0, r_tolerance=1e-13)
        assert ob.TERMINATION[s["termination_type"]] == "SUCCESS", s
        assert np.linalg.norm(x - ref) < 1e-10 * np.linalg.norm(ref)
        xi, si = o.solve(values, b, D, q_tolerance=-1.0, r_tolerance=1e-13)
        assert np.linalg.norm(xi - ref) < 1e-10 * np.linalg.norm(ref)
        o.isc_init(values, D, b)
        v = np.random.default_rng(seed).standard_normal(st.num_cols_f)
        np.testing.assert_allclose(o.isc_multiply(v), S @ v, atol=1e-12 * np.abs(S).max() * np.abs(v).sum())


# ---- the independent Snavely / Jet restatement (oracle/snavely_reference.py) --------------------------
# Pins the evaluator oracle WITHOUT any product code: hand-computed projections, complex-step derivatives
# of the plain residual, and the exact-zero switch of AngleAxisRotatePoint (rotation.h:799-801).

def test_snavely_reference_known_answers():
    from oracle import snavely_reference as sr
    # aa = 0 (Taylor branch, R = I): p = X + t = (0.1, 0.2, -2); xp = 0.05, yp = 0.1, r2 = 0.0125;
    # distortion = 1 + 0.0125 (0.1 + 0.01 * 0.0125) = 1.0012515625; f = 100 -> (5.0062578125, 10.012515625)
    cam = np.array([[0, 0, 0, 0, 0, 0, 100.0, 0.1, 0.01]])
    rx, ry = sr.snavely_residual_jets(cam, np.array([[0.1, 0.2, -2.0]]), np.zeros((1, 2)))
    assert abs(rx.a[0] - 5.0062578125) < 1e-14 and abs(ry.a[0] - 10.012515625) < 1e-14
    # rotation by pi/2 about z maps (0.2, -0.1, -2) to (0.1, 0.2, -2): the same image point
    cam2 = cam.copy()
    cam2[0, 2] = np.pi / 2
    rx2, ry2 = sr.snavely_residual_jets(cam2, np.array([[0.2, -0.1, -2.0]]), np.zeros((1, 2)))
    assert abs(rx2.a[0] - 5.0062578125) < 1e-13 and abs(ry2.a[0] - 10.012515625) < 1e-13
    # translation and observation enter linearly: t_z = -1 with X_z = -1 gives the same depth
    cam3 = cam.copy()
    cam3[0, 5] = -1.0
    rx3, ry3 = sr.snavely_residual_jets(cam3, np.array([[0.1, 0.2, -1.0]]), np.array([[5.0, 10.0]]))
    assert abs(rx3.a[0] - 0.0062578125) < 1e-13 and abs(ry3.a[0] - 0.012515625) < 1e-13
    # d pred_x / d focal = distortion * xp, by hand
    assert abs(rx.v[0, 6] - 1.0012515625 * 0.05) < 1e-15


def test_snavely_reference_jets_match_complex_step():
    from oracle import snavely_reference as sr
    rng = np.random.default_rng(5)
    n = 200
    cam = np.concatenate([0.3 * rng.standard_normal((n, 3)), rng.standard_normal((n, 3)),
                          rng.uniform(500, 1500, (n, 1)), 1e-2 * rng.standard_normal((n, 2))], axis=1)
    pt = rng.standard_normal((n, 3))
    pt[:, 2] -= 6.0
    obs = rng.standard_normal((n, 2))
    rx, ry = sr.snavely_residual_jets(cam, pt, obs)
    plain = sr.snavely_residual_plain(cam, pt, obs)
    assert np.abs(plain[:, 0] - rx.a).max() < 1e-10 and np.abs(plain[:, 1] - ry.a).max() < 1e-10
    h = 1e-30
    for k in range(12):
        c, x = cam.astype(complex), pt.astype(complex)
        if k < 9:
            c[:, k] += 1j * h
        else:
            x[:, k - 9] += 1j * h
        d = sr.snavely_residual_plain(c, x, obs).imag / h
        scale = max(1.0, np.abs(d).max())
        assert np.abs(d[:, 0] - rx.v[:, k]).max() < 1e-11 * scale, k
        assert np.abs(d[:, 1] - ry.v[:, k]).max() < 1e-11 * scale, k


def test_snavely_reference_zero_rotation_branch_is_exact_zero_only():
    """rotation.h:799-801 switches on fpclassify(theta) != FP_ZERO: |aa| = 1e-9 still takes Rodrigues."""
    from oracle import snavely_reference as sr
    base = np.array([[0, 0, 0, 0.1, -0.2, 0.3, 800.0, 1e-3, 1e-4]])
    pt, obs = np.array([[0.4, -0.3, -5.0]]), np.zeros((1, 2))
    rx0, ry0 = sr.snavely_residual_jets(base, pt, obs)
    assert np.all(np.isfinite(rx0.v)) and np.all(np.isfinite(ry0.v))
    # at exactly zero: d p / d aa = -hat(pt) (first-order branch)
    tiny = base.copy()
    tiny[0, 0:3] = [6e-10, -5e-10, 4e-10]
    rx1, ry1 = sr.snavely_residual_jets(tiny, pt, obs)
    assert abs(rx1.a[0] - rx0.a[0]) < 1e-6 and np.abs(rx1.v - rx0.v).max() < 1e-5 * np.abs(rx0.v).max()
    c, x = tiny.astype(complex), pt.astype(complex)
    c[0, 1] += 1e-30j
    d = sr.snavely_residual_plain(c, x, obs).imag / 1e-30
    assert abs(d[0, 0] - rx1.v[0, 1]) < 1e-6 * abs(d[0, 0])


@pytest.mark.parametrize("zero_cameras", [False, True])
def test_host_evaluator_matches_snavely_reference(zero_cameras):
    """The product's host evaluator (csrc/synthetic_bal.cc + csrc/snavely_model.h, the code the device
    kernel shares) against the independent restatement: residuals, cost and every Jacobian block."""
    from oracle import snavely_reference as sr
    from tests import util
    m = util.pkg()
    p = m.SyntheticBAL(12, 400, 1700, seed=9)
    cams = p.cameras.copy()
    if zero_cameras:  # exact-zero and tiny rotations: both branches of AngleAxisRotatePoint
        c = cams.reshape(-1, 9)
        c[0, 0:3] = 0.0
        c[1, 0:3] = [3e-9, -2e-9, 1e-9]
    cost, residuals, values = p.evaluate(cameras=cams)
    rcost, rres, rvalues = sr.evaluate(p.point_row_start, p.cam_idx, p.observations, cams, p.points)
    assert abs(cost - rcost) <= 1e-13 * rcost
    assert util.rel_err(residuals, rres) < 1e-13
    # per-block comparison: tiny-rotation blocks carry 1/theta cancellation, so compare column-wise at 1e-7
    assert util.rel_err(values, rvalues) < (1e-7 if zero_cameras else 1e-12)
    c2, r2, _ = sr.evaluate(p.point_row_start, p.cam_idx, p.observations, cams, p.points, jacobian=False)
    assert abs(c2 - rcost) <= 1e-13 * rcost and util.rel_err(r2, rres) < 1e-13
