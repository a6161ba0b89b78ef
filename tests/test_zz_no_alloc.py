"""Collected LAST (file name): the allocation-free entry point cb200_block_diagonal_apply_static (operands staged
in a file-scope static device array; no handle, no cudaMalloc) against the existing entry point
cb200_schur_jacobi_apply, through the C ABI, on two calls in a row (the array keeps the first call's contents).
Passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from tests import util

OP_TOL = 1e-12


@pytest.mark.gpu
def test_static_scratch_entry_point_matches_schur_jacobi_apply():
    m = util.pkg()
    rng = np.random.default_rng(17)
    for C, P, N, seed in ((40, 900, 4000, 1), (7, 60, 260, 2)):  # the second, smaller call runs over the first one's leftovers
        p = m.SyntheticBAL(C, P, N, seed=seed, geometry=False)
        values, b, D = p.random_values(seed)
        D = np.abs(D) + 1.0
        g = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=m.SCHUR_JACOBI)
        g.upload(values, b, D)
        blocks = g.schur_jacobi(invert=True)
        x, y0 = rng.standard_normal(9 * C), rng.standard_normal(9 * C)
        want = g.schur_jacobi_apply(x, y0.copy())
        g.close()
        for rep in range(2):
            got = m.block_diagonal_apply_static(blocks, x, y0.copy())
            assert util.rel_err(got, want) < OP_TOL, (C, rep)
    # sizes beyond the static array are rejected with a clear message, nothing is touched
    n = 16385
    y = np.zeros(9 * n)
    with pytest.raises(m.B200Error, match="exceeds the static scratch"):
        m.block_diagonal_apply_static(np.zeros(81 * n), np.zeros(9 * n), y)
    assert not y.any()


def test_static_scratch_entry_point_is_exported():
    m = util.pkg()
    assert hasattr(m.solver.lib(), "cb200_block_diagonal_apply_static")
