#!/usr/bin/env python
"""Every kind of launch the library makes, once, on a small problem: the workload for
`compute-sanitizer --tool memcheck` (no torch in the process: ctypes + numpy only).

    compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_sweep.py

(compute-sanitizer is closed on the pool this round ran on; the script was run plain there, exit 0:
profiles/r2_launch_sweep_plain.log.)

Covers: solve from host buffers (chunked upload, plain launches, then graph capture and replay), the four
preconditioners, the seam operators (E'E inverse, rhs, S*x, back substitution, PSE operator, block Jacobi),
points longer than one tile and longer than 256 rows, duplicate (point, camera) pairs, deterministic mode,
the explicit Schur complement, the device evaluator + LM loop, and the allocation-free entry point.
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def lm_like(p):
    cost, b, values = p.evaluate()
    scale = 1.0 / (1.0 + np.sqrt(p.squared_column_norm(values)))
    p.scale_columns(values, scale)
    D = np.sqrt(np.clip(p.squared_column_norm(values), 1e-6, 1e32) / 1e4)
    return values, b, D


def main():
    pkg = importlib.import_module("ceres-solver-ceres-solver_b200")
    p = pkg.SyntheticBAL(31, 3000, 14000, seed=7, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0,
                         dup_pairs=4)
    values, b, D = lm_like(p)
    st = p.structure
    rng = np.random.default_rng(0)
    xv = rng.standard_normal(st.num_cols_f)
    ref = None
    for code in (pkg.SCHUR_JACOBI, pkg.SCHUR_POWER_SERIES_EXPANSION, pkg.JACOBI, pkg.IDENTITY):
        for det in (False, True):
            g = pkg.IterativeSchurComplementSolver(st, preconditioner_type=code, max_num_iterations=60,
                                                   deterministic=det)
            for rep in range(3):  # plain, captured, replayed
                x, s = g.solve(values, b, D, 1e-3, -1.0)
            g.upload(values, b, D)
            g.isc_init()
            g.ete_inverse(); g.isc_rhs(); g.isc_multiply(xv); g.isc_back_substitute(xv)
            y = np.zeros_like(xv)
            if code == pkg.SCHUR_JACOBI:
                g.schur_jacobi(invert=True)
                g.schur_jacobi_apply(xv, y)
            if code == pkg.SCHUR_POWER_SERIES_EXPANSION:
                g.isc_pse_operator(xv, y)
                g.pse_apply(xv, 5, 0.0)
            g.close()
            if code == pkg.SCHUR_JACOBI and not det:
                ref = x.copy()
            print(f"precond {code} deterministic {det}: {s['num_iterations']} iterations, |x| {np.linalg.norm(x):.6e}",
                  flush=True)
    # long points: one tile < rows <= 256 (chunk tiles of the stream kernel), rows > 256 (point_long_kernel),
    # single-row and empty points
    C2 = 300
    lens = np.array([3, 1, 0, 300, 1, 1, 257, 100, 0, 33, 64, 256, 2, 5, 31, 32] + [1] * 40 + [3] * 30)
    prs = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    cam = np.concatenate([np.sort(rng.choice(C2, size=L, replace=False)) for L in lens if L > 0]).astype(np.int32)
    st2 = pkg.BlockStructure.bal(prs, cam, C2)
    N2, P2 = len(cam), len(lens)
    v2, b2 = rng.standard_normal(24 * N2), rng.standard_normal(2 * N2)
    D2 = np.abs(rng.standard_normal(3 * P2 + 9 * C2)) + 0.5
    for code in (pkg.SCHUR_JACOBI, pkg.SCHUR_POWER_SERIES_EXPANSION):
        for det in (False, True):
            g = pkg.IterativeSchurComplementSolver(st2, preconditioner_type=code, max_num_iterations=40, deterministic=det)
            for rep in range(3):
                x2, s2 = g.solve(v2, b2, D2, -1.0, 1e-8)
            g.close()
            print(f"ragged / long points, precond {code} deterministic {det}: {s2['num_iterations']} iterations", flush=True)
    # explicit Schur complement
    g = pkg.IterativeSchurComplementSolver(st, max_num_iterations=60, use_explicit_schur_complement=True)
    xe, se = g.solve(values, b, D, 1e-3, -1.0)
    g.close()
    print(f"explicit: {se['num_iterations']} iterations, rel diff to implicit {np.linalg.norm(xe - ref) / np.linalg.norm(ref):.2e}",
          flush=True)
    # device evaluator + LM loop
    g = pkg.IterativeSchurComplementSolver(st, max_num_iterations=100)
    g.bal_attach(p.observations)
    g.bal_set_parameters(p.cameras, p.points)
    out = g.lm_minimize(eta=1e-2, max_num_iterations=3)
    g.close()
    print(f"lm: {len(out['iterations'])} iterations, cost {out['final_cost']:.6e}", flush=True)
    # allocation-free entry point
    nb = 257
    blocks = rng.standard_normal((nb, 81))
    xx = rng.standard_normal(nb * 9)
    yy = np.zeros(nb * 9)
    pkg.block_diagonal_apply_static(blocks, xx, yy)
    want = np.einsum("nij,nj->ni", blocks.reshape(nb, 9, 9), xx.reshape(nb, 9)).reshape(-1)
    assert np.allclose(yy, want, rtol=1e-13, atol=1e-13)
    print("static entry point ok", flush=True)


if __name__ == "__main__":
    main()
