#!/usr/bin/env python
"""use_explicit_schur_complement against the implicit operator on one LM-shaped linear solve
(SURVEY 8(f) row 3): size of the explicit reduced camera matrix, time to form it
(SchurEliminator::Eliminate), time of one S x on it against the fused implicit S x, and the two
complete solves (same iteration counts, solutions compared).

    python tools/explicit_schur_benchmark.py --shape ladybug [--eta 1e-2] [--out profiles/...json]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import util  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="ladybug")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--eta", type=float, default=1e-2)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    pkg = importlib.import_module("ceres-solver-ceres-solver_b200")
    p = pkg.SyntheticBAL.named(args.shape, scale=args.scale)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    N, P, C = p.N, p.P, p.C
    t0 = time.perf_counter()
    ge = pkg.IterativeSchurComplementSolver(p.structure, preconditioner_type=pkg.SCHUR_JACOBI, max_num_iterations=500,
                                            use_explicit_schur_complement=True)
    create_explicit_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    gi = pkg.IterativeSchurComplementSolver(p.structure, preconditioner_type=pkg.SCHUR_JACOBI, max_num_iterations=500)
    create_implicit_s = time.perf_counter() - t0
    row_ptr, cols = pkg.explicit_plan(p.structure)
    lens = np.diff(p.point_row_start).astype(np.int64)
    pairs = int((lens * (lens + 1) // 2).sum())
    cells = len(cols)
    out = {"shape": args.shape, "C": C, "P": P, "N": N, "eta": args.eta,
           "upper_cells": cells, "density_of_upper_triangle": cells / (C * (C + 1) / 2),
           "S_bytes": cells * 648, "J_bytes": 192 * N, "outer_products": pairs,
           "create_s": {"explicit": create_explicit_s, "implicit": create_implicit_s}}
    for g in (ge, gi):
        g.upload(values, residuals, D)
        g.isc_init()
    ms_elim, _ = ge.time_op(13, warmup=1, iters=args.iters)
    ms_form, _ = ge.time_op(14, warmup=1, iters=args.iters)
    ms_spmv, _ = ge.time_op(15, warmup=2, iters=4 * args.iters)
    ms_sx, _ = gi.time_op(6, warmup=2, iters=4 * args.iters)
    ms_setup_implicit, _ = gi.time_op(1, warmup=1, iters=args.iters)
    ms_sj, _ = gi.time_op(3, warmup=1, iters=args.iters)
    out["ms"] = {"eliminate_total": ms_elim, "outer_product_kernel": ms_form, "explicit_Sx": ms_spmv,
                 "implicit_Sx": ms_sx, "implicit_setup_pass": ms_setup_implicit, "implicit_schur_jacobi": ms_sj}
    # bytes: every cell read twice (row list + column list), 81 RED.F64 per outer product
    out["explicit_Sx_GBs"] = (2 * cells - C) * 648 / ms_spmv / 1e6
    out["outer_product_Gred_per_s"] = pairs * 81 / ms_form / 1e6
    se = ge.solve_resident(args.eta, -1.0)
    si = gi.solve_resident(args.eta, -1.0)
    reps = 5
    te = ti = 0.0
    for _ in range(reps):
        s = ge.solve_resident(args.eta, -1.0)
        te += s["setup_ms"] + s["pcg_ms"] + s["back_substitute_ms"]
        s = gi.solve_resident(args.eta, -1.0)
        ti += s["setup_ms"] + s["pcg_ms"] + s["back_substitute_ms"]
    xe, xi = ge.download_solution(), gi.download_solution()
    out["solve"] = {"explicit": {k: se[k] for k in ("num_iterations", "setup_ms", "pcg_ms", "back_substitute_ms")},
                    "implicit": {k: si[k] for k in ("num_iterations", "setup_ms", "pcg_ms", "back_substitute_ms")},
                    "explicit_ms": te / reps, "implicit_ms": ti / reps, "rel_diff_of_solutions": util.rel_err(xe, xi)}
    its = max(se["num_iterations"], 1)
    out["break_even_pcg_iterations"] = (ms_elim - ms_setup_implicit - ms_sj) / max(ms_sx - ms_spmv, 1e-9)
    print(json.dumps(out, indent=1))
    if args.out:
        json.dump(out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
