#!/usr/bin/env python
"""Run one device op of the solver back to back on a synthetic BAL shape (for ncu / quick timing).

    python tools/profile_op.py --shape venice --op 6 --iters 5
ops: see cb200_time_op in include/ceres_b200_schur.h
"""
import argparse
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="venice")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--op", type=int, nargs="+", default=[6])
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--precond", type=int, default=2)
    ap.add_argument("--nrep", type=int, nargs="+", default=[0])
    ap.add_argument("--pcg", type=int, default=0, help="time a solve with exactly this many PCG iterations")
    ap.add_argument("--period", type=int, default=0)
    ap.add_argument("--deterministic", action="store_true", help="cb200_options::deterministic")
    ap.add_argument("--lm-values", action="store_true",
                    help="the bench's problem (Snavely Jacobian, Jacobi-scaled, LM diagonal) instead of N(0,1) values")
    ap.add_argument("--solves-first", type=int, default=0, help="run this many LM-shaped solves before timing the ops")
    args = ap.parse_args()
    pkg = importlib.import_module("ceres-solver-ceres-solver_b200")
    if args.lm_values:
        p = pkg.SyntheticBAL.named(args.shape, scale=args.scale, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
        cost, b, values = p.evaluate()
        jscale = 1.0 / (1.0 + np.sqrt(p.squared_column_norm(values)))
        p.scale_columns(values, jscale)
        D = np.sqrt(np.clip(p.squared_column_norm(values), 1e-6, 1e32) / 1e4)
    else:
        p = pkg.SyntheticBAL.named(args.shape, scale=args.scale, geometry=False)
        values, b, D = p.random_values(1)
        D = np.abs(D) + 1.0
    N, P, C = p.N, p.P, p.C
    bytes_sx = 192 * N + 4 * N + 4 * P + 72 * P + 144 * C
    for nrep in args.nrep:
        g = pkg.IterativeSchurComplementSolver(p.structure, preconditioner_type=args.precond, max_num_iterations=50,
                                               y_replicas=nrep, status_check_period=args.period,
                                               deterministic=args.deterministic)
        g.upload(values, b, D)
        g.isc_init()
        for _ in range(args.solves_first):
            g.solve_resident(1e-2, -1.0)
        for op in args.op:
            ms, launches = g.time_op(op, warmup=args.warmup, iters=args.iters)
            print(f"nrep {nrep} op {op}: {ms:.4f} ms/launch-group ({launches} launches); "
                  f"vs B_Sx={bytes_sx/1e9:.3f} GB -> {bytes_sx/ms/1e6:.0f} GB/s", flush=True)
        if args.pcg:
            g.set_iterations(args.pcg, args.pcg)
            for _ in range(3):
                sm = g.solve_resident(-1.0, -1.0)
            print(f"pcg {sm['num_iterations']} iterations: setup {sm['setup_ms']:.3f} ms, pcg {sm['pcg_ms']:.3f} ms "
                  f"({sm['pcg_ms'] / sm['num_iterations'] * 1e3:.1f} us/iteration), back {sm['back_substitute_ms']:.3f} ms, "
                  f"launches {sm['kernel_launches']}", flush=True)
        g.close()


if __name__ == "__main__":
    main()
