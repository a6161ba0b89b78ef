#!/usr/bin/env python
"""The reference's per-op measurement protocol (internal/ceres/evaluation_benchmark.cc:333-608) on this path:
every product it times, on the BAL structure with N(0,1) values (its :89-99,147-153), device-resident, CUDA
events around back-to-back launches. Bytes per op are SURVEY.md section 8(d)'s algorithmic figures.

    python tools/evaluation_benchmark.py --shape venice [--cpu]   (--cpu: the oracle port on the host cores too)
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="venice")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    pkg = importlib.import_module("ceres-solver-ceres-solver_b200")
    p = pkg.SyntheticBAL.named(args.shape, geometry=False)
    values, b, D = p.random_values(1)
    D = np.abs(D) + 1.0
    N, P, C = p.N, p.P, p.C
    B_sx = 192 * N + 4 * N + 76 * P + 144 * C
    rows = [  # (reference benchmark, op of cb200_time_op, algorithmic bytes, what runs here)
        ("PMVRightMultiplyAndAccumulateF", 9, 144 * N + 4 * N + 32 * N + 72 * C, "pmv_f_kernel (seam)"),
        ("PMVLeftMultiplyAndAccumulateF", 10, 144 * N + 4 * N + 16 * N + 72 * C, "pmv_f_kernel (seam)"),
        ("PMVRightMultiplyAndAccumulateE", 7, 48 * N + 32 * N + 24 * P + 4 * P, "pmv_e_kernel (seam)"),
        ("PMVLeftMultiplyAndAccumulateE", 8, 48 * N + 16 * N + 48 * P + 4 * P, "pmv_e_kernel (seam)"),
        ("PMVUpdateBlockDiagonalEtE (+ inverse, rhs, weights)", 1, 208 * N + 4 * N + 76 * P + 24 * N + 48 * P,
         "point_stream_kernel<Rhs, BUILD>"),
        ("PMVUpdateBlockDiagonalFtF", 11, 144 * N + 4 * N + 648 * C, "camera_block_kernel<false>"),
        ("ISCRightMultiplyAndAccumulate (4 products + temporaries in the reference)", 6, B_sx,
         "point_stream_kernel<Sx>: one fused pass"),
        ("PSEPreconditionerRightMultiplyAndAccumulate (5 terms)", 12, 5 * B_sx + 12 * 648 * C,
         "PseApply: 5 x point_stream_kernel<PseZ> + 9x9 block applies"),
        ("SchurJacobi Update (eliminator + invert)", 3, 172 * N + 648 * C, "camera_block_kernel<true> + invert"),
        ("BackSubstitute", 4, 212 * N + 72 * P, "point_stream_kernel<Back>"),
    ]
    g = pkg.IterativeSchurComplementSolver(p.structure, preconditioner_type=pkg.SCHUR_POWER_SERIES_EXPANSION,
                                           max_num_iterations=50)
    g.upload(values, b, D)
    g.isc_init()
    out = []
    print(f"{args.shape}: C={C} P={P} N={N}")
    print(f"{'reference benchmark':78s} {'ms':>8s} {'GB/s':>7s}  kernel")
    for name, op, nbytes, kernel in rows:
        ms, launches = g.time_op(op, warmup=2, iters=args.iters)
        out.append({"benchmark": name, "ms": ms, "algorithmic_bytes": nbytes, "gbs": nbytes / ms / 1e6,
                    "kernel": kernel, "launches": launches})
        print(f"{name:78s} {ms:8.4f} {nbytes / ms / 1e6:7.0f}  {kernel}", flush=True)
    g.close()
    if args.cpu:
        from oracle import binding as ob
        from tests import util
        o = ob.Oracle(util.oracle_structure(p.structure), num_threads=os.cpu_count() or 1,
                      preconditioner_type=ob.SCHUR_JACOBI, max_num_iterations=50)
        o.isc_init(values, D, b)
        x = np.random.default_rng(0).standard_normal(9 * C)
        t0 = time.perf_counter()
        for _ in range(3):
            o.isc_multiply(x)
        cpu_ms = (time.perf_counter() - t0) / 3 * 1e3
        out.append({"benchmark": "ISCRightMultiplyAndAccumulate, oracle port on the host", "ms": cpu_ms,
                    "threads": os.cpu_count()})
        print(f"{'ISCRightMultiplyAndAccumulate, oracle port, ' + str(os.cpu_count()) + ' threads':78s} {cpu_ms:8.2f}")
    if args.out:
        json.dump({"shape": args.shape, "C": C, "P": P, "N": N, "rows": out}, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
