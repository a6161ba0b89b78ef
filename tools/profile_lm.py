#!/usr/bin/env python
"""Run the device-resident LM loop on a synthetic BAL shape (for ncu / quick timing)."""
import argparse
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="venice")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--iterations", type=int, default=3)
    ap.add_argument("--eta", type=float, default=1e-2)
    ap.add_argument("--repeat", type=int, default=2)
    ap.add_argument("--explicit", action="store_true", help="use_explicit_schur_complement")
    ap.add_argument("--switch", type=int, default=0, help="explicit_switch_iterations")
    args = ap.parse_args()
    pkg = importlib.import_module("ceres-solver-ceres-solver_b200")
    p = pkg.SyntheticBAL.named(args.shape, scale=args.scale, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    g = pkg.IterativeSchurComplementSolver(p.structure, max_num_iterations=500,
                                           use_explicit_schur_complement=args.explicit,
                                           explicit_switch_iterations=args.switch)
    g.bal_attach(p.observations)
    for _ in range(args.repeat):
        g.bal_set_parameters(p.cameras, p.points)
        t0 = time.perf_counter()
        r = g.lm_minimize(eta=args.eta, max_num_iterations=args.iterations)
        dt = time.perf_counter() - t0
        n = len(r["iterations"]) - 1
        print(f"{n} LM iterations in {dt*1e3:.2f} ms: evaluate {r['evaluate_ms']:.2f} ms, linear solver "
              f"{r['linear_solver_ms']:.2f} ms, {r['num_linear_solver_iterations']} PCG iterations; "
              f"cost {r['initial_cost']:.6e} -> {r['final_cost']:.6e}; {r['message']}")
        for it in r["iterations"]:
            print("  ", it["iteration"], f"{it['cost']:.6e}", it["step_is_successful"], it["linear_solver_iterations"],
                  f"{it['iteration_ms']:.2f} ms")


if __name__ == "__main__":
    main()
