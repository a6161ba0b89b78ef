#!/usr/bin/env python
"""Run bench.py on every BASELINE.json configuration that fits one GPU and keep one JSON line per configuration
under profiles/ (VERDICT r1 'Next round' 7): problem-16 shape, Ladybug shape, Venice shape with
SCHUR_POWER_SERIES_EXPANSION, Final-13682 shape. (Venice + SCHUR_JACOBI is the driver's own bench line.)

    python tools/r2_sweep.py [--out profiles] [--only NAME ...]
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CONFIGS = {
    "problem16": ["--shape", "problem-16", "--steps", "200"],
    "ladybug": ["--shape", "ladybug", "--steps", "100"],
    "venice_spse": ["--shape", "venice", "--precond", "spse", "--steps", "20", "--no-explicit"],
    "final13682": ["--shape", "final-13682", "--steps", "10", "--no-cusparse", "--no-explicit", "--lm-iterations", "5"],
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles"))
    ap.add_argument("--only", nargs="*", default=None)
    args = ap.parse_args()
    for name, flags in CONFIGS.items():
        if args.only and name not in args.only:
            continue
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--warmup", "3"] + flags,
                           capture_output=True, text=True)
        line = next((l for l in r.stdout.splitlines() if l.startswith("{")), None)
        if line is None:
            print(f"{name}: FAILED\n{r.stderr[-2000:]}", flush=True)
            continue
        d = json.loads(line)
        with open(os.path.join(args.out, f"r2_bench_{name}.json"), "w") as f:
            json.dump(d, f, indent=1)
        print(f"{name}: {d['value']:.1f} {d['unit']} resident, {d['e2e']['value']:.1f} e2e, roofline frac "
              f"{d['roofline']['frac']:.3f}, parity {d['parity'] and d['parity']['rel_err_vs_oracle']:.2e} "
              f"(iterations equal: {d['parity'] and d['parity']['iterations_equal']}), cpu {d['cpu_baseline']['value']:.2f}",
              flush=True)


if __name__ == "__main__":
    main()
