#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py into per-solve tables
(profiles/r2_bench_launches_one_solve.txt).   python tools/launch_list_summary.py launches.csv out.txt"""
import csv
import sys


def main():
    src, dst = sys.argv[1], sys.argv[2]
    with open(src) as f:
        lines = [l for l in f if not l.startswith("==")]
    r = csv.reader(lines)
    hdr = next(r)
    data = [x for x in r if len(x) == len(hdr)]
    names = [(x[4].split("(")[0].replace("void ", ""), float(x[-1]) / 1e3, x[8]) for x in data]
    starts = [i for i, (n, t, g) in enumerate(names) if n.startswith("point_stream_kernel<2, 1, 0>")]

    def show(a, b, title):
        tot = sum(t for n, t, g in names[a:b])
        out = [f"# {title}: {b - a} launches, {tot:.1f} us (ncu gpu__time_duration.sum: cold-cache, serialised -- compare shares)",
               "#      us  share          grid  kernel"]
        for n, t, g in names[a:b]:
            out.append(f"{t:9.2f} {100 * t / tot:5.1f}%  {g:>12}  {n}")
        out.append(f"{tot:9.2f} total")
        return out

    # chunked end-to-end solves: runs of consecutive E'E / rhs passes (one per upload chunk)
    runs, i = [], 0
    while i < len(starts):
        j = i
        while j + 1 < len(starts) and starts[j + 1] - starts[j] <= 2:
            j += 1
        if j - i >= 4:
            runs.append((i, j))
        i = j + 1
    in_run = {k for a, b in runs for k in range(a, b + 1)}
    singles = [k for k in range(len(starts)) if k not in in_run]
    text = []
    if len(singles) >= 5:
        a = starts[singles[4]]
        b = starts[singles[4] + 1] if singles[4] + 1 < len(starts) else len(names)
        text += show(a, b, "one HBM-resident solve of bench.py (Venice shape, SCHUR_JACOBI, 2 PCG iterations; replayed from CUDA graphs)")
    if runs:
        a = starts[runs[-1][0]]
        nxt = runs[-1][1] + 1
        b = starts[nxt] if nxt < len(starts) else len(names)
        text += [""] + show(a, b, "one end-to-end solve (cb200_solve from host buffers: E'E / rhs pass per upload chunk, then the rest)")
    open(dst, "w").write("\n".join(text) + "\n")
    print("\n".join(text))


if __name__ == "__main__":
    main()
