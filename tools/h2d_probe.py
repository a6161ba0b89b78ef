#!/usr/bin/env python
"""What the HOST side of the box gives: plain pinned-memory H2D copies (torch, no product code) running
concurrently on subsets of the GPUs. Explains the end-to-end scaling of bench.py's `e2e` (every step uploads the
Jacobian): the per-GPU copy rate falls when several GPUs behind the same PCIe root / switch copy at once.

    python tools/h2d_probe.py [--mb 512] [--out profiles/r2_h2d_probe.json]
"""
import argparse
import json
import time

import torch


def run(devices, nbytes, reps=6):
    bufs, devs, streams = [], [], []
    for d in devices:
        torch.cuda.set_device(d)
        bufs.append(torch.empty(nbytes, dtype=torch.uint8, pin_memory=True))
        devs.append(torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{d}"))
        streams.append(torch.cuda.Stream(device=d))
    best = None
    for _ in range(reps):
        for d in devices:
            torch.cuda.synchronize(d)
        t0 = time.perf_counter()
        for h, g, s in zip(bufs, devs, streams):
            with torch.cuda.stream(s):
                g.copy_(h, non_blocking=True)
        for d in devices:
            torch.cuda.synchronize(d)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    agg = len(devices) * nbytes / best / 1e9
    return {"devices": list(devices), "aggregate_gbs": agg, "per_gpu_gbs": agg / len(devices)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=512)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    n = torch.cuda.device_count()
    subsets = [[0]]
    if n >= 2:
        subsets += [[0, 1]]
    if n >= 4:
        subsets += [[0, 1, 2, 3], [0, 2]]
    if n >= 8:
        subsets += [[4, 5, 6, 7], [0, 4], [0, 1, 4, 5], list(range(8))]
    out = [run(s, args.mb << 20) for s in subsets]
    for r in out:
        print(f"GPUs {r['devices']}: {r['per_gpu_gbs']:.1f} GB/s per GPU, {r['aggregate_gbs']:.1f} GB/s aggregate", flush=True)
    if args.out:
        json.dump({"what": "concurrent pinned H2D copies, torch only", "mb_per_gpu": args.mb, "results": out},
                  open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
