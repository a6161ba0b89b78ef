#!/usr/bin/env python
"""bench.py — ITERATIVE_SCHUR linear solves per second on a synthetic Venice-shape BAL problem.

A step = the linear solve of one Levenberg-Marquardt iteration (ImplicitSchurComplement::Init,
SCHUR_JACOBI update, PCG to q_tolerance = eta = 1e-2, back substitution) on the Jacobi-scaled
Snavely Jacobian of the problem, exactly what LevenbergMarquardtStrategy::ComputeStep hands to
LinearSolver::Solve. `value` is measured with values / b / D resident in HBM; `e2e` goes through
cb200_solve with pinned HOST buffers (H2D of J, b, D and D2H of x inside the timed region).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--shape venice]
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
PKG = "ceres-solver-ceres-solver_b200"
METRIC = "itschur_linear_solves_per_sec"
UNIT = "linear solves/s"   # one step = the ITERATIVE_SCHUR linear solve of one LM iteration (lm_loop carries LM iterations/s)
FIXED_PCG = 50        # fixed-work PCG arm: min = max = 50 iterations, tolerances off (evaluation_benchmark.cc protocol)
ETA = 1e-2            # bundle_adjuster.cc:114
MAX_PCG = 500         # bundle_adjuster.cc:120
RADIUS = 1e4          # Solver::Options::initial_trust_region_radius


def make_problem(shape, scale):
    pkg = importlib.import_module(PKG)
    p = pkg.SyntheticBAL.named(shape, scale=scale, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    cost, residuals, values = p.evaluate()
    jscale = 1.0 / (1.0 + np.sqrt(p.squared_column_norm(values)))  # trust_region_minimizer.cc:263-279
    p.scale_columns(values, jscale)
    D = np.sqrt(np.clip(p.squared_column_norm(values), 1e-6, 1e32) / RADIUS)  # lm_strategy.cc:79-95
    return pkg, p, values, residuals, D


def algorithmic_bytes_sx(N, P, C):
    """J values once + camera ids + row starts + (E'E)^-1 blocks + x in / y out."""
    return 192 * N + 4 * N + 4 * P + 72 * P + 2 * 72 * C


def bind_to_gpu_numa_node(device):
    """Multi-rank runs: keep this process (and therefore the pages of the pinned host buffers it allocates
    next) on the CPUs NVML reports as local to its GPU. Ranks started by torchrun float over both sockets;
    a pinned buffer on the far socket makes every H2D copy cross the inter-socket link, and the timed value
    is the max over ranks. Returns the CPU list, or None when NVML / the affinity call is unavailable."""
    if os.environ.get("CB200_NO_NUMA_BIND"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(device)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return cpus
    except Exception:
        pass
    return None


class ClockSampler:
    """SM clock and throttle reasons sampled (NVML, every 50 ms, own thread) DURING the timed region. (Not more
    often: an NVML query holds driver locks for a good fraction of a millisecond, and with several ranks a launch
    delayed on one of them stalls every GPU at the next exchange.)"""

    def __init__(self, device):
        import threading
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(device)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None
            return
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def _run(self):
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for n, bit in names.items():
                    if mask & bit:
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.05)

    def stop(self):
        if self._thread is not None:
            self._stop.set()
            self._thread.join()
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def run_reference(args, rank):
    """The reference's own CPU implementation of the path. The reference cannot be compiled here
    (every translation unit of the path needs Eigen, absent from the image), so this arm times the
    oracle port of it on the host cores, all threads."""
    if rank != 0:
        return
    from oracle import binding as ob
    pkg, p, values, residuals, D = make_problem(args.shape, args.scale)
    st = p.structure
    cores = os.cpu_count()
    ost = ob.Structure(st.row_size, st.row_pos, st.cell_ptr, st.cell_block_id, st.cell_position, st.col_size,
                       st.col_pos, st.num_eliminate_blocks)
    o = ob.Oracle(ost, num_threads=cores, preconditioner_type=PRECOND[args.precond], max_num_iterations=MAX_PCG,
                  native=True)
    # bounded sample: at ~1 s per solve on 16 cores the driver's K steps are capped so the run ends in
    # about a minute; every timed step is a full solve of the same problem
    its = 0
    t0 = time.perf_counter()
    o.solve(values, residuals, D, ETA, -1.0)
    first = time.perf_counter() - t0
    timed = max(1, min(args.steps, int(60.0 / max(first, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(timed):
        _, s = o.solve(values, residuals, D, ETA, -1.0)
        its += s["num_iterations"]
    dt = time.perf_counter() - t0
    v = timed / dt
    # fixed-work PCG (evaluation_benchmark.cc's protocol): min = max = K iterations, both tests off
    of = ob.Oracle(ost, num_threads=cores, preconditioner_type=PRECOND[args.precond], min_num_iterations=FIXED_PCG,
                   max_num_iterations=FIXED_PCG, native=True)
    t0 = time.perf_counter()
    _, sf = of.solve(values, residuals, D, -1.0, -1.0)
    dtf = time.perf_counter() - t0
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT,
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / timed,
            "steps_timed": timed, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args, p), "pcg_iterations_per_step": its / timed,
            "pcg_iters_per_sec": its / dt,
            "pcg_iters_per_sec_fixed": sf["num_iterations"] / dtf,
            "pcg_fixed": {"iterations": sf["num_iterations"], "seconds": dtf,
                          "note": "one solve with min = max = %d PCG iterations, q / r tests off (includes its setup)" % FIXED_PCG},
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{timed} full solves of the same problem (1 warm-up)"},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


PRECOND = {"identity": 0, "jacobi": 1, "schur_jacobi": 2, "spse": 3}
PRECOND_NAME = {"identity": "IDENTITY", "jacobi": "JACOBI", "schur_jacobi": "SCHUR_JACOBI",
                "spse": "SCHUR_POWER_SERIES_EXPANSION"}


def workload_config(args, p):
    return {"workload": f"synthetic BAL {args.shape}-shape x{args.scale:g}: {p.C} cameras, {p.P} points, "
                        f"{p.N} observations; ITERATIVE_SCHUR + {PRECOND_NAME[args.precond]}, eta={ETA}, first LM iteration "
                        f"(Jacobi-scaled Snavely Jacobian, radius {RADIUS:g})",
            "cameras": p.C, "points": p.P, "observations": p.N,
            "l2": "inputs (24*N doubles) exceed the 126 MB L2; no flush needed"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--shape", default="venice")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-lm", action="store_true", help="skip the device-resident LM loop measurement")
    ap.add_argument("--no-cusparse", action="store_true", help="skip the cuSPARSE CSR comparison of S*x")
    ap.add_argument("--no-explicit", action="store_true",
                    help="skip the use_explicit_schur_complement comparison (SURVEY 8(f) row 3)")
    ap.add_argument("--no-fixed", action="store_true", help="skip the fixed-work PCG arm (short runs under ncu)")
    ap.add_argument("--lm-iterations", type=int, default=10)
    ap.add_argument("--precond", default="schur_jacobi", choices=["schur_jacobi", "spse", "jacobi", "identity"])
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)

    numa_cpus = bind_to_gpu_numa_node(local_rank) if world > 1 else None
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    uid = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pkg, p, values, residuals, D = make_problem(args.shape, args.scale)
    if world > 1:
        t = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            t = torch.frombuffer(bytearray(pkg.nccl_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(t, 0)
        uid = bytes(t.cpu().numpy().tobytes())

    def pinned(a):
        t = torch.empty(a.shape, dtype=torch.float64, pin_memory=True)
        t.numpy()[...] = a
        return t

    hv, hb, hD = pinned(values), pinned(residuals), pinned(D)
    hx = torch.zeros(3 * p.P + 9 * p.C, dtype=torch.float64, pin_memory=True)
    g = pkg.IterativeSchurComplementSolver(p.structure, preconditioner_type=PRECOND[args.precond],
                                           max_num_iterations=MAX_PCG, device_id=local_rank, rank=rank,
                                           world_size=world, nccl_unique_id=uid)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def stats(ms_list):
        a = np.asarray(ms_list, dtype=np.float64)
        return {"mean": float(a.mean()), "median": float(np.median(a)), "p95": float(np.percentile(a, 95)),
                "min": float(a.min()), "max": float(a.max())}

    def measure(handle, steps, warmup, barrier):
        """(resident per-step ms list, e2e per-step ms list, last resident summary, last e2e summary, iterations,
        launches): `steps` timed solves with inputs resident, then `steps` through the C ABI with host buffers."""
        handle.upload_ptr(hv.data_ptr(), hb.data_ptr(), hD.data_ptr())
        for _ in range(warmup):
            sr = handle.solve_resident(ETA, -1.0)
        barrier()
        res, its, launches = [], 0, 0
        for _ in range(steps):
            sr = handle.solve_resident(ETA, -1.0)
            res.append(sr["total_ms"])
            its += sr["num_iterations"]
            launches += sr["kernel_launches"]
        barrier()
        for _ in range(max(2, warmup)):
            handle.solve_ptr(hv.data_ptr(), hb.data_ptr(), hD.data_ptr(), ETA, -1.0, hx.data_ptr())
        barrier()
        e2e = []
        for _ in range(steps):
            se = handle.solve_ptr(hv.data_ptr(), hb.data_ptr(), hD.data_ptr(), ETA, -1.0, hx.data_ptr())
            e2e.append(se["total_ms"])
        barrier()
        return res, e2e, sr, se, its, launches

    # ---- HBM-resident arm, then end to end through the C ABI with host buffers ---------------------
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    t0 = time.perf_counter()
    res_ms, e2e_list, s, s2, its, launches = measure(g, args.steps, args.warmup, barrier)
    wall_ms = max_over_ranks(1e3 * (time.perf_counter() - t0))
    clocks = sampler.stop() if sampler else None
    # per step: the slowest rank (the step is over when every rank has its x)
    if world > 1:
        t = torch.tensor([res_ms, e2e_list], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res_ms, e2e_list = t[0].tolist(), t[1].tolist()
    dev_ms = float(np.sum(res_ms))
    e2e_ms = float(np.sum(e2e_list))
    e2e_wall_ms = wall_ms
    phases = {k: s[k] for k in ("setup_ms", "pcg_ms", "back_substitute_ms")}
    n_loc = g.row_end - g.row_begin
    p_loc = g.point_end - g.point_begin
    h2d = 8 * (24 * n_loc + 2 * n_loc + 3 * p_loc + 9 * p.C)
    d2h = 8 * (3 * p_loc + 9 * p.C)
    x_gpu = hx.numpy().copy()

    # ---- one-off cost of page-locking the caller's Jacobian (what the adapter does to BlockSparseMatrix::values()) --
    host_register_ms = None
    if world == 1 and rank == 0:
        t0 = time.perf_counter()
        pkg.host_register(values)
        host_register_ms = 1e3 * (time.perf_counter() - t0)
        pkg.host_unregister(values)

    # ---- fixed-work PCG (SURVEY 8(d) "Solver settings": min = max = K iterations, q / r tests off) ----------
    g.set_iterations(FIXED_PCG, FIXED_PCG)
    for _ in range(0 if args.no_fixed else 2):
        g.solve_resident(-1.0, -1.0)
    barrier()
    fixed_ms, fixed_its = 0.0, 0
    for _ in range(1 if args.no_fixed else 3):
        sf = g.solve_resident(-1.0, -1.0)
        fixed_ms += sf["total_ms"]
        fixed_its += sf["num_iterations"]
    fixed_ms = max_over_ranks(fixed_ms)
    g.set_iterations(0, MAX_PCG)
    nfix = 1 if args.no_fixed else 3
    pcg_fixed = {"iterations_per_solve": fixed_its / nfix, "ms_per_solve": fixed_ms / nfix,
                 "ms_per_pcg_iteration": (fixed_ms - nfix * (s["setup_ms"])) / max(fixed_its, 1),
                 "note": "3 solves with min = max = %d PCG iterations, q / r tests off; the rate counts the whole "
                         "solve (setup and back substitution included)" % FIXED_PCG}

    # ---- roofline of the dominant kernel (fused S*x), timed alone with CUDA events ----------------
    sx_ms, _ = g.time_op(6, warmup=3, iters=20)
    sx_ms = max_over_ranks(sx_ms)
    bytes_sx = algorithmic_bytes_sx(n_loc, p_loc, p.C)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic, traffic_source = None, None  # dram__bytes_read + dram__bytes_write of one launch, from the committed ncu capture
    for name in ("r2_sx_traffic.json", "r1_sx_traffic.json"):
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", name)))
            if world == 1 and args.shape == t.get("shape") and args.scale == 1.0:
                traffic = t["traffic_bytes_per_launch"]
                traffic_source = f"profiles/{name} (one ncu --set full capture of this kernel on this problem; not re-measured in this run)"
                break
        except Exception:
            pass
    # the same copy probe MEASURED_PEAKS.json was taken with, on THIS box (boxes of the pool differ by several %)
    copy_gbs = None
    try:
        a = torch.empty(1 << 30, dtype=torch.bfloat16, device="cuda")
        bq = torch.empty_like(a)
        best = 1e9
        for _ in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            bq.copy_(a)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        copy_gbs = 2 * a.numel() * 2 / (best * 1e-3) / 1e9
        del a, bq
        torch.cuda.empty_cache()
    except Exception:
        pass
    achieved = bytes_sx / (sx_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "point_stream_kernel<kModeSx> (+ point_long_kernel for points of > 256 rows)", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_source,
                "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback",
                "bytes_per_launch": bytes_sx, "ms_per_launch": sx_ms,
                "hbm_copy_gbs_this_box": copy_gbs,
                "frac_of_this_box_copy": (achieved / copy_gbs) if copy_gbs else None}

    # ---- the reference's CUDA SpMV method (cuSPARSE CSR), re-created: what the fused kernel replaces -------
    if world == 1 and not args.no_cusparse:
        try:
            cs = cusparse_sx(p, values, D)
            ours = g.isc_multiply(cs["x"])  # same operator, same x: the comparison is like for like
            roofline["cusparse_csr_reference_method"] = {
                "sx_ms": cs["sx_ms"], "speedup_of_fused_kernel": cs["sx_ms"] / sx_ms, "method": cs["method"],
                "rel_diff_of_results": float(np.linalg.norm(ours - cs["y"]) / np.linalg.norm(cs["y"]))}
            del cs
            torch.cuda.empty_cache()
        except Exception as e:  # a comparison, never a reason to lose the bench line
            roofline["cusparse_csr_reference_method"] = {"unavailable": repr(e)[:200]}

    # ---- widened boundary: whole LM iterations on the device-resident problem ----
    lm = None
    if not args.no_lm:
        g.bal_attach(p.observations)
        barrier()
        t0 = time.perf_counter()
        g.bal_set_parameters(p.cameras, p.points)          # H2D of the parameter blocks
        r = g.lm_minimize(eta=ETA, max_num_iterations=args.lm_iterations, initial_trust_region_radius=RADIUS)
        cam_out, pts_out = g.bal_get_parameters()          # D2H of the result
        barrier()
        wall = max_over_ranks(time.perf_counter() - t0)
        n_it = len(r["iterations"]) - 1
        lm = {"lm_iterations": n_it, "lm_iters_per_sec": n_it / wall, "ms_per_lm_iteration": 1e3 * wall / max(n_it, 1),
              "initial_cost": r["initial_cost"], "final_cost": r["final_cost"], "termination": r["message"],
              "successful_steps": r["num_successful_steps"] - 1, "pcg_iterations": r["num_linear_solver_iterations"],
              "evaluate_ms": r["evaluate_ms"], "linear_solver_ms": r["linear_solver_ms"],
              "h2d_bytes": 8 * (9 * p.C + 3 * p_loc), "d2h_bytes": 8 * (9 * p.C + 3 * p_loc),
              "note": "cb200_lm_minimize: device Snavely evaluator + ITERATIVE_SCHUR solve + step evaluation per "
                      "iteration; timed from host parameter buffers to host parameter buffers"}

    # ---- the same solve with use_explicit_schur_complement (explicit block-sparse S) ----
    explicit = None
    if world == 1 and not args.no_explicit and args.precond == "schur_jacobi":
        try:
            g.upload_ptr(hv.data_ptr(), hb.data_ptr(), hD.data_ptr())
            si = g.solve_resident(ETA, -1.0)
            xi = g.download_solution()
            ge = pkg.IterativeSchurComplementSolver(p.structure, preconditioner_type=PRECOND[args.precond],
                                                    max_num_iterations=MAX_PCG, device_id=local_rank,
                                                    use_explicit_schur_complement=True)
            ge.upload_ptr(hv.data_ptr(), hb.data_ptr(), hD.data_ptr())
            ge.solve_resident(ETA, -1.0)
            ms = 0.0
            for _ in range(5):
                se = ge.solve_resident(ETA, -1.0)
                ms += (se["setup_ms"] + se["pcg_ms"] + se["back_substitute_ms"]) / 5
            xe = ge.download_solution()
            form_ms, _ = ge.time_op(13, warmup=1, iters=5)
            spmv_ms, _ = ge.time_op(15, warmup=2, iters=20)
            cells = len(pkg.explicit_plan(p.structure)[1])
            explicit = {"ms_per_step": ms, "steps_per_sec": 1e3 / ms, "pcg_iterations": se["num_iterations"],
                        "implicit_pcg_iterations": si["num_iterations"], "eliminate_ms": form_ms,
                        "explicit_sx_ms": spmv_ms, "implicit_sx_ms": sx_ms, "upper_cells": cells,
                        "S_bytes": 648 * cells, "rel_diff_of_solutions": float(np.linalg.norm(xe - xi) / np.linalg.norm(xi)),
                        "note": "same handle type created with use_explicit_schur_complement: SchurEliminator::Eliminate "
                                "into a block-sparse S + PCG on it; slower than the implicit operator at this shape "
                                "(DESIGN.md 7.2), reported for completeness"}
            ge.close()
        except Exception as e:  # a comparison, never a reason to lose the bench line
            explicit = {"unavailable": repr(e)[:200]}

    # ---- parity of THIS run against the oracle (every N): each rank compares what it wrote of x -----------------
    parity = None
    cpu = None
    if not args.no_cpu_baseline:
        xo = np.zeros_like(x_gpu)
        so_its = torch.zeros(1, dtype=torch.float64, device="cuda")
        if rank == 0:
            cpu, xo, so = cpu_baseline(p, values, residuals, D, PRECOND[args.precond], solves=3 if world == 1 else 1,
                                       lm=(world == 1 and not args.no_lm))
            so_its[0] = so["num_iterations"]
        if world > 1:
            txo = torch.from_numpy(xo).cuda()
            dist.broadcast(txo, 0)
            dist.broadcast(so_its, 0)
            xo = txo.cpu().numpy()
        ne = 3 * p.P
        sl = np.r_[3 * g.point_begin:3 * g.point_end, ne:ne + 9 * p.C]
        err = float(np.linalg.norm(x_gpu[sl] - xo[sl]) / np.linalg.norm(xo[sl]))
        err = max_over_ranks(err)
        parity = {"rel_err_vs_oracle": err, "tolerance": 1e-9,
                  "iterations_equal": bool(int(so_its.item()) == s2["num_iterations"]),
                  "pcg_iterations": s2["num_iterations"], "oracle_pcg_iterations": int(so_its.item()),
                  "what": "x of the last timed end-to-end solve (every rank's own point entries + the camera entries) "
                          "against the oracle's x for the same problem"}

    # ---- the same N GPUs behind ONE handle in ONE process (cb200_options::num_gpus: what a single-process
    # ---- ceres::Solve reaches through the adapter); measured on rank 0 while the other ranks wait ---------------
    single = None
    if world > 1:
        g.close()
        g = None
        torch.cuda.empty_cache()
        barrier()
        # (the other ranks wait on the rendezvous store, on the CPU: a NCCL barrier would keep a polling kernel on
        #  their GPUs, time-slicing against the very GPUs rank 0 is about to measure)
        store = dist.distributed_c10d._get_default_store()
        if rank != 0:
            store.wait(["cb200_single_process_done"])
        if rank == 0:
            try:
                gm = pkg.IterativeSchurComplementSolver(p.structure, preconditioner_type=PRECOND[args.precond],
                                                        max_num_iterations=MAX_PCG, num_gpus=world)
                r1, e1, sa, sb, its1, l1 = measure(gm, args.steps, args.warmup, torch.cuda.synchronize)
                xs = hx.numpy()
                single = {"value": len(r1) / (1e-3 * float(np.sum(r1))), "unit": UNIT, "ms_per_step": stats(r1),
                          "e2e": {"value": len(e1) / (1e-3 * float(np.sum(e1))), "unit": UNIT, "ms_per_step": stats(e1),
                                  "h2d_ms": sb["h2d_ms"]},
                          "pcg_iterations_per_step": its1 / len(r1), "gpu_launches_per_gpu": l1,
                          "phases_ms_last_step": {k: sa[k] for k in ("setup_ms", "pcg_ms", "back_substitute_ms")},
                          "rel_err_vs_oracle": (float(np.linalg.norm(xs - xo) / np.linalg.norm(xo)) if parity else None),
                          "note": "one process, one handle with num_gpus = %d (worker thread per GPU, sums over NVLink peer "
                                  "memory); same problem, same steps" % world}
                gm.close()
            except Exception as e:
                single = {"unavailable": repr(e)[:300]}
            store.set("cb200_single_process_done", "1")
        barrier()

    if rank == 0:
        line = {"metric": METRIC, "value": args.steps / (dev_ms * 1e-3), "unit": UNIT,
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": workload_config(args, p),
                "ms_per_step_stats": stats(res_ms),
                "pcg_iterations_per_step": its / args.steps, "pcg_iters_per_sec": its / (dev_ms * 1e-3),
                "pcg_iters_per_sec_fixed": 1e3 * FIXED_PCG / pcg_fixed["ms_per_solve"], "pcg_fixed": pcg_fixed,
                "wall_ms_per_step_both_arms": wall_ms / (2 * args.steps), "phases_ms_last_step": phases,
                "termination": s["message"],
                "e2e": {"value": args.steps / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms / args.steps, "ms_per_step_stats": stats(e2e_list),
                        "host_cpus_bound_to_gpu_numa_node": len(numa_cpus) if numa_cpus else None,
                        "h2d_ms": s2["h2d_ms"], "d2h_ms": s2["d2h_ms"],
                        "h2d_gbs_per_gpu": h2d / (s2["h2d_ms"] * 1e-3) / 1e9 if s2["h2d_ms"] > 0 else None,
                        "host_register_ms": host_register_ms,
                        "note": "cb200_solve from pinned host buffers: Jacobian copied in 16 point-range chunks, the "
                                "E'E / rhs pass of a chunk runs while the next one crosses PCIe; ms = first H2D byte to "
                                "last D2H byte (CUDA events)"},
                "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "lm_loop": lm,
                "explicit_schur": explicit, "parity": parity, "single_process_multi_gpu": single}
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def cusparse_sx(p, values, D, iters=10):
    """SURVEY 8(d) "optional GPU cross-check baseline": S*x the way the reference's own CUDA path would do it --
    E and F as scalar CSR matrices and one cuSPARSE SpMV per product (CudaSparseMatrix::Right/LeftMultiplyAndAccumulate,
    cuda_sparse_matrix.cc:185-222, over the CRS views of cuda_partitioned_block_sparse_crs_view.cc:105-119), re-created
    with torch.sparse CSR (= cusparseSpMV). Library code, a comparison only: nothing of the product runs here. The
    transposes are stored as their own CSR matrices, which favours this baseline (the reference multiplies with
    CUSPARSE_OPERATION_TRANSPOSE)."""
    import torch
    dev = torch.device("cuda")
    N, P, C = p.N, p.P, p.C
    v = torch.from_numpy(values).to(dev)
    cam = torch.from_numpy(p.cam_idx.astype("int64")).to(dev)
    pt = torch.repeat_interleave(torch.arange(P, device=dev), torch.from_numpy(np.diff(p.point_row_start).astype("int64")).to(dev))
    E = torch.sparse_csr_tensor((torch.arange(2 * N + 1, device=dev) * 3).to(torch.int32),
                                ((pt.repeat_interleave(2) * 3)[:, None] + torch.arange(3, device=dev)[None, :]).reshape(-1).to(torch.int32),
                                v[:6 * N].clone(), size=(2 * N, 3 * P))
    F = torch.sparse_csr_tensor((torch.arange(2 * N + 1, device=dev) * 9).to(torch.int32),
                                ((cam.repeat_interleave(2) * 9)[:, None] + torch.arange(9, device=dev)[None, :]).reshape(-1).to(torch.int32),
                                v[6 * N:].clone(), size=(2 * N, 9 * C))
    Et = E.to_sparse_coo().t().coalesce().to_sparse_csr()
    Ft = F.to_sparse_coo().t().coalesce().to_sparse_csr()
    # block-diagonal (E'E + D_e^2)^-1 as P dense 3x3 blocks, built with library calls as well
    e = v[:6 * N].reshape(N, 2, 3)
    ete = torch.zeros(P, 3, 3, dtype=torch.float64, device=dev).index_add_(0, pt, e.transpose(1, 2) @ e)
    Dt = torch.from_numpy(D).to(dev)
    ete += torch.diag_embed(Dt[:3 * P].reshape(P, 3) ** 2)
    inv = torch.linalg.inv(ete)
    df2 = Dt[3 * P:] ** 2
    x = torch.randn(9 * C, dtype=torch.float64, device=dev)
    def sx():
        t = F @ x
        u = Et @ t
        w = (inv @ u.reshape(P, 3, 1)).reshape(-1)
        t = t - E @ w
        return df2 * x + Ft @ t
    y = sx()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        y = sx()
    b.record()
    torch.cuda.synchronize()
    return {"sx_ms": a.elapsed_time(b) / iters, "x": x.cpu().numpy(), "y": y.cpu().numpy(),
            "method": "torch.sparse CSR (cusparseSpMV) x4 + batched 3x3 apply; scalar CSR of E, F and their stored transposes"}


def cpu_baseline(p, values, residuals, D, precond, solves=3, lm=True):
    """Oracle port of the reference's CPU path on all host threads: one warm-up solve, then the mean of `solves`
    full solves of the same problem. Returns (baseline object, x of the last solve, its summary): the x is what
    the GPU result of this run is checked against (the line's `parity`)."""
    from oracle import binding as ob
    st = p.structure
    cores = os.cpu_count()
    ost = ob.Structure(st.row_size, st.row_pos, st.cell_ptr, st.cell_block_id, st.cell_position, st.col_size,
                       st.col_pos, st.num_eliminate_blocks)
    o = ob.Oracle(ost, num_threads=cores, preconditioner_type=precond, max_num_iterations=MAX_PCG, native=True)
    x, s = o.solve(values, residuals, D, ETA, -1.0)  # warm-up (page faults, thread start)
    t0 = time.perf_counter()
    for _ in range(solves):
        x, s = o.solve(values, residuals, D, ETA, -1.0)
    dt = (time.perf_counter() - t0) / solves
    out = {"value": 1.0 / dt, "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"mean of {solves} warm full solves of the same problem ({s['num_iterations']} PCG iterations, "
                     f"{dt:.2f} s each)",
           "pcg_iters_per_sec": s["num_iterations"] / dt}
    if lm:
        # the same widening as lm_loop: whole LM iterations (host evaluator + oracle solve + step evaluation).
        # A Python harness around the oracle: reported for orientation, NOT a baseline to divide lm_loop by.
        from oracle import lm_reference as lmr
        t0 = time.perf_counter()
        r = lmr.minimize(p, p.cameras, p.points, lambda v, b, Dv, e: o.solve(v, b, Dv, e, -1.0), eta=ETA,
                         max_num_iterations=2, initial_trust_region_radius=RADIUS, independent_evaluator=False)
        dt = time.perf_counter() - t0
        n_it = len(r["iterations"]) - 1
        out["lm_loop"] = {"lm_iterations": n_it, "lm_iters_per_sec": n_it / dt, "final_cost": r["final_cost"],
                          "sample": f"{n_it} LM iterations, {dt:.1f} s (OpenMP evaluator, numpy column passes in Python, "
                                    f"oracle solve): a harness, not a tuned baseline"}
    return out, x, s


if __name__ == "__main__":
    main()
