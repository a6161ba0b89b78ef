"""World-size-2 test of the multi-GPU decomposition on CPU with the gloo backend.

The product shards points with cb200_shard_plan (host-only entry point: the same ShardPoints
cb200_create applies) and sums camera-sized partial results over ranks (NCCL on the GPU box).
Here each rank restates its shard's share with the oracle on the sub-problem of its own rows,
all-reduces with gloo, and the sum must equal the un-sharded oracle: reduced rhs, S x, and the
Schur-Jacobi blocks. Back substitution needs no communication: each rank's slice of y_e must equal
the full one."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import util


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _sub_problem(m, p, values, b, D, p0, p1, r0, r1):
    """The rows [r0, r1) of points [p0, p1) as a stand-alone <2,3,9> problem (all cameras kept)."""
    N = p.N
    prs = p.point_row_start[p0:p1 + 1] - r0
    cam = p.cam_idx[r0:r1]
    st = m.BlockStructure.bal(prs, cam, p.C)
    v = np.concatenate([values[6 * r0:6 * r1], values[6 * N + 18 * r0:6 * N + 18 * r1]])
    Dsub = np.concatenate([D[3 * p0:3 * p1], D[3 * p.P:]])
    return st, v, b[2 * r0:2 * r1].copy(), Dsub


def _worker(rank, world, port, out):
    from oracle import binding as ob
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = util.pkg()
        p = m.SyntheticBAL(11, 300, 1400, seed=4, geometry=False)
        values, b, D = p.random_values(4)
        D = np.abs(D) + 0.5
        pb, rb = m.shard_plan(p.structure, world)
        p0, p1, r0, r1 = int(pb[rank]), int(pb[rank + 1]), int(rb[rank]), int(rb[rank + 1])
        st, v, bs, Ds = _sub_problem(m, p, values, b, D, p0, p1, r0, r1)
        # camera diagonal is added once, after the reduction (replicated work on every rank)
        Ds_nocam = Ds.copy()
        Ds_nocam[3 * (p1 - p0):] = 0.0
        o = ob.Oracle(util.oracle_structure(st), preconditioner_type=ob.SCHUR_JACOBI)
        o.isc_init(v, Ds_nocam, bs)
        x = np.random.default_rng(9).standard_normal(9 * p.C)
        Df2 = D[3 * p.P:] ** 2
        parts = {"rhs": o.isc_rhs(), "sx": o.isc_multiply(x),
                 "sj": o.schur_jacobi(v, Ds_nocam, invert=False).reshape(-1)}
        # use_explicit_schur_complement: every rank forms the cells of its own points in the GLOBAL block
        # structure (cb200_explicit_plan, host only); the cells are summed over the ranks
        grow, gcols = m.explicit_plan(p.structure)
        lrow, lcols, lpos, lvals, lrhs = o.explicit_eliminate(v, bs, Ds_nocam)
        cells = np.zeros((len(gcols), 81))
        for i in range(p.C):
            idx = grow[i] + np.searchsorted(gcols[grow[i]:grow[i + 1]], lcols[lrow[i]:lrow[i + 1]])
            assert np.array_equal(gcols[idx], lcols[lrow[i]:lrow[i + 1]])  # a shard's pairs are a subset
            cells[idx] = lvals.reshape(-1, 81)[lrow[i]:lrow[i + 1]]
        parts["cells"] = cells.reshape(-1)
        parts["rhs_explicit"] = lrhs
        for k in parts:
            t = torch.from_numpy(parts[k].copy())
            dist.all_reduce(t)
            parts[k] = t.numpy()
        parts["sx"] = parts["sx"] + Df2 * x
        parts["sj"] = (parts["sj"].reshape(-1, 9, 9) + np.stack([np.diag(d) for d in Df2.reshape(-1, 9)])).reshape(-1)
        back = o.isc_back_substitute(x)[:3 * (p1 - p0)]
        # un-sharded oracle
        full = ob.Oracle(util.oracle_structure(p.structure), preconditioner_type=ob.SCHUR_JACOBI)
        full.isc_init(values, D, b)
        want = {"rhs": full.isc_rhs(), "sx": full.isc_multiply(x),
                "sj": full.schur_jacobi(values, D, invert=False).reshape(-1)}
        _, _, _, fvals, frhs = full.explicit_eliminate(values, b, D)
        gc = parts["cells"].reshape(-1, 9, 9)
        gc[grow[:-1]] += np.stack([np.diag(d) for d in Df2.reshape(-1, 9)])  # D_f^2 once, after the sum
        want["cells"], want["rhs_explicit"] = fvals, frhs
        errs = {k: util.rel_err(parts[k], want[k]) for k in want}
        errs["back"] = util.rel_err(back, full.isc_back_substitute(x)[3 * p0:3 * p1])
        errs["rows"] = float(abs((r1 - r0) - p.N / world) / p.N)
        out[rank] = errs
    finally:
        dist.destroy_process_group()


def test_point_sharded_sum_equals_unsharded_world2():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out))
             for r in range(world)]
    for pr in procs:
        pr.start()
    for pr in procs:
        pr.join(180)
        assert pr.exitcode == 0
    for r in range(world):
        errs = out[r]
        for k in ("rhs", "sx", "sj", "back", "cells", "rhs_explicit"):
            assert errs[k] < 1e-12, (r, k, errs)
        assert errs["rows"] < 0.02, errs  # balanced observation counts



def test_shard_plan_properties():
    m = util.pkg()
    p = m.SyntheticBAL(9, 500, 2600, seed=2, geometry=False)
    for world in (1, 2, 3, 8):
        pb, rb = m.shard_plan(p.structure, world)
        assert pb[0] == 0 and pb[-1] == p.P and rb[0] == 0 and rb[-1] == p.N
        assert np.all(np.diff(pb) >= 0)
        assert np.array_equal(rb, p.point_row_start[pb])
        assert np.abs(np.diff(rb) - p.N / world).max() <= np.diff(p.point_row_start).max()
