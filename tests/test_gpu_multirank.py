"""Two-GPU parity: one process per GPU, points sharded, NCCL all-reduce of the camera-sized partials
(SURVEY 8(e)). Each rank's solution (its own points' entries + all camera entries) must match the
un-sharded oracle within the same 1e-9 bar, with equal iteration counts. Skipped with < 2 GPUs."""
import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from tests import util

pytestmark = pytest.mark.gpu


def _worker(rank, world, uid, ptype, out):
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug", scale=0.05)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    g = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=ptype, max_num_iterations=300,
                                         device_id=rank, rank=rank, world_size=world, nccl_unique_id=uid)
    x = np.full(p.structure.num_cols, np.nan)
    x, s = g.solve(values, residuals, D, 1e-3, -1.0, x=x)
    sx = g.isc_multiply(np.arange(9 * p.C, dtype=np.float64) / (9 * p.C))
    out[rank] = (x, s["num_iterations"], s["termination_type"], g.point_begin, g.point_end, sx)
    g.close()


@pytest.mark.parametrize("ptype", [2, 3])
def test_two_gpu_solve_matches_oracle(ptype):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from oracle import binding as ob
    m = util.pkg()
    world = 2
    uid = m.nccl_unique_id()
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    procs = [ctx.Process(target=_worker, args=(r, world, uid, ptype, out)) for r in range(world)]
    for pr in procs:
        pr.start()
    for pr in procs:
        pr.join(300)
        assert pr.exitcode == 0
    p = m.SyntheticBAL.named("ladybug", scale=0.05)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    o = ob.Oracle(util.oracle_structure(p.structure), num_threads=8, preconditioner_type=ptype,
                  max_num_iterations=300)
    xo, so = o.solve(values, residuals, D, 1e-3, -1.0)
    o.isc_init(values, D, residuals)
    sxo = o.isc_multiply(np.arange(9 * p.C, dtype=np.float64) / (9 * p.C))
    ne = 3 * p.P
    covered = np.zeros(p.P, dtype=bool)
    for r in range(world):
        x, its, term, pb, pe, sx = out[r]
        assert its == so["num_iterations"] and term == so["termination_type"]
        assert util.rel_err(x[ne:], xo[ne:]) < 1e-9
        assert util.rel_err(x[3 * pb:3 * pe], xo[3 * pb:3 * pe]) < 1e-9
        assert util.rel_err(sx, sxo) < 1e-12
        # entries of other ranks' points are left untouched
        mask = np.ones(ne, dtype=bool)
        mask[3 * pb:3 * pe] = False
        assert np.all(np.isnan(x[:ne][mask]))
        covered[pb:pe] = True
    assert covered.all()
