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


def _worlds():
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return sorted({2, n}) if n >= 2 else [2]


@pytest.mark.parametrize("world", _worlds())
@pytest.mark.parametrize("ptype", [2, 3])
def test_two_gpu_solve_matches_oracle(ptype, world):
    """world = 2 and world = every GPU of the box (8 on the scaling box)."""
    if torch.cuda.device_count() < world:
        pytest.skip("needs 2 GPUs")
    from oracle import binding as ob
    m = util.pkg()
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


def _lm_worker(rank, world, uid, out):
    m = util.pkg()
    p = m.SyntheticBAL(31, 3000, 14000, seed=7, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    g = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=500,
                                         device_id=rank, rank=rank, world_size=world, nccl_unique_id=uid)
    g.bal_attach(p.observations)
    g.bal_set_parameters(p.cameras, p.points)
    r = g.lm_minimize(eta=1e-2, max_num_iterations=6)
    cam, pts = g.bal_get_parameters()
    out[rank] = (r["initial_cost"], r["final_cost"], [(i["iteration"], i["linear_solver_iterations"], i["cost"])
                                                      for i in r["iterations"]], cam, pts, g.point_begin, g.point_end)
    g.close()


def test_two_gpu_lm_loop_matches_reference_loop():
    """The device trust-region loop with the points sharded over two GPUs (evaluator, column norms,
    gradient, candidate cost and model cost all reduced over ranks) against the CPU loop."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from oracle import binding as ob
    from oracle import lm_reference as lmr
    m = util.pkg()
    world = 2
    uid = m.nccl_unique_id()
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    procs = [ctx.Process(target=_lm_worker, args=(r, world, uid, out)) for r in range(world)]
    for pr in procs:
        pr.start()
    for pr in procs:
        pr.join(300)
        assert pr.exitcode == 0
    p = m.SyntheticBAL(31, 3000, 14000, seed=7, rotation_sigma=2e-3, translation_sigma=1.0, point_sigma=1.0)
    o = ob.Oracle(util.oracle_structure(p.structure), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI,
                  max_num_iterations=500)
    want = lmr.minimize(p, p.cameras, p.points, lambda v, r, D, e: o.solve(v, r, D, e, -1.0), eta=1e-2,
                        max_num_iterations=6)
    pts_all = np.full(3 * p.P, np.nan)
    for r in range(world):
        initial, final, its, cam, pts, pb, pe = out[r]
        assert abs(initial - want["initial_cost"]) <= 1e-12 * want["initial_cost"]
        assert abs(final - want["final_cost"]) <= 1e-8 * want["final_cost"]
        assert len(its) == len(want["iterations"])
        for (it, ls, cost), b in zip(its, want["iterations"]):
            assert it == b["iteration"] and ls == b["linear_solver_iterations"]
            assert abs(cost - b["cost"]) <= 1e-8 * b["cost"]
        assert util.rel_err(cam, want["cameras"]) < 1e-7
        pts_all[3 * pb:3 * pe] = pts[3 * pb:3 * pe]
    assert util.rel_err(pts_all, want["points"]) < 1e-7


def _explicit_worker(rank, world, uid, out):
    m = util.pkg()
    p = m.SyntheticBAL.named("ladybug", scale=0.05)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    g = m.IterativeSchurComplementSolver(p.structure, preconditioner_type=m.SCHUR_JACOBI, max_num_iterations=300,
                                         device_id=rank, rank=rank, world_size=world, nccl_unique_id=uid,
                                         use_explicit_schur_complement=True)
    x = np.full(p.structure.num_cols, np.nan)
    x, s = g.solve(values, residuals, D, 1e-3, -1.0, x=x)
    cells, rhs = g.explicit_eliminate()
    sx = g.explicit_multiply(np.arange(9 * p.C, dtype=np.float64) / (9 * p.C))
    out[rank] = (x, s["num_iterations"], s["termination_type"], g.point_begin, g.point_end, cells, rhs, sx)
    g.close()


def test_two_gpu_explicit_schur_matches_oracle():
    """use_explicit_schur_complement with the points sharded over two GPUs: every rank forms the outer
    products of its own points, the cells are summed over the ranks (NCCL), PCG on the explicit matrix
    runs replicated and bit-identical on both."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from oracle import binding as ob
    m = util.pkg()
    world = 2
    uid = m.nccl_unique_id()
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    procs = [ctx.Process(target=_explicit_worker, args=(r, world, uid, out)) for r in range(world)]
    for pr in procs:
        pr.start()
    for pr in procs:
        pr.join(300)
        assert pr.exitcode == 0
    p = m.SyntheticBAL.named("ladybug", scale=0.05)
    values, residuals, D, _, _ = util.lm_like_problem(p)
    o = ob.Oracle(util.oracle_structure(p.structure), num_threads=8, preconditioner_type=ob.SCHUR_JACOBI,
                  max_num_iterations=300)
    xo, so = o.explicit_solve(values, residuals, D, 1e-3, -1.0)
    _, _, _, want_cells, want_rhs = o.explicit_eliminate(values, residuals, D)
    sxo = o.explicit_multiply(np.arange(9 * p.C, dtype=np.float64) / (9 * p.C))
    ne = 3 * p.P
    for r in range(world):
        x, its, term, pb, pe, cells, rhs, sx = out[r]
        assert its == so["num_iterations"] and term == so["termination_type"]
        assert util.rel_err(cells.reshape(-1), want_cells) < 1e-12
        assert util.rel_err(rhs, want_rhs) < 1e-12
        assert util.rel_err(sx, sxo) < 1e-12
        assert util.rel_err(x[ne:], xo[ne:]) < 1e-9
        assert util.rel_err(x[3 * pb:3 * pe], xo[3 * pb:3 * pe]) < 1e-9
    # replicated PCG: both ranks hold the same camera solution bit for bit
    assert np.array_equal(out[0][0][ne:], out[1][0][ne:])
    assert np.array_equal(out[0][5], out[1][5])
