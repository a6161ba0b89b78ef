"""BAL problem files: csrc/bal_io.h against the numpy
restatement of examples/bal_problem.cc, and the bundle_adjuster CLI's flag handling."""
import os
import subprocess

import numpy as np
import pytest

from tests import util
from oracle import bal_reference as br


@pytest.fixture(scope="module")
def small():
    m = util.pkg()
    p = m.SyntheticBAL(7, 60, 260, seed=11)
    return m, p, m.BALFile.from_problem(p)


def test_write_read_round_trip(small, tmp_path):
    m, p, f = small
    path = tmp_path / "problem.txt"
    f.write(path)
    head = open(path).readline().split()
    assert head == ["7", "60", "260"]  # "C P N" header (bal_problem.cc:83-85)
    g = m.BALFile.read(path)
    assert (g.camera_index == f.camera_index).all() and (g.point_index == f.point_index).all()
    # parameters are written with %.16g, observations with %g (bal_problem.cc:155-182)
    assert util.rel_err(g.cameras, f.cameras) < 1e-15 and util.rel_err(g.points, f.points) < 1e-15
    assert np.abs(g.observations - f.observations).max() <= 5e-6 * np.abs(f.observations).max()
    q = g.problem()
    assert (q.point_row_start == p.point_row_start).all() and (q.cam_idx == p.cam_idx).all()


def test_group_by_point_is_stable(small):
    m, p, f = small
    rng = np.random.default_rng(5)
    perm = rng.permutation(f.num_observations)  # a file that lists its observations in any order
    g = m.BALFile(f.camera_index[perm], f.point_index[perm], f.observations.reshape(-1, 2)[perm].reshape(-1),
                  f.cameras, f.points)
    q = g.problem()
    assert (q.point_row_start == p.point_row_start).all()
    for j in range(p.P):  # the rows of a point keep their file order
        rows = slice(q.point_row_start[j], q.point_row_start[j + 1])
        assert (q.cam_idx[rows] == g.camera_index[g.point_index == j]).all()
    c0, _, _ = p.evaluate(jacobian=False)
    c1, _, _ = q.evaluate(jacobian=False)
    assert abs(c0 - c1) <= 1e-12 * c0


def test_normalize_matches_reference_restatement(small):
    m, p, f = small
    g = m.BALFile(f.camera_index, f.point_index, f.observations, f.cameras.copy(), f.points.copy())
    g.normalize()
    cams, pts = br.normalize(f.cameras, f.points)
    assert util.rel_err(g.points, pts.reshape(-1)) < 1e-14
    assert util.rel_err(g.cameras, cams.reshape(-1)) < 1e-13
    # median absolute deviation is 100 afterwards, and the reprojections are unchanged up to the scale
    med = np.array([br.upper_median(pts[:, i]) for i in range(3)])
    assert abs(br.upper_median(np.abs(pts - med).sum(axis=1)) - 100.0) < 1e-9
    c0, _, _ = p.evaluate(jacobian=False)
    c1, _, _ = g.problem().evaluate(jacobian=False)
    assert abs(c0 - c1) <= 1e-9 * c0  # the camera model is invariant to a similarity of the scene


def test_perturb(small):
    m, p, f = small
    def fresh():
        return m.BALFile(f.camera_index, f.point_index, f.observations, f.cameras.copy(), f.points.copy())
    a, b = fresh(), fresh()
    a.perturb(0.0, 0.0, 0.0)  # nothing drawn; cameras go through centre and back (bal_problem.cc:326-331)
    assert util.rel_err(a.cameras, f.cameras) < 1e-13 and (a.points == f.points).all()
    a, b = fresh(), fresh()
    a.perturb(0.01, 0.2, 0.5)
    b.perturb(0.01, 0.2, 0.5)
    assert (a.cameras == b.cameras).all() and (a.points == b.points).all()  # default-seeded mt19937
    d = (a.points - f.points)
    assert 0.35 < d.std() < 0.65 and abs(d.mean()) < 0.15
    # intrinsics untouched; rotation noise is drawn with POINT sigma (bal_problem.cc:318-319)
    ca, cf = a.cameras.reshape(-1, 9), f.cameras.reshape(-1, 9)
    assert (ca[:, 6:] == cf[:, 6:]).all()
    assert 0.2 < (ca[:, :3] - cf[:, :3]).std() < 0.9
    c = fresh()
    c.perturb(0.0, 0.0, 0.5)  # rotation_sigma == 0: rotations stay
    assert np.abs(c.cameras.reshape(-1, 9)[:, :3] - cf[:, :3]).max() < 1e-15
    assert (c.points == a.points).all()  # same first draws


def test_ply_and_errors(small, tmp_path):
    m, p, f = small
    f.write_ply(tmp_path / "a.ply")
    lines = open(tmp_path / "a.ply").read().splitlines()
    assert lines[0] == "ply" and lines[2] == "element vertex 67" and lines[9] == "end_header"
    assert len(lines) == 10 + 67 and lines[10].endswith("0 255 0") and lines[-1].endswith("255 255 255")
    centre = np.array(lines[10].split()[:3], float)
    assert np.allclose(centre, br.camera_center(f.cameras[:9]), rtol=1e-5, atol=1e-5)
    with pytest.raises(ValueError):
        m.BALFile.read(tmp_path / "missing.txt")
    (tmp_path / "bad.txt").write_text("2 3 4\n0 0 1.0 2.0\n0 7 1.0 2.0\n")
    with pytest.raises(ValueError, match="out of range"):
        m.BALFile.read(tmp_path / "bad.txt")
    (tmp_path / "short.txt").write_text("1 1 1\n0 0 1.0 2.0\n1 2 3\n")
    with pytest.raises(ValueError, match="camera parameters"):
        m.BALFile.read(tmp_path / "short.txt")


def test_cli_refuses_what_it_does_not_implement(small, tmp_path):
    m, p, f = small
    cli = m._build.build_cli()
    f.write(tmp_path / "p.txt")
    def run(*args):
        return subprocess.run([cli, *args], capture_output=True, text=True, timeout=60)
    r = run()
    assert r.returncode == 1 and "Usage: bundle_adjuster --input=bal_problem" in r.stderr
    for args, what in [(("--linear_solver=sparse_schur",), "iterative_schur"),
                       (("--preconditioner", "cluster_jacobi"), "not implemented"),
                       (("--robustify",), "not implemented"),
                       (("--trust_region_strategy=dogleg",), "levenberg_marquardt"),
                       (("--no_such_flag=1",), "unknown flag")]:
        r = run(f"--input={tmp_path / 'p.txt'}", *args)
        assert r.returncode == 1 and what in r.stderr, (args, r.stderr)
    r = run(f"--input={tmp_path / 'nope.txt'}")
    assert r.returncode == 1 and "unable to open file" in r.stderr
