"""CPU tests of the product's host side: the C-ABI library loads and exports every symbol
include/ceres_b200_schur.h declares, structure validation (DetectStructure rules,
detect_structure.cc:39-120) rejects what is not <2,3,9> before touching the GPU, there is no
CPU fallback, and the synthetic BAL generator produces the Jacobian of the Snavely model."""
import ctypes
import os
import re

import numpy as np
import pytest

REFERENCE_BLOCK_STRUCTURE = "/root/reference/internal/ceres/block_structure.h"

from tests import util
from tests.golden import kat_problems as kp


def test_library_exports_every_declared_symbol():
    m = util.pkg()
    L = m.solver.lib()
    header = open(os.path.join(util.ROOT, "include", "ceres_b200_schur.h")).read()
    declared = set(re.findall(r"\b(cb200_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    assert declared == set(m.solver.SYMBOLS), declared ^ set(m.solver.SYMBOLS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.cb200_version() >= 100


def test_unsupported_structures_are_rejected():
    m = util.pkg()
    # the reference's own KAT problems are scalar / mixed blocks: not <2,3,9>
    for pid in (2, 4, 6):
        st_kw, *_ = kp.flat(pid)
        st = m.BlockStructure(**st_kw)
        with pytest.raises(m.B200Error, match="cb200 error -2"):
            m.IterativeSchurComplementSolver(st)
    p = m.SyntheticBAL(5, 20, 60, geometry=False)
    st = p.structure
    with pytest.raises(m.B200Error, match="error -1"):
        m.IterativeSchurComplementSolver(st, preconditioner_type=7)  # CLUSTER_* not supported
    bad = m.BlockStructure.bal(p.point_row_start, p.cam_idx, 5)
    bad.cell_block_id[0::2] = bad.cell_block_id[0::2][::-1].copy()  # rows not grouped by point
    with pytest.raises(m.B200Error, match="error -2"):
        m.IterativeSchurComplementSolver(bad)


@pytest.mark.skipif(util.has_gpu(), reason="needs a machine without a GPU")
def test_no_cpu_fallback():
    m = util.pkg()
    p = m.SyntheticBAL(5, 20, 60, geometry=False)
    with pytest.raises(m.B200Error, match="error -3"):
        m.IterativeSchurComplementSolver(p.structure)


def test_synthetic_structure_shapes_and_determinism():
    m = util.pkg()
    p = m.SyntheticBAL(50, 2000, 9000, seed=1, dup_pairs=3)
    q = m.SyntheticBAL(50, 2000, 9000, seed=1, dup_pairs=3)
    assert np.array_equal(p.cam_idx, q.cam_idx) and np.array_equal(p.cameras, q.cameras)
    lens = np.diff(p.point_row_start)
    assert lens.sum() == 9000 and lens.min() >= 2 and p.cam_idx.max() < 50
    dup = sum(len(set(p.cam_idx[a:b])) != b - a for a, b in zip(p.point_row_start[:-1], p.point_row_start[1:]))
    assert dup == 3
    with pytest.raises(ValueError):
        m.SyntheticBAL(3, 10, 100)  # more observations than point-camera pairs


def test_synthetic_jacobian_is_the_derivative():
    m = util.pkg()
    p = m.SyntheticBAL(6, 40, 150, seed=2)
    cost, r0, values = p.evaluate()
    assert abs(cost - 0.5 * r0 @ r0) < 1e-9 * cost
    J = util.oracle_structure(p.structure).dense(values)
    rng = np.random.default_rng(0)
    d = rng.standard_normal(J.shape[1]) * 1e-6
    dp, dc = d[:3 * p.P], d[3 * p.P:]
    _, r1, _ = p.evaluate(cameras=p.cameras + dc, points=p.points + dp, jacobian=False)
    _, r2, _ = p.evaluate(cameras=p.cameras - dc, points=p.points - dp, jacobian=False)
    assert util.rel_err((r1 - r2) / 2, J @ d) < 1e-6
    # column helpers agree with the dense matrix
    assert util.rel_err(p.squared_column_norm(values), (J * J).sum(axis=0)) < 1e-13
    x = rng.standard_normal(J.shape[1])
    assert util.rel_err(p.jacobian_times(values, x), J @ x) < 1e-13
    v2 = p.scale_columns(values.copy(), 1.0 / (1.0 + x * x))
    assert util.rel_err(util.oracle_structure(p.structure).dense(v2), J / (1.0 + x * x)) < 1e-13


def test_adapter_header_flattens_a_ceres_block_structure(tmp_path):
    """include/ceres_b200_adapter.h (the reference-side binding of INTEGRATION.md) compiled against a
    stand-in for ceres/block_structure.h; the tile plan covers every row once; cb200_create is
    reached through the C ABI from C++ (and refuses to run without a GPU)."""
    import subprocess
    m = util.pkg()
    lib = m._build.build_cuda()
    exe = str(tmp_path / "adapter_driver")
    real = []
    if os.path.exists(REFERENCE_BLOCK_STRUCTURE):  # flatten the reference's own type where its header is present
        inc = tmp_path / "real_headers" / "ceres"
        inc.mkdir(parents=True, exist_ok=True)
        (inc / "block_structure.h").write_text(f'#include "{REFERENCE_BLOCK_STRUCTURE}"\n')
        real = ["-DCB200_REAL_BLOCK_STRUCTURE", "-I" + str(tmp_path / "real_headers"),
                "-I" + os.path.join(util.ROOT, "tests", "cpp", "standin_ceres")]
    subprocess.run(["g++", "-std=c++17", "-O1"] + real + ["-o", exe, os.path.join(util.ROOT, "tests", "cpp", "adapter_driver.cc"),
                    lib, "-Wl,-rpath," + os.path.dirname(lib)], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    kv = dict(t.split("=") for t in out.split())
    assert kv["flatten"] == "0" and kv["canonical"] == "1" and kv["bad"] == "0", out
    N = int(kv["N"])
    assert int(kv["rows_in_tiles"]) == N - 45 - 300 and kv["long"] == "1" and kv["empty"] == "1", out
    # the 45-row point: two first-pass chunks and two second-pass chunks in one group
    assert kv["chunk1"] == "45" and kv["chunk2"] == "45", out
    assert int(kv["groups"]) == int(kv["tiles"]) - 3, out
    sh = [int(t) for t in kv["shard"].split(",")]
    assert sh[0] == 0 and sh[3] == 40 and sh[0] <= sh[1] <= sh[2] <= sh[3], out
    assert kv["create"] == ("0" if util.has_gpu() else "-3"), out


@pytest.mark.parametrize("shape", [(6, 40, 150, 0), (31, 800, 3500, 4), (3, 50, 120, 2)])
def test_explicit_schur_block_structure_matches_init_storage(shape):
    """cb200_explicit_plan (host only) == SparseSchurComplementSolver::InitStorage +
    BlockRandomAccessSparseMatrix layout as restated by the oracle."""
    from oracle import binding as ob
    m = util.pkg()
    C, P, N, dup = shape
    p = m.SyntheticBAL(C, P, N, seed=3, dup_pairs=dup)
    cost, residuals, values = p.evaluate()
    row_ptr, cols = m.explicit_plan(p.structure)
    o = ob.Oracle(util.oracle_structure(p.structure))
    o_row_ptr, o_cols, o_pos, _, _ = o.explicit_eliminate(values, residuals, None)
    assert np.array_equal(row_ptr, o_row_ptr) and np.array_equal(cols, o_cols)
    assert np.array_equal(o_pos, 81 * np.arange(len(cols)))  # cell k lives at 81 k
    assert all(cols[row_ptr[i]] == i for i in range(C))  # diagonal cell first in every row


def test_explicit_schur_option_validation():
    """solver.cc:277-290: SCHUR_JACOBI only, no SPSE initialisation (checked before any device work)."""
    m = util.pkg()
    p = m.SyntheticBAL(5, 20, 60, geometry=False)
    with pytest.raises(m.B200Error, match="only supports SCHUR_JACOBI"):
        m.IterativeSchurComplementSolver(p.structure, preconditioner_type=m.JACOBI,
                                         use_explicit_schur_complement=True)
    with pytest.raises(m.B200Error, match="does not support use_spse_initialization"):
        m.IterativeSchurComplementSolver(p.structure, use_spse_initialization=True,
                                         use_explicit_schur_complement=True)


def build_adapter_class_driver(tmp_path):
    """The adapter class against stand-ins for linear_solver.h / block_sparse_matrix.h / absl's check.h (the
    include set of the v2.3.0 tree: abseil, no glog) and, where the reference tree is present (this
    container, not the GPU box), against the reference's OWN internal/ceres/block_structure.h (Eigen-free;
    only the generated export.h is stubbed) instead of the stand-in of the same name."""
    import subprocess
    m = util.pkg()
    lib = m._build.build_cuda()
    exe = str(tmp_path / "adapter_class_driver")
    standin = os.path.join(util.ROOT, "tests", "cpp", "standin_ceres")
    includes = []
    if os.path.exists(REFERENCE_BLOCK_STRUCTURE):
        real = tmp_path / "real_headers" / "ceres"
        real.mkdir(parents=True, exist_ok=True)
        (real / "block_structure.h").write_text(f'#include "{REFERENCE_BLOCK_STRUCTURE}"\n')
        includes.append("-I" + str(tmp_path / "real_headers"))
    includes.append("-I" + standin)
    subprocess.run(["g++", "-std=c++17", "-O1"] + includes +
                   ["-o", exe, os.path.join(util.ROOT, "tests", "cpp", "adapter_class_driver.cc"), lib,
                    "-Wl,-rpath," + os.path.dirname(lib)], check=True)
    return exe


@pytest.mark.skipif(util.has_gpu(), reason="needs a machine without a GPU")
def test_adapter_class_compiles_and_reports_fatal_error_without_a_gpu(tmp_path):
    """B200IterativeSchurComplementSolver (include/ceres_b200_adapter.h, the class INTEGRATION.md adds to
    linear_solver.cc) compiled against stand-ins for linear_solver.h / block_sparse_matrix.h and driven
    through LinearSolver::Solve: without a device it must come back with FATAL_ERROR and the library's
    message, never fall back to a CPU path."""
    import subprocess
    exe = build_adapter_class_driver(tmp_path)
    for mode in ("i", "e"):
        out = subprocess.run([exe, mode], check=True, capture_output=True, text=True).stdout
        assert "termination=3" in out and "no CUDA device" in out, out


def test_adapter_class_delegates_out_of_scope_structures(tmp_path):
    """A block structure outside the device path's scope (here a camera-prior row: one F cell, no E cell;
    iterative_schur_complement_solver.cc handles it, DetectStructure falls back to dynamic sizes) goes to the
    solver LinearSolver::Create returned before the patch, on the first call and on every later one; without
    a fallback it is FATAL_ERROR with the library's reason. Decided before any device work: runs anywhere."""
    import subprocess
    exe = build_adapter_class_driver(tmp_path)
    out = subprocess.run([exe, "u"], check=True, capture_output=True, text=True).stdout
    assert "delegated_calls=2" in out and "x0=42" in out and "termination=0" in out, out
    assert "delegated-to-reference-solver" in out, out
    out = subprocess.run([exe, "U"], check=True, capture_output=True, text=True).stdout
    assert "termination=3" in out and "cells" in out, out


def test_no_schur_complement_blocks_is_unsupported():
    """num_schur_complement_blocks == 0 (iterative_schur_complement_solver.cc:86-95; the shape of
    LinearLeastSquaresProblem 3: E blocks only): outside the <2,3,9> scope -> CB200_ERR_UNSUPPORTED_STRUCTURE
    from cb200_create before any device work (the adapter then delegates to the reference solver)."""
    m = util.pkg()
    P, N = 4, 8
    rows = np.arange(N)
    st = m.BlockStructure(np.full(N, 2), 2 * rows, np.arange(N + 1), rows // 2, 6 * rows, np.full(P, 3),
                          3 * np.arange(P), P)
    with pytest.raises(m.B200Error, match="cb200 error -2"):
        m.IterativeSchurComplementSolver(st)


def test_explicit_plan_invariants_host_only(tmp_path):
    """csrc/explicit_plan.h on a random structure with unsorted cameras, duplicate (point, camera) rows, empty
    points and 40-row points: block pairs == InitStorage's set, column lists cover every strictly-upper cell
    once, every (u <= v) pair of every point lands in the cell (camera u, camera v) with the duplicate flag
    exactly where two rows share a camera, and the work items cover every 16-row chunk once."""
    import subprocess
    exe = str(tmp_path / "explicit_plan_driver")
    subprocess.run(["g++", "-std=c++17", "-O1", "-pthread", "-o", exe,
                    os.path.join(util.ROOT, "tests", "cpp", "explicit_plan_driver.cc")], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    kv = dict(t.split("=") for t in out.split())
    assert int(kv["cells"]) > 37 and int(kv["dup_pairs"]) > 0 and int(kv["items"]) > 400, out
    for k in ("bad_structure", "bad_transpose", "bad_pairs", "bad_order", "bad_items"):
        assert kv[k] == "0", out


def test_inprocess_group_rendezvous(tmp_path):
    """The host rendezvous of a multi-GPU handle's shards (InProcGroup, csrc/multi_gpu.h) under real threads:
    barriers in lock step, rank-ordered sums bit-identical on every shard, Abort releases waiting shards."""
    import subprocess
    exe = str(tmp_path / "inproc_group_driver")
    subprocess.run(["g++", "-std=c++17", "-O1", "-pthread", "-o", exe,
                    os.path.join(util.ROOT, "tests", "cpp", "inproc_group_driver.cc")], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True, timeout=60).stdout
    assert out.strip() == "order_errors=0 differing=0 value_ok=1 released=2", out


def test_round2_option_validation_needs_no_device():
    """cb200_options added in round 2 are validated before any device work: num_gpus excludes world_size > 1,
    bounds on num_gpus / h2d_chunks, deterministic excludes the explicit matrix. (Without a GPU a VALID
    configuration then fails with CB200_ERR_CUDA: there is no CPU fallback.)"""
    m = util.pkg()
    p = m.SyntheticBAL(5, 20, 60, geometry=False)
    with pytest.raises(m.B200Error, match="excludes world_size"):
        m.IterativeSchurComplementSolver(p.structure, num_gpus=2, world_size=2, rank=0, nccl_unique_id=b"\0" * 128)
    with pytest.raises(m.B200Error, match="bad num_gpus"):
        m.IterativeSchurComplementSolver(p.structure, num_gpus=17)
    with pytest.raises(m.B200Error, match="bad h2d_chunks"):
        m.IterativeSchurComplementSolver(p.structure, h2d_chunks=1000)
    with pytest.raises(m.B200Error, match="deterministic mode covers the implicit operator only"):
        m.IterativeSchurComplementSolver(p.structure, deterministic=True, use_explicit_schur_complement=True)
    if not util.has_gpu():
        with pytest.raises(m.B200Error, match="cb200 error -3"):
            m.IterativeSchurComplementSolver(p.structure, num_gpus=2)
        with pytest.raises(m.B200Error, match="cb200 error -3"):
            m.block_diagonal_apply_static(np.zeros(81), np.zeros(9), np.zeros(9))
