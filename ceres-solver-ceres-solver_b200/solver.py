"""ctypes binding of libceres_b200_schur.so and a host-side mirror of the
reference's solver interface for the ITERATIVE_SCHUR path.

`IterativeSchurComplementSolver` mirrors
internal/ceres/iterative_schur_complement_solver.h (Solve(A, b, per_solve_options, x)
-> Summary) on top of the C ABI in include/ceres_b200_schur.h. There is no CPU
fallback: without the CUDA library (or a GPU) construction raises.
"""
import ctypes
import os

import numpy as np

from . import _build

IDENTITY, JACOBI, SCHUR_JACOBI, SCHUR_POWER_SERIES_EXPANSION = 0, 1, 2, 3
TERMINATION = {0: "SUCCESS", 1: "NO_CONVERGENCE", 2: "FAILURE", 3: "FATAL_ERROR"}

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int)


class CStructure(ctypes.Structure):
    _fields_ = [("num_row_blocks", ctypes.c_int), ("row_block_size", _ip), ("row_block_pos", _ip),
                ("cell_ptr", _ip), ("cell_block_id", _ip), ("cell_position", _ip),
                ("num_col_blocks", ctypes.c_int), ("col_block_size", _ip), ("col_block_pos", _ip)]


class COptions(ctypes.Structure):
    _fields_ = [("num_eliminate_blocks", ctypes.c_int), ("preconditioner_type", ctypes.c_int),
                ("min_num_iterations", ctypes.c_int), ("max_num_iterations", ctypes.c_int),
                ("residual_reset_period", ctypes.c_int), ("use_spse_initialization", ctypes.c_int),
                ("max_num_spse_iterations", ctypes.c_int), ("spse_tolerance", ctypes.c_double),
                ("device_id", ctypes.c_int), ("rank", ctypes.c_int), ("world_size", ctypes.c_int),
                ("nccl_unique_id", ctypes.c_char_p), ("y_replicas", ctypes.c_int),
                ("status_check_period", ctypes.c_int), ("use_explicit_schur_complement", ctypes.c_int),
                ("explicit_switch_iterations", ctypes.c_int), ("num_gpus", ctypes.c_int),
                ("device_ids", ctypes.c_int * 16), ("deterministic", ctypes.c_int),
                ("cuda_graphs", ctypes.c_int), ("h2d_chunks", ctypes.c_int)]


class CSummary(ctypes.Structure):
    _fields_ = [("residual_norm", ctypes.c_double), ("num_iterations", ctypes.c_int),
                ("termination_type", ctypes.c_int), ("message", ctypes.c_char * 512),
                ("h2d_ms", ctypes.c_float), ("setup_ms", ctypes.c_float), ("pcg_ms", ctypes.c_float),
                ("back_substitute_ms", ctypes.c_float), ("d2h_ms", ctypes.c_float),
                ("kernel_launches", ctypes.c_int), ("total_ms", ctypes.c_float)]


class CLmOptions(ctypes.Structure):
    _fields_ = [("max_num_iterations", ctypes.c_int), ("initial_trust_region_radius", ctypes.c_double),
                ("max_trust_region_radius", ctypes.c_double), ("min_trust_region_radius", ctypes.c_double),
                ("min_relative_decrease", ctypes.c_double), ("min_lm_diagonal", ctypes.c_double),
                ("max_lm_diagonal", ctypes.c_double), ("eta", ctypes.c_double),
                ("function_tolerance", ctypes.c_double), ("gradient_tolerance", ctypes.c_double),
                ("parameter_tolerance", ctypes.c_double), ("jacobi_scaling", ctypes.c_int),
                ("max_num_consecutive_invalid_steps", ctypes.c_int)]


class CLmIteration(ctypes.Structure):
    _fields_ = [("iteration", ctypes.c_int), ("step_is_valid", ctypes.c_int), ("step_is_successful", ctypes.c_int),
                ("cost", ctypes.c_double), ("cost_change", ctypes.c_double), ("gradient_max_norm", ctypes.c_double),
                ("gradient_norm", ctypes.c_double), ("step_norm", ctypes.c_double),
                ("relative_decrease", ctypes.c_double), ("trust_region_radius", ctypes.c_double),
                ("eta", ctypes.c_double), ("linear_solver_iterations", ctypes.c_int),
                ("iteration_ms", ctypes.c_float)]


class CLmSummary(ctypes.Structure):
    _fields_ = [("initial_cost", ctypes.c_double), ("final_cost", ctypes.c_double), ("num_iterations", ctypes.c_int),
                ("num_successful_steps", ctypes.c_int), ("num_unsuccessful_steps", ctypes.c_int),
                ("termination_type", ctypes.c_int), ("message", ctypes.c_char * 256),
                ("total_ms", ctypes.c_float), ("evaluate_ms", ctypes.c_float), ("linear_solver_ms", ctypes.c_float),
                ("num_linear_solver_iterations", ctypes.c_int)]


LM_DEFAULTS = dict(max_num_iterations=50, initial_trust_region_radius=1e4, max_trust_region_radius=1e16,
                   min_trust_region_radius=1e-32, min_relative_decrease=1e-3, min_lm_diagonal=1e-6,
                   max_lm_diagonal=1e32, eta=1e-1, function_tolerance=1e-6, gradient_tolerance=1e-10,
                   parameter_tolerance=1e-8, jacobi_scaling=1, max_num_consecutive_invalid_steps=5)
LM_TERMINATION = {0: "CONVERGENCE", 1: "NO_CONVERGENCE", 2: "FAILURE"}

# every symbol include/ceres_b200_schur.h declares: name -> (restype, argtypes)
_H = ctypes.c_void_p
SYMBOLS = {
    "cb200_last_error": (ctypes.c_char_p, []),
    "cb200_version": (ctypes.c_int, []),
    "cb200_nccl_unique_id": (ctypes.c_int, [ctypes.c_char_p]),
    "cb200_create": (ctypes.c_int, [ctypes.POINTER(CStructure), ctypes.POINTER(COptions),
                                    ctypes.POINTER(_H)]),
    "cb200_destroy": (None, [_H]),
    "cb200_num_rows": (ctypes.c_int, [_H]),
    "cb200_num_cols": (ctypes.c_int, [_H]),
    "cb200_num_cols_e": (ctypes.c_int, [_H]),
    "cb200_num_cols_f": (ctypes.c_int, [_H]),
    "cb200_shard": (ctypes.c_int, [_H, _ip, _ip, _ip, _ip]),
    "cb200_shard_plan": (ctypes.c_int, [ctypes.POINTER(CStructure), ctypes.c_int, ctypes.c_int, _ip, _ip]),
    "cb200_explicit_plan": (ctypes.c_int, [ctypes.POINTER(CStructure), ctypes.c_int, _ip, _ip, _ip]),
    "cb200_explicit_num_cells": (ctypes.c_int, [_H]),
    "cb200_explicit_eliminate": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_explicit_multiply": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_solve": (ctypes.c_int, [_H, _dp, _dp, _dp, ctypes.c_double, ctypes.c_double, _dp,
                                   ctypes.POINTER(CSummary)]),
    "cb200_upload": (ctypes.c_int, [_H, _dp, _dp, _dp]),
    "cb200_solve_resident": (ctypes.c_int, [_H, ctypes.c_double, ctypes.c_double,
                                            ctypes.POINTER(CSummary)]),
    "cb200_download_solution": (ctypes.c_int, [_H, _dp]),
    "cb200_clear_diagonal": (ctypes.c_int, [_H]),
    "cb200_set_iterations": (ctypes.c_int, [_H, ctypes.c_int, ctypes.c_int]),
    "cb200_set_trace": (ctypes.c_int, [_H, _dp, ctypes.c_int]),
    "cb200_pmv_right_e": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_pmv_right_f": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_pmv_left_e": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_pmv_left_f": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_isc_init": (ctypes.c_int, [_H]),
    "cb200_isc_ete_inverse": (ctypes.c_int, [_H, _dp]),
    "cb200_isc_ftf_inverse": (ctypes.c_int, [_H, _dp]),
    "cb200_isc_rhs": (ctypes.c_int, [_H, _dp]),
    "cb200_isc_multiply": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_isc_pse_operator": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_isc_back_substitute": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_schur_jacobi_update": (ctypes.c_int, [_H, ctypes.c_int, _dp]),
    "cb200_schur_jacobi_apply": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_pse_apply": (ctypes.c_int, [_H, ctypes.c_int, ctypes.c_double, _dp, _dp]),
    "cb200_block_diagonal_apply_static": (ctypes.c_int, [_dp, _dp, _dp, ctypes.c_int]),
    "cb200_time_op": (ctypes.c_int, [_H, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                     ctypes.POINTER(ctypes.c_float), _ip]),
    "cb200_bal_attach": (ctypes.c_int, [_H, _dp]),
    "cb200_bal_set_parameters": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_bal_get_parameters": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_bal_evaluate": (ctypes.c_int, [_H, ctypes.c_int, _dp, _dp]),
    "cb200_download_jacobian": (ctypes.c_int, [_H, _dp, _dp]),
    "cb200_host_register": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t]),
    "cb200_host_unregister": (ctypes.c_int, [ctypes.c_void_p]),
    "cb200_lm_minimize": (ctypes.c_int, [_H, ctypes.POINTER(CLmOptions), ctypes.POINTER(CLmSummary),
                                         ctypes.POINTER(CLmIteration), ctypes.c_int]),
}

_LIB = None


def lib():
    """Load libceres_b200_schur.so (building it if the sources are newer)."""
    global _LIB
    if _LIB is None:
        # CB200_LIBRARY: a differently-built copy of the same library (kernel experiments, tools/build_variant.py)
        path = os.environ.get("CB200_LIBRARY") or _build.build_cuda()
        L = ctypes.CDLL(path)
        for name, (res, args) in SYMBOLS.items():
            f = getattr(L, name)  # AttributeError if the library lacks a declared symbol
            f.restype = res
            f.argtypes = args
        _LIB = L
    return _LIB


class B200Error(RuntimeError):
    pass


def host_register(array):
    """Page-lock a numpy array in place (cb200_host_register): uploads from it then run at PCIe rate."""
    _check(lib().cb200_host_register(array.ctypes.data, array.nbytes))


def host_unregister(array):
    _check(lib().cb200_host_unregister(array.ctypes.data))


def _check(rc):
    if rc != 0:
        raise B200Error(f"cb200 error {rc}: {lib().cb200_last_error().decode()}")


def _d(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"], "need contiguous float64"
    return a.ctypes.data_as(_dp)


def _i(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(_ip)


class BlockStructure:
    """Flattened CompressedRowBlockStructure (block_structure.h:52-96)."""

    def __init__(self, row_size, row_pos, cell_ptr, cell_block_id, cell_position, col_size, col_pos,
                 num_eliminate_blocks):
        a = lambda v: np.ascontiguousarray(v, dtype=np.int32)
        self.row_size, self.row_pos, self.cell_ptr = a(row_size), a(row_pos), a(cell_ptr)
        self.cell_block_id, self.cell_position = a(cell_block_id), a(cell_position)
        self.col_size, self.col_pos = a(col_size), a(col_pos)
        self.num_eliminate_blocks = int(num_eliminate_blocks)

    @property
    def num_rows(self):
        return int(self.row_pos[-1] + self.row_size[-1]) if len(self.row_size) else 0

    @property
    def num_cols(self):
        return int(self.col_pos[-1] + self.col_size[-1])

    @property
    def num_cols_e(self):
        return int(self.col_size[:self.num_eliminate_blocks].sum())

    @property
    def num_cols_f(self):
        return self.num_cols - self.num_cols_e

    def c_struct(self):
        return CStructure(len(self.row_size), _i(self.row_size), _i(self.row_pos), _i(self.cell_ptr),
                          _i(self.cell_block_id), _i(self.cell_position), len(self.col_size),
                          _i(self.col_size), _i(self.col_pos))

    @staticmethod
    def bal(point_row_start, cam_idx, num_cameras, interleaved=False):
        """<2,3,9> structure of a bundle-adjustment Jacobian whose rows are grouped by
        point. Default value layout = the evaluator's (all E cells, then all F cells,
        block_jacobian_writer.cc:102-165); interleaved=True gives [E_r F_r] per row
        (fake_bundle_adjustment_jacobian.cc:82-88)."""
        prs = np.asarray(point_row_start, dtype=np.int64)
        cam = np.asarray(cam_idx, dtype=np.int64)
        P, N, C = len(prs) - 1, len(cam), int(num_cameras)
        rows = np.arange(N, dtype=np.int64)
        row_point = np.repeat(np.arange(P, dtype=np.int64), np.diff(prs))
        cell_block_id = np.empty(2 * N, dtype=np.int64)
        cell_block_id[0::2] = row_point
        cell_block_id[1::2] = P + cam
        cell_position = np.empty(2 * N, dtype=np.int64)
        if interleaved:
            cell_position[0::2] = 24 * rows
            cell_position[1::2] = 24 * rows + 6
        else:
            cell_position[0::2] = 6 * rows
            cell_position[1::2] = 6 * N + 18 * rows
        col_size = np.concatenate([np.full(P, 3), np.full(C, 9)])
        col_pos = np.concatenate([[0], np.cumsum(col_size)[:-1]])
        return BlockStructure(np.full(N, 2), 2 * rows, 2 * np.arange(N + 1), cell_block_id,
                              cell_position, col_size, col_pos, P)


class IterativeSchurComplementSolver:
    """Device-resident ITERATIVE_SCHUR solver for one <2,3,9> block structure.

    Mirrors IterativeSchurComplementSolver (iterative_schur_complement_solver.cc:55-157):
    construct once per Solve() with the LinearSolver::Options fields, call solve() once per
    LM iteration with new values / b / D."""

    def __init__(self, structure, preconditioner_type=SCHUR_JACOBI, min_num_iterations=0,
                 max_num_iterations=500, residual_reset_period=10, use_spse_initialization=False,
                 max_num_spse_iterations=5, spse_tolerance=0.1, device_id=0, rank=0, world_size=1,
                 nccl_unique_id=None, y_replicas=0, status_check_period=0,
                 use_explicit_schur_complement=False, explicit_switch_iterations=0, num_gpus=0, device_ids=None,
                 deterministic=False, cuda_graphs=0, h2d_chunks=0):
        self.L = lib()
        self.structure = structure
        cs = structure.c_struct()
        opt = COptions(structure.num_eliminate_blocks, preconditioner_type, min_num_iterations,
                       max_num_iterations, residual_reset_period, int(use_spse_initialization),
                       max_num_spse_iterations, spse_tolerance, device_id, rank, world_size,
                       nccl_unique_id, y_replicas, status_check_period, int(use_explicit_schur_complement),
                       explicit_switch_iterations, num_gpus, (ctypes.c_int * 16)(*(device_ids or [])),
                       int(deterministic), cuda_graphs, h2d_chunks)
        h = _H()
        _check(self.L.cb200_create(ctypes.byref(cs), ctypes.byref(opt), ctypes.byref(h)))
        self.h = h
        self.num_rows = self.L.cb200_num_rows(h)
        self.num_cols = self.L.cb200_num_cols(h)
        self.num_cols_e = self.L.cb200_num_cols_e(h)
        self.num_cols_f = self.L.cb200_num_cols_f(h)
        v = [ctypes.c_int() for _ in range(4)]
        _check(self.L.cb200_shard(h, *[ctypes.byref(x) for x in v]))
        self.point_begin, self.point_end, self.row_begin, self.row_end = [x.value for x in v]
        self._trace = None

    def close(self):
        if getattr(self, "h", None):
            self.L.cb200_destroy(self.h)
            self.h = None

    __del__ = close

    @staticmethod
    def _summary(s):
        return {"num_iterations": s.num_iterations, "termination_type": s.termination_type,
                "message": s.message.decode(), "h2d_ms": s.h2d_ms, "setup_ms": s.setup_ms,
                "pcg_ms": s.pcg_ms, "back_substitute_ms": s.back_substitute_ms, "d2h_ms": s.d2h_ms,
                "kernel_launches": s.kernel_launches, "total_ms": s.total_ms}

    # --- LinearSolver::Solve ---------------------------------------------------------------
    def solve(self, values, b, D, q_tolerance=0.0, r_tolerance=0.0, x=None):
        if x is None:
            x = np.zeros(self.num_cols)
        s = CSummary()
        _check(self.L.cb200_solve(self.h, _d(values), _d(b), _d(D), q_tolerance, r_tolerance, _d(x),
                                  ctypes.byref(s)))
        out = self._summary(s)
        if self._trace is not None:
            out["trace"] = self._trace[:min(len(self._trace), max(s.num_iterations, 0))]
        return x, out

    def upload(self, values=None, b=None, D=None):
        _check(self.L.cb200_upload(self.h, _d(values), _d(b), _d(D)))

    def upload_ptr(self, values_ptr=0, b_ptr=0, D_ptr=0):
        """Same as upload() with raw host addresses (e.g. pinned torch tensors' data_ptr())."""
        c = lambda p: ctypes.cast(ctypes.c_void_p(p), _dp) if p else None
        _check(self.L.cb200_upload(self.h, c(values_ptr), c(b_ptr), c(D_ptr)))

    def solve_ptr(self, values_ptr, b_ptr, D_ptr, q_tolerance, r_tolerance, x_ptr):
        c = lambda p: ctypes.cast(ctypes.c_void_p(p), _dp) if p else None
        s = CSummary()
        _check(self.L.cb200_solve(self.h, c(values_ptr), c(b_ptr), c(D_ptr), q_tolerance,
                                  r_tolerance, c(x_ptr), ctypes.byref(s)))
        return self._summary(s)

    def clear_diagonal(self):
        _check(self.L.cb200_clear_diagonal(self.h))

    def solve_resident(self, q_tolerance=0.0, r_tolerance=0.0):
        s = CSummary()
        _check(self.L.cb200_solve_resident(self.h, q_tolerance, r_tolerance, ctypes.byref(s)))
        return self._summary(s)

    def download_solution(self, x=None):
        if x is None:
            x = np.zeros(self.num_cols)
        _check(self.L.cb200_download_solution(self.h, _d(x)))
        return x

    def set_iterations(self, min_num_iterations, max_num_iterations):
        _check(self.L.cb200_set_iterations(self.h, min_num_iterations, max_num_iterations))

    def set_trace(self, max_trace):
        self._trace = np.zeros((max_trace, self.num_cols_f)) if max_trace else None
        _check(self.L.cb200_set_trace(self.h, _d(self._trace), max_trace))

    # --- use_explicit_schur_complement seams -----------------------------------------------
    def explicit_eliminate(self):
        """(lhs cell values [num_cells x 81], reduced rhs) of SchurEliminator::Eliminate on the
        resident values / b / D; the cell structure is explicit_plan(structure)."""
        n = self.L.cb200_explicit_num_cells(self.h)
        vals, rhs = np.zeros((n, 81)), np.zeros(self.num_cols_f)
        _check(self.L.cb200_explicit_eliminate(self.h, _d(vals), _d(rhs)))
        return vals, rhs

    def explicit_multiply(self, x):
        y = np.zeros(self.num_cols_f)
        _check(self.L.cb200_explicit_multiply(self.h, _d(np.ascontiguousarray(x, dtype=np.float64)), _d(y)))
        return y

    # --- finer-grained seams ---------------------------------------------------------------
    def pmv(self, op, x, y):
        f = [self.L.cb200_pmv_right_e, self.L.cb200_pmv_right_f, self.L.cb200_pmv_left_e,
             self.L.cb200_pmv_left_f][op]
        _check(f(self.h, _d(x), _d(y)))
        return y

    def isc_init(self):
        _check(self.L.cb200_isc_init(self.h))

    def ete_inverse(self):
        out = np.zeros(9 * (self.point_end - self.point_begin))
        _check(self.L.cb200_isc_ete_inverse(self.h, _d(out)))
        return out

    def ftf_inverse(self):
        out = np.zeros(9 * self.num_cols_f)
        _check(self.L.cb200_isc_ftf_inverse(self.h, _d(out)))
        return out

    def isc_rhs(self):
        out = np.zeros(self.num_cols_f)
        _check(self.L.cb200_isc_rhs(self.h, _d(out)))
        return out

    def isc_multiply(self, x):
        y = np.zeros(self.num_cols_f)
        _check(self.L.cb200_isc_multiply(self.h, _d(x), _d(y)))
        return y

    def isc_pse_operator(self, x, y):
        _check(self.L.cb200_isc_pse_operator(self.h, _d(x), _d(y)))
        return y

    def isc_back_substitute(self, z):
        y = np.zeros(self.num_cols)
        _check(self.L.cb200_isc_back_substitute(self.h, _d(z), _d(y)))
        return y

    def schur_jacobi(self, invert=True):
        out = np.zeros(9 * self.num_cols_f)
        _check(self.L.cb200_schur_jacobi_update(self.h, int(invert), _d(out)))
        return out

    def schur_jacobi_apply(self, x, y):
        _check(self.L.cb200_schur_jacobi_apply(self.h, _d(x), _d(y)))
        return y

    def pse_apply(self, x, max_num_spse_iterations, spse_tolerance):
        y = np.zeros(self.num_cols_f)
        _check(self.L.cb200_pse_apply(self.h, max_num_spse_iterations, spse_tolerance, _d(x), _d(y)))
        return y

    # --- widened boundary: device evaluator + trust-region loop -------------
    def bal_attach(self, observations):
        _check(self.L.cb200_bal_attach(self.h, _d(np.ascontiguousarray(observations, dtype=np.float64))))

    def bal_set_parameters(self, cameras, points):
        _check(self.L.cb200_bal_set_parameters(self.h, _d(np.ascontiguousarray(cameras, dtype=np.float64)),
                                               _d(np.ascontiguousarray(points, dtype=np.float64))))

    def bal_get_parameters(self):
        cameras = np.zeros(self.num_cols_f)
        points = np.full(self.num_cols - self.num_cols_f, np.nan)
        _check(self.L.cb200_bal_get_parameters(self.h, _d(cameras), _d(points)))
        return cameras, points

    def bal_evaluate(self, jacobian=True):
        cost = ctypes.c_double()
        gradient = np.full(self.num_cols, np.nan) if jacobian else None
        _check(self.L.cb200_bal_evaluate(self.h, int(jacobian), ctypes.byref(cost), _d(gradient)))
        return cost.value, gradient

    def download_jacobian(self):
        values = np.full(12 * self.num_rows, np.nan)
        residuals = np.full(self.num_rows, np.nan)
        _check(self.L.cb200_download_jacobian(self.h, _d(values), _d(residuals)))
        return values, residuals

    def lm_minimize(self, **options):
        o = dict(LM_DEFAULTS)
        o.update(options)
        co = CLmOptions(**o)
        cap = o["max_num_iterations"] + 2
        its = (CLmIteration * cap)()
        cs = CLmSummary()
        _check(self.L.cb200_lm_minimize(self.h, ctypes.byref(co), ctypes.byref(cs), its, cap))
        out = {k: getattr(cs, k) for k, _ in CLmSummary._fields_}
        out["message"] = cs.message.decode()
        out["termination_type"] = LM_TERMINATION[cs.termination_type]
        out["iterations"] = [{k: getattr(its[i], k) for k, _ in CLmIteration._fields_}
                             for i in range(cs.num_iterations)]
        return out

    def time_op(self, op, warmup=3, iters=10):
        ms, n = ctypes.c_float(), ctypes.c_int()
        _check(self.L.cb200_time_op(self.h, op, warmup, iters, ctypes.byref(ms), ctypes.byref(n)))
        return ms.value, n.value


def block_diagonal_apply_static(blocks, x, y):
    """y += B x through cb200_block_diagonal_apply_static (no handle, no allocation; <= 16384 blocks)."""
    _check(lib().cb200_block_diagonal_apply_static(_d(blocks), _d(x), _d(y), len(x) // 9))
    return y


def shard_plan(structure, world_size):
    """(point_bounds, row_bounds), world_size + 1 entries each: the sharding cb200_create applies."""
    pb = np.zeros(world_size + 1, dtype=np.int32)
    rb = np.zeros(world_size + 1, dtype=np.int32)
    cs = structure.c_struct()
    _check(lib().cb200_shard_plan(ctypes.byref(cs), structure.num_eliminate_blocks, world_size, _i(pb), _i(rb)))
    return pb, rb


def explicit_plan(structure):
    """(row_ptr, cols): block CRS of the upper triangle of the explicit Schur complement
    (SparseSchurComplementSolver::InitStorage), host only."""
    cs = structure.c_struct()
    n = ctypes.c_int()
    _check(lib().cb200_explicit_plan(ctypes.byref(cs), structure.num_eliminate_blocks, ctypes.byref(n), None, None))
    C = len(structure.col_size) - structure.num_eliminate_blocks
    row_ptr = np.zeros(C + 1, dtype=np.int32)
    cols = np.zeros(n.value, dtype=np.int32)
    _check(lib().cb200_explicit_plan(ctypes.byref(cs), structure.num_eliminate_blocks, None, _i(row_ptr), _i(cols)))
    return row_ptr, cols


def nccl_unique_id():
    buf = ctypes.create_string_buffer(128)
    _check(lib().cb200_nccl_unique_id(buf))
    return buf.raw
