"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_build/liboracle.so (the CPU restatement of the
reference's ITERATIVE_SCHUR path, see oracle/*.h for file:line citations).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product package never does.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = ["c_api.cc", "small_blas.h", "block_structure.h", "partitioned_view.h",
        "implicit_schur.h", "schur_jacobi.h", "solver.h", "explicit_schur.h"]

IDENTITY, JACOBI, SCHUR_JACOBI, SCHUR_POWER_SERIES_EXPANSION = 0, 1, 2, 3
TERMINATION = {0: "SUCCESS", 1: "NO_CONVERGENCE", 2: "FAILURE", 3: "FATAL_ERROR"}


class OracleSummary(ctypes.Structure):
    _fields_ = [("residual_norm", ctypes.c_double),
                ("num_iterations", ctypes.c_int),
                ("termination_type", ctypes.c_int),
                ("message", ctypes.c_char * 512)]


def build(native=False, force=False):
    """Compile the oracle. native=True builds a -march=native copy (used for the
    CPU baseline timing on the GPU box's host cores)."""
    out = os.path.join(_HERE, "_build",
                       "liboracle_native.so" if native else "liboracle.so")
    srcs = [os.path.join(_HERE, s) for s in _SRC]
    if (not force and os.path.exists(out)
            and all(os.path.getmtime(out) >= os.path.getmtime(s) for s in srcs)):
        return out
    os.makedirs(os.path.dirname(out), exist_ok=True)
    march = "native" if native else "x86-64-v3"
    cmd = ["g++", "-O3", f"-march={march}", "-std=c++17", "-fPIC", "-pthread",
           "-shared", "-o", out, os.path.join(_HERE, "c_api.cc")]
    subprocess.run(cmd, check=True, cwd=_HERE)
    return out


_LIBS = {}


def lib(native=False):
    if native in _LIBS:
        return _LIBS[native]
    path = build(native=native)
    L = ctypes.CDLL(path)
    dp = ctypes.POINTER(ctypes.c_double)
    ip = ctypes.POINTER(ctypes.c_int)
    L.oracle_create.restype = ctypes.c_void_p
    L.oracle_create.argtypes = [ctypes.c_int, ip, ip, ip, ip, ip, ctypes.c_int, ip,
                                ip, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                ctypes.c_int, ctypes.c_int, ctypes.c_double]
    L.oracle_destroy.argtypes = [ctypes.c_void_p]
    L.oracle_pmv.argtypes = [ctypes.c_void_p, dp, ctypes.c_int, dp, dp]
    L.oracle_block_diagonal.argtypes = [ctypes.c_void_p, dp, ctypes.c_int, dp]
    L.oracle_isc_init.argtypes = [ctypes.c_void_p, dp, dp, dp]
    L.oracle_isc_num_cols.argtypes = [ctypes.c_void_p]
    L.oracle_isc_rhs.argtypes = [ctypes.c_void_p, dp]
    L.oracle_isc_multiply.argtypes = [ctypes.c_void_p, dp, dp]
    L.oracle_isc_pse_operator.argtypes = [ctypes.c_void_p, dp, dp]
    L.oracle_isc_back_substitute.argtypes = [ctypes.c_void_p, dp, dp]
    L.oracle_isc_ete_inverse.argtypes = [ctypes.c_void_p, dp]
    L.oracle_isc_ftf_inverse.argtypes = [ctypes.c_void_p, dp]
    L.oracle_pse_apply.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_double,
                                   dp, dp]
    L.oracle_schur_jacobi.argtypes = [ctypes.c_void_p, dp, dp, ctypes.c_int, dp]
    L.oracle_solve.argtypes = [ctypes.c_void_p, dp, dp, dp, ctypes.c_double,
                               ctypes.c_double, dp, ctypes.POINTER(OracleSummary),
                               dp, ctypes.c_int]
    L.oracle_reduced_solution.argtypes = [ctypes.c_void_p, dp]
    L.oracle_explicit_eliminate.argtypes = [ctypes.c_void_p, dp, dp, dp, ip, ip, ip, ip, ip,
                                            ip, dp, dp]
    L.oracle_explicit_multiply.argtypes = [ctypes.c_void_p, dp, dp]
    L.oracle_explicit_solve.argtypes = [ctypes.c_void_p, dp, dp, dp, ctypes.c_double,
                                        ctypes.c_double, dp, ctypes.POINTER(OracleSummary),
                                        dp, ctypes.c_int]
    L.oracle_cg_dense.argtypes = [ctypes.c_int, dp, dp, dp, ctypes.c_int,
                                  ctypes.c_int, ctypes.c_int, ctypes.c_double,
                                  ctypes.c_double, ctypes.POINTER(OracleSummary)]
    L.oracle_llt_inverse_upper.argtypes = [dp, ctypes.c_int]
    L.oracle_inverse_small.argtypes = [dp, ctypes.c_int, dp]
    _LIBS[native] = L
    return L


def _dp(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _ip(a):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int))


class Structure:
    """Flattened CompressedRowBlockStructure (block_structure.h:52-96)."""

    def __init__(self, row_size, row_pos, cell_ptr, cell_block_id, cell_position,
                 col_size, col_pos, num_eliminate_blocks):
        a = lambda v: np.ascontiguousarray(v, dtype=np.int32)
        self.row_size, self.row_pos = a(row_size), a(row_pos)
        self.cell_ptr, self.cell_block_id = a(cell_ptr), a(cell_block_id)
        self.cell_position = a(cell_position)
        self.col_size, self.col_pos = a(col_size), a(col_pos)
        self.num_eliminate_blocks = int(num_eliminate_blocks)
        self.num_rows = int(self.row_pos[-1] + self.row_size[-1]) if len(self.row_size) else 0
        self.num_cols = int(self.col_pos[-1] + self.col_size[-1])
        ne = self.num_eliminate_blocks
        self.num_cols_e = int(self.col_size[:ne].sum())
        self.num_cols_f = self.num_cols - self.num_cols_e

    def dense(self, values):
        """Dense matrix of the block-sparse Jacobian (test helper)."""
        A = np.zeros((self.num_rows, self.num_cols))
        for r in range(len(self.row_size)):
            rs, rp = self.row_size[r], self.row_pos[r]
            for c in range(self.cell_ptr[r], self.cell_ptr[r + 1]):
                b = self.cell_block_id[c]
                cs, cp, pos = self.col_size[b], self.col_pos[b], self.cell_position[c]
                A[rp:rp + rs, cp:cp + cs] += values[pos:pos + rs * cs].reshape(rs, cs)
        return A


class Oracle:
    def __init__(self, st, num_threads=1, preconditioner_type=SCHUR_JACOBI,
                 min_num_iterations=0, max_num_iterations=500,
                 residual_reset_period=10, use_spse_initialization=False,
                 max_num_spse_iterations=5, spse_tolerance=0.1, native=False):
        self.L = lib(native)
        self.st = st
        self.h = self.L.oracle_create(
            len(st.row_size), _ip(st.row_size), _ip(st.row_pos), _ip(st.cell_ptr),
            _ip(st.cell_block_id), _ip(st.cell_position), len(st.col_size),
            _ip(st.col_size), _ip(st.col_pos), st.num_eliminate_blocks, num_threads,
            preconditioner_type, min_num_iterations, max_num_iterations,
            residual_reset_period, int(use_spse_initialization),
            max_num_spse_iterations, spse_tolerance)
        self._keep = []

    def __del__(self):
        if getattr(self, "h", None):
            self.L.oracle_destroy(self.h)
            self.h = None

    def pmv(self, values, op, x, y):
        self.L.oracle_pmv(self.h, _dp(values), op, _dp(x), _dp(y))
        return y

    def block_diagonal(self, values, which):
        n = self.L.oracle_block_diagonal(self.h, _dp(values), which, None)
        out = np.zeros(n)
        self.L.oracle_block_diagonal(self.h, _dp(values), which, _dp(out))
        return out

    def isc_init(self, values, D, b):
        self._keep = [values, D, b]
        self.L.oracle_isc_init(self.h, _dp(values), _dp(D), _dp(b))

    def isc_rhs(self):
        out = np.zeros(self.st.num_cols_f)
        self.L.oracle_isc_rhs(self.h, _dp(out))
        return out

    def isc_multiply(self, x):
        y = np.zeros(self.st.num_cols_f)
        self.L.oracle_isc_multiply(self.h, _dp(x), _dp(y))
        return y

    def isc_pse_operator(self, x, y):
        self.L.oracle_isc_pse_operator(self.h, _dp(x), _dp(y))
        return y

    def isc_back_substitute(self, z):
        y = np.zeros(self.st.num_cols)
        self.L.oracle_isc_back_substitute(self.h, _dp(z), _dp(y))
        return y

    def ete_inverse(self):
        n = self.L.oracle_isc_ete_inverse(self.h, None)
        out = np.zeros(n)
        self.L.oracle_isc_ete_inverse(self.h, _dp(out))
        return out

    def ftf_inverse(self):
        n = self.L.oracle_isc_ftf_inverse(self.h, None)
        out = np.zeros(n)
        self.L.oracle_isc_ftf_inverse(self.h, _dp(out))
        return out

    def pse_apply(self, x, max_num_spse_iterations, spse_tolerance):
        y = np.zeros(self.st.num_cols_f)
        self.L.oracle_pse_apply(self.h, max_num_spse_iterations, spse_tolerance,
                                _dp(x), _dp(y))
        return y

    def schur_jacobi(self, values, D, invert=True):
        n = self.L.oracle_schur_jacobi(self.h, _dp(values), _dp(D), int(invert), None)
        out = np.zeros(n)
        self.L.oracle_schur_jacobi(self.h, _dp(values), _dp(D), int(invert), _dp(out))
        return out

    def solve(self, values, b, D, q_tolerance=0.0, r_tolerance=0.0, max_trace=0,
              x_init=None):
        x = np.zeros(self.st.num_cols) if x_init is None else x_init.copy()
        s = OracleSummary()
        trace = np.zeros((max_trace, self.st.num_cols_f)) if max_trace else None
        self.L.oracle_solve(self.h, _dp(values), _dp(b), _dp(D), q_tolerance,
                            r_tolerance, _dp(x), ctypes.byref(s), _dp(trace),
                            max_trace)
        summary = {"num_iterations": s.num_iterations,
                   "termination_type": s.termination_type,
                   "message": s.message.decode()}
        if max_trace:
            summary["trace"] = trace[:min(max_trace, max(s.num_iterations, 0))]
        return x, summary

    # -- use_explicit_schur_complement = true (oracle/explicit_schur.h) --
    def explicit_eliminate(self, values, b, D):
        """Block CRS (upper triangle) of the explicit Schur complement + reduced rhs."""
        nb, nc, nv = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        self.L.oracle_explicit_eliminate(self.h, _dp(values), _dp(b), _dp(D), ctypes.byref(nb),
                                         ctypes.byref(nc), ctypes.byref(nv), None, None, None,
                                         None, None)
        row_ptr = np.zeros(nb.value + 1, dtype=np.int32)
        cols = np.zeros(nc.value, dtype=np.int32)
        pos = np.zeros(nc.value, dtype=np.int32)
        vals = np.zeros(nv.value)
        rhs = np.zeros(self.st.num_cols_f)
        self.L.oracle_explicit_eliminate(self.h, _dp(values), _dp(b), _dp(D), None, None, None,
                                         _ip(row_ptr), _ip(cols), _ip(pos), _dp(vals), _dp(rhs))
        return row_ptr, cols, pos, vals, rhs

    def explicit_multiply(self, x):
        y = np.zeros(self.st.num_cols_f)
        self.L.oracle_explicit_multiply(self.h, _dp(x), _dp(y))
        return y

    def explicit_solve(self, values, b, D, q_tolerance=0.0, r_tolerance=0.0, max_trace=0):
        x = np.zeros(self.st.num_cols)
        s = OracleSummary()
        trace = np.zeros((max_trace, self.st.num_cols_f)) if max_trace else None
        self.L.oracle_explicit_solve(self.h, _dp(values), _dp(b), _dp(D), q_tolerance,
                                     r_tolerance, _dp(x), ctypes.byref(s), _dp(trace), max_trace)
        summary = {"num_iterations": s.num_iterations,
                   "termination_type": s.termination_type,
                   "message": s.message.decode()}
        if max_trace:
            summary["trace"] = trace[:min(max_trace, max(s.num_iterations, 0))]
        return x, summary

    def reduced_solution(self):
        out = np.zeros(self.st.num_cols_f)
        self.L.oracle_reduced_solution(self.h, _dp(out))
        return out


def cg_dense(A, b, x0, min_it, max_it, reset_period, q_tol, r_tol):
    L = lib()
    A = np.ascontiguousarray(A, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.ascontiguousarray(x0, dtype=np.float64).copy()
    s = OracleSummary()
    L.oracle_cg_dense(len(b), _dp(A), _dp(b), _dp(x), min_it, max_it, reset_period,
                      q_tol, r_tol, ctypes.byref(s))
    return x, {"num_iterations": s.num_iterations,
               "termination_type": s.termination_type,
               "message": s.message.decode()}
