"""Synthetic bundle-adjustment problems of the BASELINE.json shapes (ctypes binding of
csrc/synthetic_bal.cc) and the pieces of Ceres' trust-region loop that sit directly
around the linear solve. Measurement / test support: none of this is on the device
path.
"""
import ctypes
import os

import numpy as np

from . import _build
from .solver import BlockStructure

# exact BAL dataset counts: cameras, points, observations
SHAPES = {
    "problem-16": (16, 22106, 83718),
    "ladybug": (1723, 156502, 678718),
    "venice": (1778, 993923, 5001946),
    "final-13682": (13682, 4456117, 28987644),
}

_dp = ctypes.POINTER(ctypes.c_double)
_ip = ctypes.POINTER(ctypes.c_int)
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(_build.build_bal())
        L.bal_synth_structure.restype = ctypes.c_int
        L.bal_synth_structure.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_longlong,
                                          ctypes.c_ulonglong, ctypes.c_int, _ip, _ip]
        L.bal_synth_geometry.restype = None
        L.bal_synth_geometry.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_longlong, _ip, _ip,
                                         ctypes.c_ulonglong] + [ctypes.c_double] * 4 + [_dp, _dp, _dp]
        L.bal_evaluate.restype = ctypes.c_double
        L.bal_evaluate.argtypes = [ctypes.c_int, ctypes.c_longlong, _ip, _ip, _dp, _dp, _dp, _dp, _dp, _dp]
        L.bal_file_error.restype = ctypes.c_char_p
        L.bal_file_open.restype = ctypes.c_int
        L.bal_file_open.argtypes = [ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p), _ip, _ip, _ip]
        L.bal_file_fetch.restype = None
        L.bal_file_fetch.argtypes = [ctypes.c_void_p, _ip, _ip, _dp, _dp, _dp]
        L.bal_file_write.restype = ctypes.c_int
        L.bal_file_write.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, _ip, _ip, _dp, _dp, _dp]
        L.bal_file_write_ply.restype = ctypes.c_int
        L.bal_file_write_ply.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, _dp, _dp]
        L.bal_normalize.restype = None
        L.bal_normalize.argtypes = [ctypes.c_int, ctypes.c_int, _dp, _dp]
        L.bal_perturb.restype = None
        L.bal_perturb.argtypes = [ctypes.c_int, ctypes.c_int] + [ctypes.c_double] * 3 + [_dp, _dp]
        L.bal_group_by_point.restype = None
        L.bal_group_by_point.argtypes = [ctypes.c_int, ctypes.c_int, _ip, _ip, _ip]
        _LIB = L
    return _LIB


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _i(a):
    return a.ctypes.data_as(_ip)


class SyntheticBAL:
    """Structure + geometry of a synthetic BAL problem with exactly (C, P, N)."""

    def __init__(self, num_cameras, num_points, num_observations, seed=38401, dup_pairs=0,
                 geometry=True, pixel_sigma=0.5, rotation_sigma=1e-3, translation_sigma=1e-1,
                 point_sigma=1e-1):
        C, P, N = int(num_cameras), int(num_points), int(num_observations)
        self.C, self.P, self.N = C, P, N
        self.point_row_start = np.zeros(P + 1, dtype=np.int32)
        self.cam_idx = np.zeros(N, dtype=np.int32)
        rc = lib().bal_synth_structure(C, P, N, seed, dup_pairs, _i(self.point_row_start), _i(self.cam_idx))
        if rc != 0:
            raise ValueError(f"infeasible shape C={C} P={P} N={N}")
        self.structure = BlockStructure.bal(self.point_row_start, self.cam_idx, C)
        self.cameras = self.points = self.observations = None
        if geometry:
            self.cameras = np.zeros(9 * C)
            self.points = np.zeros(3 * P)
            self.observations = np.zeros(2 * N)
            lib().bal_synth_geometry(C, P, N, _i(self.point_row_start), _i(self.cam_idx), seed, pixel_sigma,
                                     rotation_sigma, translation_sigma, point_sigma, _d(self.cameras),
                                     _d(self.points), _d(self.observations))

    @classmethod
    def from_arrays(cls, num_cameras, point_row_start, cam_idx, observations, cameras, points):
        """A problem given outright (rows grouped by point), e.g. from a BAL file."""
        self = cls.__new__(cls)
        self.point_row_start = np.ascontiguousarray(point_row_start, dtype=np.int32)
        self.cam_idx = np.ascontiguousarray(cam_idx, dtype=np.int32)
        self.C, self.P, self.N = int(num_cameras), len(self.point_row_start) - 1, len(self.cam_idx)
        self.structure = BlockStructure.bal(self.point_row_start, self.cam_idx, self.C)
        self.observations = np.ascontiguousarray(observations, dtype=np.float64)
        self.cameras = np.ascontiguousarray(cameras, dtype=np.float64)
        self.points = np.ascontiguousarray(points, dtype=np.float64)
        return self

    @classmethod
    def named(cls, name, scale=1.0, **kw):
        C, P, N = SHAPES[name]
        if scale != 1.0:
            P = max(8, int(P * scale))
            N = max(2 * P, int(N * scale))
        return cls(C, P, N, **kw)

    # --- Evaluator: residuals / cost / Jacobian in the BlockJacobianWriter layout ---------
    def evaluate(self, cameras=None, points=None, jacobian=True, values=None, residuals=None):
        """Returns (cost, residuals[2N], values[24N] or None): values = [E (N x 6) | F (N x 18)]."""
        cameras = self.cameras if cameras is None else cameras
        points = self.points if points is None else points
        N = self.N
        if residuals is None:
            residuals = np.empty(2 * N)
        E = F = None
        if jacobian:
            if values is None:
                values = np.empty(24 * N)
            E, F = values[:6 * N], values[6 * N:]
        cost = lib().bal_evaluate(self.P, N, _i(self.point_row_start), _i(self.cam_idx), _d(cameras),
                                  _d(points), _d(self.observations), _d(residuals), _d(E), _d(F))
        return cost, residuals, (values if jacobian else None)

    def random_values(self, seed=0):
        """Mode B of SURVEY 8(d): N(0,1) values / b / D on the same structure
        (what evaluation_benchmark.cc:89-99,147-153 does)."""
        rng = np.random.default_rng(seed)
        values = rng.standard_normal(24 * self.N)
        b = rng.standard_normal(2 * self.N)
        D = rng.standard_normal(3 * self.P + 9 * self.C)
        return values, b, D

    # --- column norms and Jacobi scaling (block_sparse_matrix.cc:351-450) --------------------
    def squared_column_norm(self, values):
        N, P, C = self.N, self.P, self.C
        E = values[:6 * N].reshape(N, 2, 3)
        F = values[6 * N:].reshape(N, 2, 9)
        row_point = np.repeat(np.arange(P), np.diff(self.point_row_start))
        out = np.zeros(3 * P + 9 * C)
        e2 = (E * E).sum(axis=1)
        f2 = (F * F).sum(axis=1)
        for k in range(3):
            out[k:3 * P:3] = np.bincount(row_point, weights=e2[:, k], minlength=P)
        for k in range(9):
            out[3 * P + k::9] = np.bincount(self.cam_idx, weights=f2[:, k], minlength=C)
        return out

    def scale_columns(self, values, scale):
        N, P = self.N, self.P
        row_point = np.repeat(np.arange(P), np.diff(self.point_row_start))
        E = values[:6 * N].reshape(N, 2, 3)
        F = values[6 * N:].reshape(N, 2, 9)
        E *= scale[:3 * P].reshape(P, 3)[row_point][:, None, :]
        F *= scale[3 * P:].reshape(self.C, 9)[self.cam_idx][:, None, :]
        return values

    def jacobian_times(self, values, x):
        """J x (trust_region_minimizer.cc:430-438 model cost change)."""
        N, P = self.N, self.P
        row_point = np.repeat(np.arange(P), np.diff(self.point_row_start))
        E = values[:6 * N].reshape(N, 2, 3)
        F = values[6 * N:].reshape(N, 2, 9)
        xe = x[:3 * P].reshape(P, 3)[row_point]
        xf = x[3 * P:].reshape(self.C, 9)[self.cam_idx]
        return (np.einsum("nij,nj->ni", E, xe) + np.einsum("nij,nj->ni", F, xf)).reshape(-1)


class BALFile:
    """A Bundle Adjustment in the Large problem file with the semantics of the reference's
    example loader (examples/bal_problem.cc:73-333, angle-axis cameras): read / write / PLY
    export / Normalize / Perturb. The arithmetic is csrc/bal_io.h."""

    def __init__(self, camera_index, point_index, observations, cameras, points):
        self.camera_index = np.ascontiguousarray(camera_index, dtype=np.int32)
        self.point_index = np.ascontiguousarray(point_index, dtype=np.int32)
        self.observations = np.ascontiguousarray(observations, dtype=np.float64)
        self.cameras = np.ascontiguousarray(cameras, dtype=np.float64)
        self.points = np.ascontiguousarray(points, dtype=np.float64)
        self.num_cameras = len(self.cameras) // 9
        self.num_points = len(self.points) // 3
        self.num_observations = len(self.camera_index)

    @classmethod
    def read(cls, path):
        h = ctypes.c_void_p()
        c, p, n = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        if lib().bal_file_open(os.fsencode(path), ctypes.byref(h), ctypes.byref(c), ctypes.byref(p), ctypes.byref(n)):
            raise ValueError(lib().bal_file_error().decode())
        ci, pi = np.empty(n.value, np.int32), np.empty(n.value, np.int32)
        obs, cams, pts = np.empty(2 * n.value), np.empty(9 * c.value), np.empty(3 * p.value)
        lib().bal_file_fetch(h, _i(ci), _i(pi), _d(obs), _d(cams), _d(pts))
        return cls(ci, pi, obs, cams, pts)

    def write(self, path):
        if lib().bal_file_write(os.fsencode(path), self.num_cameras, self.num_points, self.num_observations,
                                _i(self.camera_index), _i(self.point_index), _d(self.observations),
                                _d(self.cameras), _d(self.points)):
            raise OSError(lib().bal_file_error().decode())

    def write_ply(self, path):
        if lib().bal_file_write_ply(os.fsencode(path), self.num_cameras, self.num_points, _d(self.cameras),
                                    _d(self.points)):
            raise OSError(lib().bal_file_error().decode())

    def normalize(self):
        lib().bal_normalize(self.num_cameras, self.num_points, _d(self.cameras), _d(self.points))

    def perturb(self, rotation_sigma, translation_sigma, point_sigma):
        lib().bal_perturb(self.num_cameras, self.num_points, rotation_sigma, translation_sigma, point_sigma,
                          _d(self.cameras), _d(self.points))

    def problem(self):
        """Rows grouped by point, as the Schur ordering arranges them (reorder_program.cc:536-539):
        a SyntheticBAL-shaped problem (structure, observations in row order, parameters)."""
        order = np.empty(self.num_observations, np.int32)
        row_start = np.empty(self.num_points + 1, np.int32)
        lib().bal_group_by_point(self.num_points, self.num_observations, _i(self.point_index), _i(order), _i(row_start))
        obs = self.observations.reshape(-1, 2)[order].reshape(-1)
        return SyntheticBAL.from_arrays(self.num_cameras, row_start, self.camera_index[order], obs,
                                        self.cameras.copy(), self.points.copy())

    @classmethod
    def from_problem(cls, p):
        """The file form of a (synthetic) problem: observations listed point by point."""
        point_index = np.repeat(np.arange(p.P, dtype=np.int32), np.diff(p.point_row_start))
        return cls(p.cam_idx, point_index, p.observations, p.cameras, p.points)
