"""B200-native ITERATIVE_SCHUR linear solve for Ceres' Levenberg-Marquardt loop.

Only what the hot path needs: csrc/ (hand-written sm_100a FP64 CUDA kernels + the
C ABI of include/ceres_b200_schur.h), solver.py (host-side mirror of the
reference's IterativeSchurComplementSolver on top of that ABI) and bal.py
(synthetic BAL problems for measurement). The directory name has hyphens, so
import it with importlib.import_module("ceres-solver-ceres-solver_b200").
"""
from . import _build  # noqa: F401
from .solver import (IDENTITY, JACOBI, SCHUR_JACOBI, SCHUR_POWER_SERIES_EXPANSION,  # noqa: F401
                     TERMINATION, B200Error, BlockStructure, IterativeSchurComplementSolver,
                     nccl_unique_id, shard_plan, explicit_plan, host_register, host_unregister,
                     block_diagonal_apply_static)
from .bal import SHAPES, BALFile, SyntheticBAL  # noqa: F401
