"""Shared helpers of the test-suite: load the product package (hyphenated directory
name), build matching oracle / product structures for synthetic BAL problems."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG_NAME = "ceres-solver-ceres-solver_b200"


def pkg():
    return importlib.import_module(PKG_NAME)


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def oracle_structure(st):
    """oracle.binding.Structure from a product BlockStructure (same flat arrays)."""
    from oracle import binding as ob
    return ob.Structure(st.row_size, st.row_pos, st.cell_ptr, st.cell_block_id, st.cell_position,
                        st.col_size, st.col_pos, st.num_eliminate_blocks)


def rel_err(a, b):
    a, b = np.asarray(a), np.asarray(b)
    d = np.linalg.norm(b)
    return np.linalg.norm(a - b) / (d if d > 0 else 1.0)


def lm_like_problem(p, radius=1e4, seed=0):
    """Jacobi-scaled Jacobian values, b = -residuals... as one LM iteration would pass them:
    values scaled by 1/(1+sqrt(colnorm2)) (trust_region_minimizer.cc:263-279),
    D = sqrt(clamp(colnorm2(J_scaled)) / radius) (levenberg_marquardt_strategy.cc:79-95)."""
    cost, residuals, values = p.evaluate()
    scale = 1.0 / (1.0 + np.sqrt(p.squared_column_norm(values)))
    p.scale_columns(values, scale)
    diag = np.clip(p.squared_column_norm(values), 1e-6, 1e32)
    D = np.sqrt(diag / radius)
    return values, residuals, D, scale, cost
