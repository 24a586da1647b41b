"""b200ba -- B200-native bundle adjustment behind the call surface of hammershock/SFM-Python.

The package directory is ``sfm-python_b200`` (not an importable name); ``import b200ba`` at the repo
root loads it (see ``b200ba.py``).  Public names mirror ``sfm_lite/bundle_adjustment.py`` and the
``scipy.optimize.least_squares`` call of ``sfm_lite/sfm.py:266-268``.
"""
from .bundle_adjustment import (BAError, BAResult, Engine, JacobianSparsity, bundle_adjust, calc_reproj_error,
                                compute_residuals, create_sparsity_matrix, engine_for, gather_observations, install,
                                least_squares, pack_Rt, pack_parameters, reset, solve_pnp, total_reproj_error, triangulate_points,
                                triangulate_points_nonlinear, triangulate_visible, unpack_parameters)
from . import synthetic
from . import dist

__all__ = ["BAError", "BAResult", "Engine", "JacobianSparsity", "bundle_adjust", "calc_reproj_error", "compute_residuals",
           "create_sparsity_matrix", "engine_for", "gather_observations", "install", "least_squares", "pack_Rt",
           "pack_parameters", "reset", "solve_pnp", "total_reproj_error", "triangulate_points", "triangulate_points_nonlinear", "triangulate_visible", "unpack_parameters",
           "synthetic", "dist"]
__version__ = "0.1.0"
