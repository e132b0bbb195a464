"""asset_asrl_b200 -- B200-native KKT evaluation for ASSET-style collocation problems.

Usage mirrors the reference package::

    import asset_asrl_b200 as ast
    vf, oc = ast.VectorFunctions, ast.OptimalControl
"""
from . import _lib  # noqa: F401  (fails loudly if the CUDA library is not built)
from . import VectorFunctions, Solvers, OptimalControl  # noqa: F401

__all__ = ["VectorFunctions", "Solvers", "OptimalControl"]
