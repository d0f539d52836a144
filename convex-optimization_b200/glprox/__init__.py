"""glprox: B200-native proximal-gradient group-LASSO solvers (host side).

Everything numerical lives in libglprox.so (csrc/, hand-written sm_100a CUDA, C ABI in
include/glprox.h).  Importing this package does not load the library; the first solve does, and
fails loudly if it is missing.
"""
from . import _lib, dataset  # noqa: F401
from .solver import (DEFAULT_OPTS, GroupLassoProblem, attach_distributed, combine_metrics,  # noqa: F401
                     exchange_handles, merge_opts, shard_bounds, solve, solve_admm_dual, solve_admm_primal, solve_alm_dual,
                     solve_group)

__all__ = ["solve", "GroupLassoProblem", "DEFAULT_OPTS", "merge_opts", "shard_bounds", "exchange_handles",
           "attach_distributed", "solve_group", "combine_metrics", "dataset", "solve_admm_dual", "solve_admm_primal", "solve_alm_dual"]
