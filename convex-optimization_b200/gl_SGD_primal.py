"""Drop-in for the reference module code/gl_SGD_primal.py (same module name, function name and positional
signature, code/gl_SGD_primal.py:9; registered in code/main.py:174-176, called at :114).  SURVEY.md 8(f) row 1:
the same two streams of A per iteration as the proximal solvers, only the row-group epilogue differs.

    x, iter, out = gl_SGD_primal(x0, A, b, mu_0, opts)

Subgradient method (x_i / ((||x_i|| < thres) + ||x_i||)) with fixed -> diminishing steps and mu-continuation;
f_hist is recorded with the STAGE mu, as in the reference.
`opts` keys and defaults are the reference's; unknown keys are ignored.
"""
from glprox.solver import solve as _solve


def gl_SGD_primal(x0, A, b, mu_0, opts):
    return _solve("SGD", x0, A, b, mu_0, opts)
