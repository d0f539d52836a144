"""Drop-in for the reference module code/gl_GD_primal.py (same module name, function name and positional
signature, code/gl_GD_primal.py:9; registered in code/main.py:174-176, called at :114).  SURVEY.md 8(f) row 1:
the same two streams of A per iteration as the proximal solvers, only the row-group epilogue differs.

    x, iter, out = gl_GD_primal(x0, A, b, mu_0, opts)

Gradient descent on the smoothed objective (delta = 1e-3) with fixed -> diminishing steps and
mu-continuation; no line search and no early stage break (the reference's break is commented out).
`opts` keys and defaults are the reference's; unknown keys are ignored.
"""
from glprox.solver import solve as _solve


def gl_GD_primal(x0, A, b, mu_0, opts):
    return _solve("GD", x0, A, b, mu_0, opts)
