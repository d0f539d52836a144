"""Drop-in for the reference module code/gl_FGD_primal.py (same module name, function name and positional
signature, code/gl_FGD_primal.py:9; registered in code/main.py:174-176, called at :114).  SURVEY.md 8(f) row 1:
the same two streams of A per iteration as the proximal solvers, only the row-group epilogue differs.

    x, iter, out = gl_FGD_primal(x0, A, b, mu_0, opts)

Nesterov-accelerated gradient on the smoothed objective  0.5||Ax-b||^2 + mu*sum_i(sqrt(||x_i||^2+delta^2)-delta)
(prox = identity), mu-continuation and the same carried backtracking step as FProxGD.
`opts` keys and defaults are the reference's; unknown keys are ignored.
"""
from glprox.solver import solve as _solve


def gl_FGD_primal(x0, A, b, mu_0, opts):
    return _solve("FGD", x0, A, b, mu_0, opts)
