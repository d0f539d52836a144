"""Drop-in for the reference module code/gl_FProxGD_primal.py (same module name, function name and
positional signature, code/gl_FProxGD_primal.py:9; registered at code/main.py:178, called at :114).

    x, iter, out = gl_FProxGD_primal(x0, A, b, mu_0, opts)

Nesterov-accelerated proximal gradient (theta = 2/(k+1), x/v/y form) with mu-continuation and a
monotone backtracking step carried across iterations, executed on a B200 by libglprox.so with
exactly one transposed and one forward stream of A per accepted iteration.
`opts` keys and defaults are the reference's (gl_FProxGD_primal.py:10-19); unknown keys are ignored.
"""
from glprox.solver import solve as _solve


def gl_FProxGD_primal(x0, A, b, mu_0, opts):
    return _solve("FProxGD", x0, A, b, mu_0, opts)
