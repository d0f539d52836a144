"""Drop-in for the reference module code/gl_ProxGD_primal.py (same module name, function name and
positional signature, code/gl_ProxGD_primal.py:9; registered at code/main.py:177, called at :114).

    x, iter, out = gl_ProxGD_primal(x0, A, b, mu_0, opts)

Proximal gradient with mu-continuation [100*mu, 10*mu, mu] and a restarted backtracking line search,
executed on a B200 by libglprox.so: per iteration A is streamed once transposed (gradient, fused
prox epilogue) and once forward per line-search trial (fused residual epilogue).
`opts` keys and defaults are the reference's (gl_ProxGD_primal.py:10-19); unknown keys are ignored.
"""
from glprox.solver import solve as _solve


def gl_ProxGD_primal(x0, A, b, mu_0, opts):
    return _solve("ProxGD", x0, A, b, mu_0, opts)
