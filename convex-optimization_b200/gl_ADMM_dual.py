"""Drop-in for the reference module code/gl_ADMM_dual.py (same module name, function name `gl_Admm_dual` and
positional signature, code/gl_ADMM_dual.py:10; registered in code/main.py:180, called at :114).  SURVEY.md 8(f)
row 4: ADMM on the dual problem; after the one-time Cholesky factor of I + rho A A^T every iteration costs the
same two streams of A as the proximal solvers.

    x, iter, out = gl_Admm_dual(x0, A, b, mu, opts)

`opts` keys and defaults are the reference's (maxit 100, thres 1e-3, tau (1+sqrt 5)/2, rho 1e2, converge_len 20);
unknown keys are ignored.  Scope of this build: l <= 8, one GPU (m > 2048 takes the blocked factorisation).
"""
from glprox.solver import solve_admm_dual as _solve


def gl_Admm_dual(x0, A, b, mu, opts):
    return _solve(x0, A, b, mu, opts)
