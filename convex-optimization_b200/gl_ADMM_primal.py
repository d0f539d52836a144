"""Drop-in for the reference module code/gl_ADMM_primal.py (same module name, function name `gl_Admm_primal` and
positional signature, code/gl_ADMM_primal.py:10; registered in code/main.py:181).  Linearised ADMM on the primal
problem.  The reference factors the n x n matrix rho I + A^T A; here the solve goes through the m x m matrix
I + A A^T / rho, so that every iteration costs the same two streams of A as the proximal solvers
(csrc/glp_admm_primal.cuh).

    x, iter, out = gl_Admm_primal(x0, A, b, mu, opts)

`opts` keys and defaults are the reference's (maxit 100, thres 1e-3, tau (1+sqrt 5)/2, rho 1e-2, eta_0 100,
converge_len 10, converge_thres 1e-5 [read, never used -- as in the reference], step_type 'fixed'); unknown keys are
ignored.  Computes in fp64 (fp32 inputs are widened for the call); l <= 8, one GPU.
"""
from glprox.solver import solve_admm_primal as _solve


def gl_Admm_primal(x0, A, b, mu, opts):
    return _solve(x0, A, b, mu, opts)
