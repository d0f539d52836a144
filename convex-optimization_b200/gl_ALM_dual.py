"""Drop-in for the reference module code/gl_ALM_dual.py (same module name, function name `gl_Alm_dual` and positional
signature, code/gl_ALM_dual.py:66; registered in code/main.py:179).  Augmented Lagrangian method on the dual: every
outer iteration runs the reference's 500 accelerated projected-gradient steps on u (solve_sub_problem, :10-63).  The
reference forms the n x n Hessian of that inner problem explicitly; here it is applied as streams of A and of the
m x m inverse Cholesky factor (csrc/glp_alm.cuh).

    x, iter, out = gl_Alm_dual(x0, A, b, mu, opts)

`opts` keys and defaults are the reference's (maxit 100, thres 1e-3, tau (1+sqrt 5)/2, rho 1e2, converge_len 20);
unknown keys are ignored.  Computes in fp64 (fp32 inputs are widened for the call); l <= 8, one GPU.
"""
from glprox.solver import solve_alm_dual as _solve


def gl_Alm_dual(x0, A, b, mu, opts):
    return _solve(x0, A, b, mu, opts)
