"""CPU oracle for the proximal-gradient group-LASSO hot path (TEST INFRASTRUCTURE ONLY).

This file is a NumPy *restatement* of the algorithm implemented by the reference solvers
  /root/reference/code/gl_ProxGD_primal.py   (ProxGD,  146 lines)
  /root/reference/code/gl_FProxGD_primal.py  (FProxGD, 161 lines)
for   min_x 0.5*||A x - b||_F^2 + mu * sum_i ||x(i,:)||_2 .

It is the checker, not the product: only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import it.  The product path (the CUDA library behind
convex-optimization_b200/) never imports anything from oracle/.

Parity pin: PINNED.  tests/golden/make_golden.py imports the *unmodified* reference functions
from /root/reference/code in the build container and stores their outputs (k, fval, f_hist, x,
support) on seeded inputs in tests/golden/*.npz; tests/test_oracle.py asserts that this
restatement reproduces every one of them (bit-for-bit in k / support, <=1e-13 relative in
f_hist), plus the published report numbers (doc/report.md:446-447).

Two restatements live here:
  * `solve(...)`            -- same pass structure as the reference (5-7 products with A per
                               iteration); this is what the CPU baseline times.
  * `solve_two_stream(...)` -- the algebraically equivalent schedule the GPU executes (A streamed
                               once forward and once transposed per iteration; A.x carried by
                               linearity).  Used as the step-level oracle
                               for the kernels and to prove the schedule equals the reference.
"""
import math
import time

import numpy as np

# --- option tables (reference: gl_ProxGD_primal.py:10-19, gl_FProxGD_primal.py:10-19) -------------
_COMMON = dict(thres=1e-3, step_type="line_search", ftol=1e-6, stable_len_threshold=70,
               maxit_line_search_iter=5)
DEFAULTS = {
    "ProxGD": dict(_COMMON, maxit=2500, alpha0=2e-3, line_search_attenuation_coeffi=0.9),
    "FProxGD": dict(_COMMON, maxit=1500, alpha0=1e-3, line_search_attenuation_coeffi=0.98),
    # "next" row f-1 of SURVEY.md 8(f): the smoothed / subgradient solvers on the same two streams
    # (gl_FGD_primal.py:10-21, gl_GD_primal.py:10-19, gl_SGD_primal.py:10-18)
    "FGD": dict(_COMMON, maxit=1500, alpha0=1e-3, line_search_attenuation_coeffi=0.98, delta=1e-6),
    "GD": dict(maxit=2500, thres=1e-3, step_type="diminishing", alpha0=1e-3, ftol=1e-5, stable_len_threshold=100,
               continuous_subgradient_flag=False, delta=1e-3, line_search_attenuation_coeffi=1.0,
               maxit_line_search_iter=0),
    "SGD": dict(maxit=2100, thres=1e-3, step_type="diminishing", alpha0=1e-3, ftol=1e-5, stable_len_threshold=100,
                continuous_subgradient_flag=False, delta=0.0, line_search_attenuation_coeffi=1.0,
                maxit_line_search_iter=0),
}
CONTINUATION = (100.0, 10.0, 1.0)  # mu stages: gl_ProxGD_primal.py:53 / gl_FProxGD_primal.py:55


def gen_data(seed=97006855, m=256, n=512, l=2, dtype=np.float64):
    """Fixture generator, restated from code/main.py:37-51 (draw order A, perm, u, x0 matters)."""
    rng = np.random.Generator(np.random.MT19937(seed=seed))
    A = rng.standard_normal(size=(m, n))
    nnz_rows = round(n * 0.1)
    picked = rng.permutation(n)[:nnz_rows]
    u = np.zeros((n, l))
    u[picked, :] = rng.standard_normal(size=(nnz_rows, l))
    b = A @ u
    x0 = rng.standard_normal(size=(n, l))
    mu = 1e-2
    if dtype != np.float64:
        A, b, x0, u = (z.astype(dtype) for z in (A, b, x0, u))
    return A, b, x0, u, mu


def row_norms(z):
    """||z(i,:)||_2 per row, shape (n,1).  numpy LA.norm(z, axis=1) == sqrt(add.reduce(z*z, 1))."""
    return np.sqrt(np.add.reduce(z * z, axis=1)).reshape(-1, 1)


def group_prox(z, t, mu, thres):
    """Prox of t*mu*sum_i||z_i|| with the reference's guarded denominator
    (gl_ProxGD_primal.py:65-71 / gl_FProxGD_primal.py:71-77): multiply first, then divide; the
    denominator gets +1.0 on rows whose norm is below `thres`."""
    rn = row_norms(z)
    shrink = np.clip(rn - t * mu, a_min=0, a_max=None)
    return z * shrink / ((rn < thres) + rn)


def objective(A, b, x, mu0):
    """real_obj_func, gl_*_primal.py:24-27: always the *final* mu (mu_0)."""
    res = A @ x - b
    return 0.5 * np.sum(res ** 2) + mu0 * np.sum(row_norms(x))


def sparsity(x):
    """gl_*_primal.py:22."""
    ax = np.abs(x)
    return np.sum(ax > 1e-6 * np.max(ax)) / x.size


def smooth_part(A, b, z):
    """g(z) = 0.5*||A z - b||^2  (gl_ProxGD_primal.py:60-61 / gl_FProxGD_primal.py:62-63)."""
    return 0.5 * np.sum((A @ z - b) ** 2)


def smooth_reg(z, delta):
    """sum_i sqrt(||z_i||^2 + delta^2) - delta   (gl_FGD_primal.py:64-67)."""
    return np.sum(np.sqrt(np.sum(z ** 2, axis=1).reshape(-1, 1) + delta * delta) - delta)


def smooth_reg_grad(z, delta):
    """z / sqrt(||z_i||^2 + delta^2)   (gl_FGD_primal.py:69-72, gl_GD_primal.py:59-63)."""
    return z / np.sqrt(np.sum(z ** 2, axis=1).reshape(-1, 1) + delta * delta)


def subgrad_reg(z, thres):
    """z / ((||z_i|| < thres) + ||z_i||)   (gl_SGD_primal.py:56-61)."""
    rn = row_norms(z)
    return z / ((rn < thres) + rn)


def solve_smooth(algo, x0, A, b, mu0, opts=None):
    """Reference-structured restatement of gl_FGD_primal / gl_GD_primal / gl_SGD_primal
    (code/gl_FGD_primal.py:56-163, code/gl_GD_primal.py:53-112, code/gl_SGD_primal.py:47-109)."""
    o = dict(DEFAULTS[algo])
    o.update(opts or {})
    maxit, ftol, alpha0, thres, rule = o["maxit"], o["ftol"], o["alpha0"], o["thres"], o["step_type"]
    if o.get("continuous_subgradient_flag") and algo in ("GD", "SGD"):   # gl_GD_primal.py:43-45, gl_SGD_primal.py:35-37
        alpha0 = 1.0 / np.max(np.linalg.eigvals(A.T @ A)).real
    delta = o["delta"]
    f_hist, f_best_hist, s_hist = [], [], []
    f_best = np.inf
    xk = np.copy(x0)
    x_last = None if algo == "FGD" else xk
    k, stable = 0, 0
    t_begin = time.time_ns()
    mu = mu0
    for factor in CONTINUATION:
        mu = factor * mu0
        inner = 0
        if algo == "FGD":
            vk, tk, stable = np.copy(xk), alpha0, 0
        while inner < maxit:
            f_now = objective(A, b, xk, mu if algo == "SGD" else mu0)   # SGD records the stage objective (:50-53)
            f_hist.append(f_now)
            f_best = min(f_best, f_now)
            f_best_hist.append(f_best)
            s_hist.append(sparsity(xk))
            k += 1
            inner += 1
            if algo == "FGD":
                with np.errstate(divide="ignore", invalid="ignore"):
                    ok = (k > 1 and abs(f_hist[k - 1] - f_hist[k - 2]) / abs(f_hist[k - 2]) < ftol
                          and abs(s_hist[k - 1] - s_hist[k - 2]) / abs(s_hist[k - 2]) < ftol)
                stable = stable + 1 if ok else 0
                if stable > o["stable_len_threshold"]:
                    break
            xk[np.abs(xk) < thres] = 0
            if algo == "FGD":
                theta = 2 / (inner + 1)
                y = (1 - theta) * xk + theta * vk
                grad = A.T @ (A @ y - b) + mu * smooth_reg_grad(y, delta)
                if rule == "line_search":
                    t = tk
                    g_y = 0.5 * np.sum((A @ y - b) ** 2) + mu * smooth_reg(y, delta)
                    for _ in range(o["maxit_line_search_iter"]):
                        x = y - t * grad
                        g_x = 0.5 * np.sum((A @ x - b) ** 2) + mu * smooth_reg(x, delta)
                        if g_x <= g_y + np.sum(grad * (x - y)) + np.sum((x - y) ** 2) / (2 * t):
                            break
                        t *= o["line_search_attenuation_coeffi"]
                else:
                    t = _open_loop_step(rule, alpha0, inner)
                x = y - t * grad
                vk = xk + (x - xk) / theta
                xk, tk = x, t
                x_last = x
            else:
                reg_g = smooth_reg_grad(xk, delta) if algo == "GD" else subgrad_reg(xk, thres)
                sub_g = A.T @ (A @ xk - b) + mu * reg_g
                if rule == "fixed" or mu > mu0:          # gl_GD_primal.py:67-70 / gl_SGD_primal.py:65-68
                    alpha = alpha0
                else:
                    alpha = _open_loop_step(rule, alpha0, inner)
                xk = xk - alpha * sub_g
                x_last = xk
    tt = _round_us(time.time_ns() - t_begin) / 1e6
    if x_last is None:
        raise UnboundLocalError("FGD ran no iteration (maxit=0), as in the reference")
    fval = objective(A, b, x_last, mu if algo == "SGD" else mu0)
    return x_last, k, {"tt": tt, "fval": fval, "f_hist": f_hist, "f_hist_best": f_best_hist}


ADMM_DUAL_DEFAULTS = dict(maxit=100, thres=1e-3, tau=(1 + math.sqrt(5)) * 0.5, rho=1e2, converge_len=20)


def solve_admm_dual(x0, A, b, mu, opts=None):
    """Reference-structured restatement of gl_Admm_dual (code/gl_ADMM_dual.py:10-103): ADMM on the dual,
    no mu-continuation; the z-update solves (I + rho A A^T) z = A(x - rho u) - b through a Cholesky factor
    computed once (:57, :63), u is the projection onto the row-wise mu-ball (:45-47, :64), x the multiplier
    update (:65); the loop stops after converge_len consecutive iterations whose primal / dual residuals
    (matrix 2-norms = largest singular values, :87-88) are both below thres."""
    o = dict(ADMM_DUAL_DEFAULTS)
    o.update(opts or {})
    maxit, thres, rho, tau, converge_len = o["maxit"], o["thres"], o["rho"], o["tau"], o["converge_len"]
    f_hist, f_best_hist = [], []
    f_best = np.inf
    x_k = np.copy(x0)
    u_k = np.zeros_like(x_k)
    t_begin = time.time_ns()
    k = 0
    L = np.linalg.cholesky(np.identity(A.shape[0]) + rho * A @ A.T)
    length = 0
    while k < maxit:
        k += 1
        z = np.linalg.solve(L.T, np.linalg.solve(L, A @ (x_k - rho * u_k) - b))
        v = x_k / rho - A.T @ z
        rn = np.linalg.norm(v, axis=1, ord=2).reshape(-1, 1)
        u = mu * v / np.clip(rn, a_min=mu, a_max=None)
        x = x_k - tau * rho * (u + A.T @ z)
        r_k = u + A.T @ z
        s_k = A @ (u_k - u)
        u_k, x_k = u, x
        f_now = objective(A, b, x_k, mu)
        f_hist.append(f_now)
        f_best = min(f_best, f_now)
        f_best_hist.append(f_best)
        if np.linalg.norm(r_k, ord=2) < thres and np.linalg.norm(s_k, ord=2) < thres:
            length += 1
        else:
            length = 0
        if length >= converge_len:
            break
    tt = _round_us(time.time_ns() - t_begin) / 1e6
    return x_k, k, {"tt": tt, "fval": objective(A, b, x_k, mu), "f_hist": f_hist, "f_hist_best": f_best_hist}


ADMM_PRIMAL_DEFAULTS = dict(maxit=100, thres=1e-3, tau=(1 + math.sqrt(5)) * 0.5, rho=1e-2, eta_0=100, converge_len=10,
                            converge_thres=1e-5, step_type="fixed")


class _ResidualStop:
    """The stop rule the three ADMM / ALM solvers share (gl_ADMM_primal.py:98-107, gl_ADMM_dual.py:87-99,
    gl_ALM_dual.py:137-147): converge_len consecutive iterations whose residual 2-norms are both below thres."""

    def __init__(self, thres, converge_len):
        self.thres, self.cap, self.run = thres, converge_len, 0

    def __call__(self, r_k, s_k):
        small = np.linalg.norm(r_k, ord=2) < self.thres and np.linalg.norm(s_k, ord=2) < self.thres
        self.run = self.run + 1 if small else 0
        return self.run >= self.cap


def solve_admm_primal(x0, A, b, mu, opts=None):
    """Reference-structured restatement of gl_Admm_primal (code/gl_ADMM_primal.py:10-117): linearised ADMM on
    min 0.5||A y - b||^2 + mu sum ||x_i||  s.t. x = y.  Once: Cholesky factor of rho I + A^T A (:62), A^T b (:63).
    Per iteration (:75-107): y solves (rho I + A^T A) y = A^T b - z + rho x through two dense solves with the factor
    (:78); x is one prox-gradient step of length eta on the augmented term (:80, prox_tf :47-51 with the guarded
    denominator); z the multiplier step (:81); residuals r = x - y, s = y - y_prev (:83-84).  Note y_0 = z_0 = x0
    (:53-55).  converge_thres is read but never used (:41)."""
    o = dict(ADMM_PRIMAL_DEFAULTS)
    o.update(opts or {})
    maxit, thres, rho, tau, eta_0 = o["maxit"], o["thres"], o["rho"], o["tau"], o["eta_0"]
    step_type = o["step_type"]
    stop = _ResidualStop(thres, o["converge_len"])
    f_hist, f_best_hist = [], []
    f_best = np.inf
    x_k = np.copy(x0)
    y_k = x_k
    z_k = x_k
    t_begin = time.time_ns()
    L = np.linalg.cholesky(rho * np.identity(A.shape[1]) + A.T @ A)
    AT_b = A.T @ b
    k = 0
    while k < maxit:
        k += 1
        if step_type == "fixed":
            eta = eta_0
        elif step_type == "diminishing":
            eta = eta_0 / np.sqrt(k)
        elif step_type == "diminishing2":
            eta = eta_0 / k
        else:
            eta = None      # the reference's set_step falls through; the next line then raises TypeError
        y = np.linalg.solve(L.T, np.linalg.solve(L, AT_b - z_k + rho * x_k))
        w = x_k - eta * rho * (x_k - y - z_k / rho)
        rn = np.linalg.norm(w, axis=1).reshape(-1, 1)
        x = w * np.clip(rn - eta * mu, a_min=0, a_max=None) / ((rn < thres) + rn)
        z = z_k - tau * rho * (x - y)
        r_k = x - y
        s_k = y - y_k
        x_k, y_k, z_k = x, y, z
        f_now = objective(A, b, x_k, mu)
        f_hist.append(f_now)
        f_best = min(f_best, f_now)
        f_best_hist.append(f_best)
        if stop(r_k, s_k):
            break
    tt = _round_us(time.time_ns() - t_begin) / 1e6
    return x_k, k, {"tt": tt, "fval": objective(A, b, x_k, mu), "f_hist": f_hist, "f_hist_best": f_best_hist}


ALM_DUAL_DEFAULTS = dict(maxit=100, thres=1e-3, tau=(1 + math.sqrt(5)) * 0.5, rho=1e2, converge_len=20)
ALM_INNER_STEPS = 500     # gl_ALM_dual.py:50
ALM_INNER_T = 1e-2        # gl_ALM_dual.py:54


def _alm_subproblem(b, rho, A, x_k, L, mu):
    """gl_ALM_dual.py:10-63: min over u in the row-wise mu-ball of the augmented Lagrangian with z eliminated -- a
    quadratic with the dense n x n Hessian Q and linear term J the reference forms explicitly (:33-40) -- by 500
    accelerated projected-gradient steps from u = 0 (gamma = 2/(k+1), fixed step 1e-2)."""
    T = np.linalg.solve(L.T, np.linalg.solve(L, np.identity(L.shape[0])))
    D = A @ x_k - b
    E = T @ D
    F = rho * T @ A
    G = np.identity(F.shape[1]) - A.T @ F
    H = A.T @ E - x_k / rho
    Q = F.T @ F + rho * G.T @ G
    J = rho * G.T @ H - F.T @ E - F.T @ b
    u_k = np.zeros_like(x_k)
    v_k = u_k
    for k in range(1, ALM_INNER_STEPS + 1):
        gamma = 2 / (k + 1)
        y = (1 - gamma) * u_k + gamma * v_k
        w = y - ALM_INNER_T * (Q @ y + J)
        u = mu * w / np.clip(np.linalg.norm(w, axis=1, ord=2).reshape(-1, 1), a_min=mu, a_max=None)
        v = u_k + (u - u_k) / gamma
        u_k, v_k = u, v
    return u_k


def solve_alm_dual(x0, A, b, mu, opts=None):
    """Reference-structured restatement of gl_Alm_dual (code/gl_ALM_dual.py:66-158): augmented Lagrangian method on
    the dual.  Once: Cholesky factor of I + rho A A^T (:114).  Per outer iteration (:119-147): u from the inner
    solver above (:122), z from the factor (:123), multiplier step x (:124), residuals r = u + A^T z,
    s = A(u_prev - u) (:126-127), the shared stop rule."""
    o = dict(ALM_DUAL_DEFAULTS)
    o.update(opts or {})
    maxit, thres, rho, tau = o["maxit"], o["thres"], o["rho"], o["tau"]
    stop = _ResidualStop(thres, o["converge_len"])
    f_hist, f_best_hist = [], []
    f_best = np.inf
    x_k = np.copy(x0)
    u_k = np.zeros_like(x_k)
    t_begin = time.time_ns()
    L = np.linalg.cholesky(np.identity(A.shape[0]) + rho * A @ A.T)
    k = 0
    while k < maxit:
        k += 1
        u = _alm_subproblem(b, rho, A, x_k, L, mu)
        z = np.linalg.solve(L.T, np.linalg.solve(L, A @ (x_k - rho * u) - b))
        x = x_k - tau * rho * (u + A.T @ z)
        r_k = u + A.T @ z
        s_k = A @ (u_k - u)
        u_k, x_k = u, x
        f_now = objective(A, b, x_k, mu)
        f_hist.append(f_now)
        f_best = min(f_best, f_now)
        f_best_hist.append(f_best)
        if stop(r_k, s_k):
            break
    tt = _round_us(time.time_ns() - t_begin) / 1e6
    return x_k, k, {"tt": tt, "fval": objective(A, b, x_k, mu), "f_hist": f_hist, "f_hist_best": f_best_hist}


def _open_loop_step(step_type, alpha0, inner):
    """Non-adaptive step rules, gl_ProxGD_primal.py:79-85 / gl_FProxGD_primal.py:82-88."""
    it_hat = max(inner, 1000) - 999
    if step_type == "fixed":
        return alpha0
    if step_type == "diminishing":
        return alpha0 / np.sqrt(it_hat)
    if step_type == "diminishing2":
        return alpha0 / it_hat
    raise ValueError("unsupported step_type %r" % (step_type,))


class _History:
    """f_hist / f_hist_best / sparsity_hist bookkeeping + the stop rule (:107-125 / :112-134)."""

    def __init__(self, ftol, stable_cap):
        self.f, self.fbest, self.s = [], [], []
        self.best = np.inf
        self.ftol, self.cap = ftol, stable_cap
        self.k = 0

    def record(self, f_now, s_now):
        self.f.append(f_now)
        self.best = min(self.best, f_now)
        self.fbest.append(self.best)
        self.s.append(s_now)
        self.k += 1

    def is_stable(self):
        k = self.k
        if k <= 1:
            return False
        with np.errstate(divide="ignore", invalid="ignore"):
            df = abs(self.f[k - 1] - self.f[k - 2]) / abs(self.f[k - 2])
            ds = abs(self.s[k - 1] - self.s[k - 2]) / abs(self.s[k - 2])
        return bool(df < self.ftol and ds < self.ftol)


def solve(algo, x0, A, b, mu0, opts=None, trace=None):
    """Reference-structured solver.  algo in {"ProxGD", "FProxGD"}.

    Returns (x, k, out) with out = {tt, fval, f_hist, f_hist_best} exactly like
    gl_ProxGD_primal.py:138-146 / gl_FProxGD_primal.py:153-161.  `trace`, if a dict, receives
    pass counters and the per-iteration accepted steps (test diagnostics only).
    """
    o = dict(DEFAULTS[algo])
    o.update(opts or {})
    maxit, ftol, alpha0, thres = o["maxit"], o["ftol"], o["alpha0"], o["thres"]
    shrink_c, ls_cap, rule = (o["line_search_attenuation_coeffi"], o["maxit_line_search_iter"],
                              o["step_type"])
    accel = algo == "FProxGD"
    if rule == "bb" and accel:
        raise ValueError("step_type 'bb' exists for ProxGD only")
    bb_t = alpha0   # opt-in BB rule (NOT in the reference): first trial step of the next line search
    hist = _History(ftol, o["stable_len_threshold"])
    steps = []
    xk = np.copy(x0)
    x_last = None if accel else xk  # FProxGD returns the last prox output (unbound if none ran)
    t_begin = time.time_ns()
    for factor in CONTINUATION:
        mu = factor * mu0
        inner, stable = 0, 0
        if accel:
            vk, tk = np.copy(xk), alpha0          # gl_FProxGD_primal.py:68-69
        while inner < maxit:
            hist.record(objective(A, b, xk, mu0), sparsity(xk))
            inner += 1
            stable = stable + 1 if hist.is_stable() else 0
            if stable > o["stable_len_threshold"]:
                break
            xk[np.abs(xk) < thres] = 0            # :127 / :136, element-wise, in place
            if accel:
                theta = 2 / (inner + 1)           # :138
                base = (1 - theta) * xk + theta * vk
            else:
                base = xk
            grad = A.T @ (A @ base - b)           # :129 / :140
            if rule in ("line_search", "bb"):
                g_base = smooth_part(A, b, base)
                t = tk if accel else (bb_t if rule == "bb" else alpha0)   # FProxGD carries t, ProxGD restarts (:91 / :94)
                for _ in range(ls_cap):
                    cand = group_prox(base - t * grad, t, mu, thres)
                    if accel:                     # gl_FProxGD_primal.py:96-99
                        d = cand - base
                        ok = smooth_part(A, b, cand) <= g_base + np.sum(grad * d) + np.sum(d ** 2) / (2 * t)
                    else:                         # gl_ProxGD_primal.py:73-74, 89-92
                        gmap = (base - cand) / t
                        ok = (smooth_part(A, b, base - t * gmap)
                              <= g_base - t * np.sum(grad * gmap) + 0.5 * t * np.sum(gmap ** 2))
                    if ok:
                        break
                    t *= shrink_c
            else:
                t = _open_loop_step(rule, alpha0, inner)
            x_new = group_prox(base - t * grad, t, mu, thres)   # :132 / :144
            steps.append(t)
            if rule == "bb":                      # BB1 = ||d||^2 / ||A d||^2, by its definition (extra product)
                d = x_new - base
                num, den = float(np.sum(d * d)), float(np.sum((A @ d) ** 2))
                bb_t = min(max(num / den, alpha0 * 1e-3), alpha0 * 1e3) if den > 0.0 and num > 0.0 else alpha0
            if accel:
                vk = xk + (x_new - xk) / theta    # :145 (xk is the thresholded iterate)
                tk = t
            xk = x_new
            x_last = x_new
    tt = _round_us(time.time_ns() - t_begin) / 1e6
    if x_last is None:
        raise UnboundLocalError("FProxGD ran no iteration (maxit=0), as in the reference")
    out = {"tt": tt, "fval": objective(A, b, x_last, mu0), "f_hist": hist.f,
           "f_hist_best": hist.fbest}
    if trace is not None:
        trace.update(steps=steps, sparsity_hist=hist.s)
    return x_last, hist.k, out


def _round_us(ns):
    """Stopwatch kMicroSecond rounding, code/utils/stopwatch.py:54-56."""
    return ns // 1000 + (1 if (ns // 100) % 10 >= 5 else 0)


def gl_ProxGD_primal(x0, A, b, mu_0, opts):
    return solve("ProxGD", x0, A, b, mu_0, opts)


def gl_FProxGD_primal(x0, A, b, mu_0, opts):
    return solve("FProxGD", x0, A, b, mu_0, opts)


# =================================================================================================
# Two-stream schedule (what the GPU executes).  One transposed and one forward stream of A per
# iteration; every other product is obtained by linearity from carried m x l vectors.
# =================================================================================================
def hard_threshold(x, thres):
    """Copy of x with |x| < thres set to zero (gl_ProxGD_primal.py:127 / gl_FProxGD_primal.py:136)."""
    out = x.copy()
    out[np.abs(out) < thres] = 0
    return out


def prox_epilogue(algo, grad, xt, v, theta, t, mu, thres, delta=0.0, full=False):
    """What the transposed pass's fused epilogue produces for one trial step t.

    Returns (xn, xnt, vn, gd, dd, reg, mx) [+ (sx, sy) with full=True]:  candidate iterate, its
    thresholded copy, the momentum candidate, the two line-search inner products, sum_i||xn_i||,
    max|xn| and the smoothed regulariser at the candidate / base point.
    FProxGD: base y=(1-theta)*xt+theta*v, xn=prox(y-t*grad), gd=<grad,xn-y>, dd=||xn-y||^2.
    ProxGD:  base xt, gd=<grad,gt>, dd=||gt||^2 with gt=(xt-xn)/t.
    FGD:     base y, gtot=grad+mu*y/sqrt(||y_i||^2+delta^2), xn=y-t*gtot, gd=<gtot,xn-y>, dd=||xn-y||^2.
    GD/SGD:  base xt, gtot=grad+mu*(smoothed | guarded sub)gradient, xn=xt-t*gtot.
    """
    accel = algo in ("FProxGD", "FGD")
    base = (1 - theta) * xt + theta * v if accel else xt
    sx = sy = 0.0
    if algo in ("FProxGD", "ProxGD"):
        xn = group_prox(base - t * grad, t, mu, thres)
        gtot = grad
    else:
        reg_g = subgrad_reg(base, thres) if algo == "SGD" else smooth_reg_grad(base, delta)
        gtot = grad + mu * reg_g
        xn = base - t * gtot
        if algo == "FGD":
            sx, sy = smooth_reg(xn, delta), smooth_reg(base, delta)
    xnt = hard_threshold(xn, thres)
    vn = xt + (xn - xt) / theta if accel else None
    if algo == "ProxGD":
        d = (base - xn) / t
    else:
        d = xn - base
    gd, dd = np.sum(gtot * d), np.sum(d ** 2)
    reg = np.sum(row_norms(xn))
    mx = np.max(np.abs(xn))
    if full:
        return xn, xnt, vn, gd, dd, reg, mx, sx, sy
    return xn, xnt, vn, gd, dd, reg, mx


def solve_two_stream(algo, x0, A, b, mu0, opts=None, trace=None, matmul=None, rmatmul=None):
    """Two-stream restatement; same return contract as `solve`.

    matmul(X) -> A @ X and rmatmul(R) -> A.T @ R can be overridden (e.g. with fp32-rounded or
    column-sharded products) to emulate the GPU arithmetic in tests.
    """
    o = dict(DEFAULTS[algo])
    o.update(opts or {})
    maxit, ftol, alpha0, thres = o["maxit"], o["ftol"], o["alpha0"], o["thres"]
    shrink_c, ls_cap, rule = (o["line_search_attenuation_coeffi"], o["maxit_line_search_iter"],
                              o["step_type"])
    if rule not in ("line_search", "fixed", "diminishing", "diminishing2"):
        raise ValueError("unsupported step_type %r" % (rule,))
    accel = algo == "FProxGD"
    fwd = matmul or (lambda X: A @ X)
    bwd = rmatmul or (lambda R: A.T @ R)
    n, l = x0.shape
    hist = _History(ftol, o["stable_len_threshold"])
    n_fwd = n_bwd = 0
    t_begin = time.time_ns()

    # ---- initial candidate: x0 itself, committed unconditionally
    xn, xnt = np.array(x0, dtype=np.float64), hard_threshold(np.array(x0, dtype=np.float64), thres)
    reg, mx = np.sum(row_norms(xn)), np.max(np.abs(xn)) if xn.size else 0.0
    prod = fwd(np.concatenate([xn, xnt], axis=1)); n_fwd += 1
    Axn, Axnt = prod[:, :l], prod[:, l:]
    rr = np.sum((Axn - b) ** 2)
    cnt = np.sum(np.abs(xn) > 1e-6 * mx)
    vn = None
    x = xt = v = Ax = Axt = Av = None
    stage, inner, stable, t_k, theta, t, trial = 0, 0, 0, alpha0, 1.0, alpha0, 0
    tested = False          # does the pending candidate carry a sufficient-decrease test?
    first_commit = True
    mu = CONTINUATION[0] * mu0
    ran_any = False
    while True:
        # -------- decision on the pending candidate (K3, part 1)
        if tested:
            g_x = 0.5 * rr
            if accel:
                ok = g_x <= g_y + gd + dd / (2 * t)
            else:
                ok = g_x <= g_y - t * gd + 0.5 * t * dd
            if not ok:
                t *= shrink_c
                trial += 1
                tested = trial < ls_cap      # the (ls_cap+1)-th step is used untested
                xn, xnt, vn, gd, dd, reg, mx = prox_epilogue(algo, grad, xt, v, theta, t, mu, thres)
                prod = fwd(np.concatenate([xn, xnt], axis=1)); n_fwd += 1
                Axn, Axnt = prod[:, :l], prod[:, l:]
                rr = np.sum((Axn - b) ** 2)
                cnt = np.sum(np.abs(xn) > 1e-6 * mx)
                continue
        # -------- commit (K3, part 2)
        if accel and not first_commit:
            v, Av = vn, Axt + (Axn - Axt) / theta
            t_k = t
        x, xt, Ax, Axt = xn, xnt, Axn, Axnt
        first_commit = False
        f_cur = 0.5 * rr + mu0 * reg
        s_cur = cnt / (n * l)
        # -------- loop tops / stage transitions (K3, part 3)
        done = False
        fresh_stage = stage == 0 and hist.k == 0 and inner == 0
        while True:
            if fresh_stage:
                inner, stable = 0, 0
                mu = CONTINUATION[stage] * mu0
                if accel:
                    v, Av, t_k = x, Ax, alpha0     # un-thresholded copy, gl_FProxGD_primal.py:68-69
                fresh_stage = False
            advance = False
            if inner >= maxit:
                advance = True
            else:
                hist.record(f_cur, s_cur)
                inner += 1
                stable = stable + 1 if hist.is_stable() else 0
                if stable > o["stable_len_threshold"]:
                    advance = True
            if not advance:
                break
            stage += 1
            if stage == len(CONTINUATION):
                done = True
                break
            fresh_stage = True
        if done:
            break
        # -------- set up the next iteration (K3, part 4): residual at the base point by linearity
        ran_any = True
        if accel:
            theta = 2 / (inner + 1)
            Abase = (1 - theta) * Axt + theta * Av
        else:
            theta = 1.0
            Abase = Axt
        r = Abase - b
        g_y = 0.5 * np.sum(r ** 2)
        if rule == "line_search":
            t, tested = (t_k if accel else alpha0), ls_cap > 0
        else:
            t, tested = _open_loop_step(rule, alpha0, inner), False
        trial = 0
        # -------- transposed stream + fused epilogue, forward stream
        grad = bwd(r); n_bwd += 1
        xn, xnt, vn, gd, dd, reg, mx = prox_epilogue(algo, grad, xt, v, theta, t, mu, thres)
        prod = fwd(np.concatenate([xn, xnt], axis=1)); n_fwd += 1
        Axn, Axnt = prod[:, :l], prod[:, l:]
        rr = np.sum((Axn - b) ** 2)
        cnt = np.sum(np.abs(xn) > 1e-6 * mx)
    tt = _round_us(time.time_ns() - t_begin) / 1e6
    if accel and not ran_any:
        raise UnboundLocalError("FProxGD ran no iteration (maxit=0), as in the reference")
    out = {"tt": tt, "fval": f_cur, "f_hist": hist.f, "f_hist_best": hist.fbest}
    if trace is not None:
        trace.update(n_fwd=n_fwd, n_bwd=n_bwd, sparsity_hist=hist.s)
    return x, hist.k, out
