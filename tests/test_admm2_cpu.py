"""gl_Admm_primal (code/gl_ADMM_primal.py) and gl_Alm_dual (code/gl_ALM_dual.py): the oracle's restatements against
golden vectors produced by the unmodified reference (tests/golden/make_golden.py, ADMMP_CASES / ALM_CASES), the
identities the CUDA path rests on (csrc/glp_admm_primal.cuh, csrc/glp_alm.cuh) checked in NumPy, and host-side
option handling."""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from oracle import glprox_oracle as orc

ADMMP = ["admmp_fixture", "admmp_seed1", "admmp_maxit7", "admmp_ragged", "admmp_l1", "admmp_l3", "admmp_dimin",
         "admmp_dimin2", "admmp_rho1", "admmp_f32in"]
ALM = ["alm_fixture", "alm_maxit3", "alm_ragged", "alm_l1", "alm_l3", "alm_loose", "alm_f32in"]


def _check_bit_identical(rec, algo, x, k, out):
    assert k == int(rec[algo + "_k"])
    assert np.array_equal(np.asarray(out["f_hist"]), rec[algo + "_f_hist"])
    assert np.array_equal(np.asarray(out["f_hist_best"]), rec[algo + "_f_hist_best"])
    assert out["fval"] == rec[algo + "_fval"] and np.array_equal(x, rec[algo + "_x"])


@pytest.mark.parametrize("name", ADMMP)
def test_oracle_admm_primal_is_bit_identical_to_the_reference(name):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    keep = (A.copy(), b.copy(), x0.copy())
    x, k, out = orc.solve_admm_primal(x0, A, b, mu, dict(rec["opts"]))
    assert np.array_equal(A, keep[0]) and np.array_equal(b, keep[1]) and np.array_equal(x0, keep[2])
    _check_bit_identical(rec, "ADMM_primal", x, k, out)


@pytest.mark.parametrize("name", [n for n in ALM if n != "alm_fixture"])
def test_oracle_alm_dual_is_bit_identical_to_the_reference(name):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    x, k, out = orc.solve_alm_dual(x0, A, b, mu, dict(rec["opts"]))
    _check_bit_identical(rec, "ALM_dual", x, k, out)


def test_published_rows_of_the_report():
    # doc/report.md:570  | ALM Dual    | 10.40 | 39 | 6.10389E-01 | 0.3467 | 4.39E-05 |
    # doc/report.md:572  | ADMM Primal | 6.23  | 63 | 6.10421E-01 | 0.0996 | 9.01E-05 |
    fro = lambda z: np.sqrt(np.sum(z * z))
    for name, algo, k, fval, sp, err in (("alm_fixture", "ALM_dual", 39, "6.10389E-01", "0.3467", "4.39E-05"),
                                         ("admmp_fixture", "ADMM_primal", 63, "6.10421E-01", "0.0996", "9.01E-05")):
        rec = load_golden(name)
        A, b, x0, u, mu = golden_inputs(rec)
        x = rec[algo + "_x"]
        assert int(rec[algo + "_k"]) == k and "%6.5E" % rec[algo + "_fval"] == fval
        assert "%6.4f" % (np.sum(np.abs(x) > 1e-6 * np.max(np.abs(x))) / x.size) == sp
        assert "%3.2E" % (fro(x - u) / (1 + fro(x))) == err


def test_alm_inner_problem_identities():
    """csrc/glp_alm.cuh: with T = (I + rho A A^T)^-1 the n x n matrices of gl_ALM_dual.py:33-40 satisfy
    Q = rho (I - rho A^T T A) and J = rho H -- what lets the CUDA path apply them as streams."""
    A, b, x0, u, mu = orc.gen_data(4, 48, 120, 2)
    rho = 1e2
    T = np.linalg.inv(np.identity(48) + rho * A @ A.T)
    D = A @ x0 - b
    E = T @ D
    F = rho * T @ A
    G = np.identity(120) - A.T @ F
    H = A.T @ E - x0 / rho
    Q = F.T @ F + rho * G.T @ G
    J = rho * G.T @ H - F.T @ E - F.T @ b
    assert np.max(np.abs(Q - rho * G)) <= 1e-9 * np.max(np.abs(Q))
    assert np.max(np.abs(J - rho * H)) <= 1e-9 * np.max(np.abs(J))


def test_admm_primal_solve_identity_and_rounding_sensitivity():
    """csrc/glp_admm_primal.cuh: (rho I + A^T A)^-1 v = (v - A^T G^-1 A v / rho) / rho with G = I + A A^T / rho.
    And why its parity bar is 5e-9 rather than 1e-10: replacing numpy.linalg.solve (LU on the triangular factors, what
    the reference calls) by triangular solves with the SAME factor already moves the objective history by > 1e-10."""
    import scipy.linalg as sl
    rec = load_golden("admmp_fixture")
    A, b, x0, u, mu = golden_inputs(rec)
    m, n = A.shape
    rho = 1e-2
    v = A.T @ b - x0 + rho * x0
    direct = np.linalg.solve(rho * np.identity(n) + A.T @ A, v)
    G = np.identity(m) + A @ A.T / rho
    wood = (v - A.T @ np.linalg.solve(G, A @ v) / rho) / rho
    assert np.max(np.abs(direct - wood)) <= 1e-9 * np.max(np.abs(direct))

    o = dict(orc.ADMM_PRIMAL_DEFAULTS)
    L = np.linalg.cholesky(rho * np.identity(n) + A.T @ A)
    ATb = A.T @ b
    x_k, y_k, z_k, fh = x0.copy(), x0, x0, []
    for k in range(1, 64):
        y = sl.solve_triangular(L.T, sl.solve_triangular(L, ATb - z_k + rho * x_k, lower=True), lower=False)
        w = x_k - o["eta_0"] * rho * (x_k - y - z_k / rho)
        rn = np.linalg.norm(w, axis=1).reshape(-1, 1)
        x = w * np.clip(rn - o["eta_0"] * mu, a_min=0, a_max=None) / ((rn < o["thres"]) + rn)
        z_k = z_k - o["tau"] * rho * (x - y)
        x_k, y_k = x, y
        fh.append(orc.objective(A, b, x_k, mu))
    ref = rec["ADMM_primal_f_hist"]
    dev = np.max(np.abs(np.asarray(fh) - ref) / np.abs(ref))
    assert 1e-10 < dev < 5e-9, dev


def test_edge_cases_of_the_restatements():
    A, b, x0, u, mu = orc.gen_data(3, 32, 64, 2)
    for fn in (orc.solve_admm_primal, orc.solve_alm_dual):
        x, k, out = fn(x0, A, b, mu, {"maxit": 0})
        assert k == 0 and np.array_equal(x, x0) and out["f_hist"] == []
        assert out["fval"] == orc.objective(A, b, x0, mu)
    with pytest.raises(TypeError):      # the reference's set_step returns None for an unknown name (:66-72)
        orc.solve_admm_primal(x0, A, b, mu, {"step_type": "line_search"})


def test_host_option_handling():
    import math
    from glprox import solver
    assert solver.ADMM_PRIMAL_DEFAULT_OPTS == {"maxit": 100, "thres": 1e-3, "tau": (1 + math.sqrt(5)) * 0.5, "rho": 1e-2,
                                               "eta_0": 100, "converge_len": 10, "converge_thres": 1e-5,
                                               "step_type": "fixed"}
    assert solver.ADMM_PRIMAL_DEFAULT_OPTS == orc.ADMM_PRIMAL_DEFAULTS
    assert solver.ALM_DUAL_DEFAULT_OPTS == orc.ALM_DUAL_DEFAULTS == solver.ADMM_DUAL_DEFAULT_OPTS
    import gl_ADMM_primal, gl_ALM_dual
    assert callable(gl_ADMM_primal.gl_Admm_primal) and callable(gl_ALM_dual.gl_Alm_dual)
    with pytest.raises(ValueError):
        gl_ADMM_primal.gl_Admm_primal(np.zeros((5, 2)), np.zeros((3, 4)), np.zeros((3, 2)), 0.1, {})
    with pytest.raises(ValueError):
        gl_ALM_dual.gl_Alm_dual(np.zeros((4, 2)), np.zeros((3, 4)), np.zeros((2, 2)), 0.1, {})
