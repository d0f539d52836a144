"""the oracle's restatement of gl_Admm_dual (code/gl_ADMM_dual.py) against golden vectors
produced by the unmodified reference (tests/golden/make_golden.py, ADMM_CASES), plus host-side option handling."""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from oracle import glprox_oracle as orc

CASES = ["admm_fixture", "admm_seed1", "admm_seed2", "admm_maxit7", "admm_ragged", "admm_l1", "admm_rho10",
         "admm_loose", "admm_f32in"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_admm_dual_is_bit_identical_to_the_reference(name):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    keep = (A.copy(), b.copy(), x0.copy())
    x, k, out = orc.solve_admm_dual(x0, A, b, mu, dict(rec["opts"]))
    assert np.array_equal(A, keep[0]) and np.array_equal(b, keep[1]) and np.array_equal(x0, keep[2])
    assert k == int(rec["ADMM_dual_k"])
    assert np.array_equal(np.asarray(out["f_hist"]), rec["ADMM_dual_f_hist"])
    assert np.array_equal(np.asarray(out["f_hist_best"]), rec["ADMM_dual_f_hist_best"])
    assert out["fval"] == rec["ADMM_dual_fval"] and np.array_equal(x, rec["ADMM_dual_x"])


def test_published_row_of_the_report():
    # doc/report.md:571  | ADMM Dual | 2.12 | 71 | 6.10786E-01 | 0.0996 | 1.98E-04 |
    rec = load_golden("admm_fixture")
    A, b, x0, u, mu = golden_inputs(rec)
    assert int(rec["ADMM_dual_k"]) == 71 and "%6.5E" % rec["ADMM_dual_fval"] == "6.10786E-01"
    x = rec["ADMM_dual_x"]
    assert "%6.4f" % (np.sum(np.abs(x) > 1e-6 * np.max(np.abs(x))) / x.size) == "0.0996"
    fro = lambda z: np.sqrt(np.sum(z * z))
    assert "%3.2E" % (fro(x - u) / (1 + fro(x))) == "1.98E-04"


def test_admm_edge_cases_of_the_restatement():
    A, b, x0, u, mu = orc.gen_data(3, 32, 64, 2)
    x, k, out = orc.solve_admm_dual(x0, A, b, mu, {"maxit": 0})
    assert k == 0 and np.array_equal(x, x0) and out["f_hist"] == []
    assert out["fval"] == orc.objective(A, b, x0, mu)
