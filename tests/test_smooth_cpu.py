"""CPU tests of SURVEY.md 8(f) row 1 -- gl_FGD_primal / gl_GD_primal / gl_SGD_primal on the hot path's two
streams: the oracle restatement against golden vectors of the unmodified reference, and the device state
machine (csrc/glp_decide.h, host build) driven through the GPU schedule."""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from gpu_emulator import emulate
from oracle import glprox_oracle as orc

SMOOTH = ["smooth_maxit40", "smooth_ragged", "smooth_l3", "smooth_fixed", "smooth_dim2", "smooth_seed2", "smooth_csf",
          "smooth_csf_ragged"]
# FGD (delta = 1e-6: an almost non-smooth gradient, no prox) is chaotic under ANY re-rounding of the products
# on long runs -- NumPy with a reversed summation order moves k from 2034 to 2045 on the fixture -- so its long
# cases are held to the final objective only; GD / SGD and short FGD runs are held to the whole trajectory.
FGD_LONG = {"smooth_fixed", "smooth_dim2", "smooth_seed2", "smooth_fixture", "smooth_csf"}


@pytest.mark.parametrize("algo", ["FGD", "GD", "SGD"])
@pytest.mark.parametrize("name", SMOOTH)
def test_oracle_restatement_is_bit_identical_to_reference(name, algo):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    x, k, out = orc.solve_smooth(algo, x0, A, b, mu, rec["opts"])
    assert k == int(rec[algo + "_k"])
    assert np.array_equal(np.asarray(out["f_hist"]), rec[algo + "_f_hist"])
    assert np.array_equal(x, rec[algo + "_x"]) and out["fval"] == rec[algo + "_fval"]


@pytest.mark.parametrize("algo", ["FGD", "GD", "SGD"])
@pytest.mark.parametrize("name", SMOOTH)
def test_device_state_machine_reproduces_reference(name, algo):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    x, k, out, info = emulate(algo, x0, A, b, mu, rec["opts"])
    ref = rec[algo + "_f_hist"]
    fh = np.asarray(out["f_hist"])
    if algo == "FGD" and name in FGD_LONG:
        assert abs(k - int(rec[algo + "_k"])) <= 0.15 * int(rec[algo + "_k"])   # the stop rule fires on a chaotic tail
        assert abs(out["fval"] - rec[algo + "_fval"]) <= 1e-6 * abs(rec[algo + "_fval"])
        assert np.max(np.abs(fh[:100] - ref[:100]) / np.abs(ref[:100])) <= 1e-10
        return
    assert k == int(rec[algo + "_k"])
    assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-11
    assert abs(out["fval"] - rec[algo + "_fval"]) <= 1e-11 * abs(rec[algo + "_fval"])
    assert np.max(np.abs(x - rec[algo + "_x"])) <= 1e-9


def test_smooth_option_semantics():
    from glprox import solver
    assert solver.merge_opts("GD", {})["delta"] == 1e-3 and solver.merge_opts("FGD", {})["delta"] == 1e-6
    assert solver.merge_opts("SGD", {"maxit": 5})["maxit"] == 5 and solver.merge_opts("SGD", {})["maxit"] == 2100
    with pytest.raises(ValueError):
        solver.pack_opts("GD", solver.merge_opts("GD", {"step_type": "line_search"}))
    # continuous_subgradient_flag is honoured (alpha0 <- 1 / lambda_max(A^T A) by the handle before packing), not refused
    o = solver.pack_opts("SGD", solver.merge_opts("SGD", {"continuous_subgradient_flag": True}))
    assert o.alpha0 == 1e-3
    o = solver.pack_opts("FGD", solver.merge_opts("FGD", {}))
    assert (o.maxit, o.delta, o.line_search_attenuation_coeffi, o.maxit_line_search_iter) == (1500, 1e-6, 0.98, 5)
