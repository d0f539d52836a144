"""GPU parity tests of gl_Admm_primal and gl_Alm_dual through the drop-in modules against golden vectors of the
unmodified reference.  gl_Alm_dual: same iteration count, objective history <= 1e-10.  gl_Admm_primal: same iteration
count, objective history <= 5e-9 -- the reference's own iteration amplifies any re-rounding of its solve to ~4e-10
(tests/test_admm2_cpu.py::test_admm_primal_solve_identity_and_rounding_sensitivity)."""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu


def _primal():
    import gl_ADMM_primal
    return gl_ADMM_primal.gl_Admm_primal


def _alm():
    import gl_ALM_dual
    return gl_ALM_dual.gl_Alm_dual


def _check(rec, algo, x, k, out, tol_f, tol_x):
    assert isinstance(k, int) and k == int(rec[algo + "_k"])
    fh, ref = np.asarray(out["f_hist"]), rec[algo + "_f_hist"]
    assert fh.shape == ref.shape and np.max(np.abs(fh - ref) / np.abs(ref)) <= tol_f
    assert np.array_equal(np.asarray(out["f_hist_best"]), np.minimum.accumulate(fh))
    assert abs(out["fval"] - rec[algo + "_fval"]) <= tol_f * abs(rec[algo + "_fval"])
    assert np.max(np.abs(x - rec[algo + "_x"])) <= tol_x * max(1.0, np.max(np.abs(rec[algo + "_x"])))
    ax = np.abs(x)
    assert np.array_equal(np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1)), rec[algo + "_support_rule"])
    assert abs(out["glprox"]["sparsity"] - float(rec[algo + "_sparsity"])) <= 1e-12


@pytest.mark.parametrize("name", ["admmp_fixture", "admmp_seed1", "admmp_maxit7", "admmp_ragged", "admmp_l1", "admmp_l3",
                                  "admmp_dimin", "admmp_dimin2", "admmp_rho1"])
def test_fp64_admm_primal_matches_the_reference(name):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    keep = (A.copy(), b.copy(), x0.copy())
    x, k, out = _primal()(x0, A, b, mu, dict(rec["opts"]))
    assert np.array_equal(A, keep[0]) and np.array_equal(b, keep[1]) and np.array_equal(x0, keep[2])
    _check(rec, "ADMM_primal", x, k, out, 5e-9, 2e-9)
    assert out["glprox"]["n_tpass"] == k and out["glprox"]["n_fwd"] == k + 1   # A is streamed twice per iteration


@pytest.mark.parametrize("name", ["alm_fixture", "alm_maxit3", "alm_ragged", "alm_l1", "alm_l3", "alm_loose"])
def test_fp64_alm_dual_matches_the_reference(name):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    keep = (A.copy(), b.copy(), x0.copy())
    x, k, out = _alm()(x0, A, b, mu, dict(rec["opts"]))
    assert np.array_equal(A, keep[0]) and np.array_equal(b, keep[1]) and np.array_equal(x0, keep[2])
    _check(rec, "ALM_dual", x, k, out, 1e-10, 1e-9)


@pytest.mark.parametrize("which", ["ADMM_primal", "ALM_dual"])
def test_fp32_inputs_are_widened_and_match_the_fp64_reference_on_rounded_inputs(which):
    rec = load_golden("admmp_f32in" if which == "ADMM_primal" else "alm_f32in")
    A, b, x0, u, mu = golden_inputs(rec)
    fn = _primal() if which == "ADMM_primal" else _alm()
    x, k, out = fn(x0.astype(np.float32), A.astype(np.float32), b.astype(np.float32), mu, dict(rec["opts"]))
    assert x.dtype == np.float32 and k == int(rec[which + "_k"])
    fh, ref = np.asarray(out["f_hist"]), rec[which + "_f_hist"]
    assert np.max(np.abs(fh - ref) / np.abs(ref)) <= (5e-9 if which == "ADMM_primal" else 1e-10)


def test_edge_cases():
    from glprox import _lib
    A, b, x0, u, mu = orc.gen_data(3, 32, 64, 2)
    for fn, ofn in ((_primal(), orc.solve_admm_primal), (_alm(), orc.solve_alm_dual)):
        x, k, out = fn(x0, A, b, mu, {"maxit": 0})
        assert k == 0 and np.array_equal(x, x0) and out["f_hist"] == []
        assert abs(out["fval"] - orc.objective(A, b, x0, mu)) <= 1e-12 * abs(out["fval"])
        x2, k2, _ = fn(x0, A, b, mu, {"maxit": 3, "no_such_key": 1})       # unknown keys are ignored
        xr, kr, _ = ofn(x0, A, b, mu, {"maxit": 3})
        assert k2 == kr and np.max(np.abs(x2 - xr)) <= 1e-9
        A9, b9, x09, _, _ = orc.gen_data(3, 32, 64, 9)
        with pytest.raises(_lib.GlproxError):                              # l > 8 is refused loudly
            fn(x09, A9, b9, mu, {"maxit": 2})
    with pytest.raises(TypeError):
        _primal()(x0, A, b, mu, {"step_type": "line_search"})


def test_handle_path_and_streaming_size():
    """fp64 1024 x 4096 (8 panels of the blocked factorisation) against the oracle; fp32 handle -> widened twin."""
    import torch
    from glprox import GroupLassoProblem
    A, b, x0, u, mu = orc.gen_data(11, 1024, 4096, 2)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x, k, out = prob.solve_admm_primal(torch.from_numpy(x0).cuda(), mu, {"maxit": 12})
    xr, kr, outr = orc.solve_admm_primal(x0, A, b, mu, {"maxit": 12})
    ref = np.asarray(outr["f_hist"])
    assert k == kr and np.max(np.abs(np.asarray(out["f_hist"]) - ref) / np.abs(ref)) <= 5e-8
    x, k, out = prob.solve_alm_dual(torch.from_numpy(x0).cuda(), mu, {"maxit": 2})
    xr, kr, outr = orc.solve_alm_dual(x0, A, b, mu, {"maxit": 2})
    ref = np.asarray(outr["f_hist"])
    assert k == kr and np.max(np.abs(np.asarray(out["f_hist"]) - ref) / np.abs(ref)) <= 1e-9
    prob.close()
    p32 = GroupLassoProblem(torch.from_numpy(A[:256, :512].astype(np.float32)).cuda(), torch.from_numpy(b[:256].astype(np.float32)).cuda())
    x, k, out = p32.solve_admm_primal(torch.from_numpy(x0[:512].astype(np.float32)).cuda(), mu, {"maxit": 4})
    assert x.dtype == torch.float32 and k == 4
    p32.close()
