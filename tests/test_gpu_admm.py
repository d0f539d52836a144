"""GPU parity tests of SURVEY.md 8(f) row 4: gl_Admm_dual through the drop-in module against golden vectors of
the unmodified reference (fp64: 1e-10 on the objective history, same iteration count; fp32: 1e-5)."""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu


def _fn():
    import gl_ADMM_dual
    return gl_ADMM_dual.gl_Admm_dual


@pytest.mark.parametrize("name", ["admm_fixture", "admm_seed1", "admm_seed2", "admm_maxit7", "admm_ragged", "admm_l1",
                                  "admm_rho10", "admm_loose"])
def test_fp64_admm_dual_matches_the_reference(name):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    keep = (A.copy(), b.copy(), x0.copy())
    x, k, out = _fn()(x0, A, b, mu, dict(rec["opts"]))
    assert np.array_equal(A, keep[0]) and np.array_equal(b, keep[1]) and np.array_equal(x0, keep[2])
    assert isinstance(k, int) and k == int(rec["ADMM_dual_k"])
    fh, ref = np.asarray(out["f_hist"]), rec["ADMM_dual_f_hist"]
    assert fh.shape == ref.shape and np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10
    assert np.array_equal(np.asarray(out["f_hist_best"]), np.minimum.accumulate(fh))
    assert abs(out["fval"] - rec["ADMM_dual_fval"]) <= 1e-10 * abs(rec["ADMM_dual_fval"])
    assert np.max(np.abs(x - rec["ADMM_dual_x"])) <= 1e-9
    ax = np.abs(x)
    assert np.array_equal(np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1)), rec["ADMM_dual_support_rule"])
    assert abs(out["glprox"]["sparsity"] - float(rec["ADMM_dual_sparsity"])) <= 1e-12
    assert out["glprox"]["n_tpass"] == k   # one transposed and one forward stream per iteration (+1 forward)
    assert out["glprox"]["n_fwd"] == k + 1


def test_fp32_admm_dual_matches_fp64_reference_on_rounded_inputs():
    rec = load_golden("admm_f32in")
    A, b, x0, u, mu = golden_inputs(rec)
    x, k, out = _fn()(x0.astype(np.float32), A.astype(np.float32), b.astype(np.float32), mu, dict(rec["opts"]))
    assert x.dtype == np.float32
    ref_f = rec["ADMM_dual_fval"]
    assert abs(out["fval"] - ref_f) <= 1e-5 * abs(ref_f)
    assert abs(orc.objective(A, b, x.astype(np.float64), mu) - ref_f) <= 1e-5 * abs(ref_f)
    assert abs(k - int(rec["ADMM_dual_k"])) <= 3   # the residual test sits at thres = 1e-3: fp32 may cross it a step apart


def test_admm_edge_cases():
    from glprox import _lib
    A, b, x0, u, mu = orc.gen_data(3, 32, 64, 2)
    x, k, out = _fn()(x0, A, b, mu, {"maxit": 0})
    assert k == 0 and np.array_equal(x, x0) and out["f_hist"] == []
    assert abs(out["fval"] - orc.objective(A, b, x0, mu)) <= 1e-12 * abs(out["fval"])
    # unknown keys are ignored; out of scope shapes are refused loudly, not computed wrongly
    x2, k2, _ = _fn()(x0, A, b, mu, {"maxit": 5, "no_such_key": 1})
    xr, kr, _ = orc.solve_admm_dual(x0, A, b, mu, {"maxit": 5})
    assert k2 == kr and np.max(np.abs(x2 - xr)) <= 1e-10
    A9, b9, x09, _, _ = orc.gen_data(3, 32, 64, 9)
    with pytest.raises(_lib.GlproxError):
        _fn()(x09, A9, b9, mu, {})


def test_admm_on_a_device_resident_problem_and_in_the_driver_table(capsys):
    import torch
    from glprox import GroupLassoProblem
    rec = load_golden("admm_fixture")
    A, b, x0, u, mu = golden_inputs(rec)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    for rep in range(2):   # the factor buffer is reused
        x, k, out = prob.solve_admm_dual(torch.from_numpy(x0).cuda(), mu, {})
        assert k == 71 and abs(out["fval"] - rec["ADMM_dual_fval"]) <= 1e-10 * rec["ADMM_dual_fval"]
    # the proximal solvers still work on the same handle afterwards
    x2, k2, out2 = prob.solve("FProxGD", torch.from_numpy(x0).cuda(), mu, {"maxit": 5})
    assert k2 == 15
    met = prob.metrics(x, torch.from_numpy(u).cuda(), mu)
    assert "%3.2E" % met["err_exact"] == "1.98E-04"      # doc/report.md:571
    prob.close()
    import main as driver
    rows = driver.main(["--solvers", "ADMM_dual", "--log-level", "WARNING"])
    assert rows[0]["solver"] == "ADMM Dual" and rows[0]["iter"] == "   71" and rows[0]["optval"] == "6.10786E-01"
    assert rows[0]["sparsity"] == "0.0996" and rows[0]["err-to-exact"] == "1.98E-04"


@pytest.mark.parametrize("name", ["admm_fixture", "admm_ragged", "admm_l1", "admm_rho10"])
def test_blocked_path_matches_the_reference_on_the_goldens(name, monkeypatch):
    """The streaming-size schedule (csrc/glp_dense.cuh, glp_admm_host.inl: blocked fp64 Cholesky, Y = L^-T, the factor
    applied as two streams of Y through K2 / K1) forced onto the golden cases: same k, f_hist within 1e-10, same support."""
    monkeypatch.setenv("GLPROX_ADMM_BLOCKED", "1")
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    x, k, out = _fn()(x0, A, b, mu, dict(rec["opts"]))
    assert k == int(rec["ADMM_dual_k"])
    fh, ref = np.asarray(out["f_hist"]), rec["ADMM_dual_f_hist"]
    assert fh.shape == ref.shape and np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10
    assert abs(out["fval"] - rec["ADMM_dual_fval"]) <= 1e-10 * abs(rec["ADMM_dual_fval"])
    assert np.max(np.abs(x - rec["ADMM_dual_x"])) <= 1e-9
    ax = np.abs(x)
    assert np.array_equal(np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1)), rec["ADMM_dual_support_rule"])


def test_blocked_path_fp64_parity_at_1024_x_4096(monkeypatch):
    """parity with oracle.solve_admm_dual (code/gl_ADMM_dual.py:57-99) <= 1e-10 in fp64 at
    1024 x 4096 on the blocked schedule (8 diagonal blocks of 128)."""
    monkeypatch.setenv("GLPROX_ADMM_BLOCKED", "1")
    A, b, x0, u, mu = orc.gen_data(5, 1024, 4096, 2)
    opts = {"maxit": 40}
    x, k, out = _fn()(x0, A, b, mu, dict(opts))
    xr, kr, outr = orc.solve_admm_dual(x0, A, b, mu, dict(opts))
    assert k == kr
    fh, ref = np.asarray(out["f_hist"]), np.asarray(outr["f_hist"])
    assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10
    assert abs(out["fval"] - outr["fval"]) <= 1e-10 * abs(outr["fval"])
    assert np.max(np.abs(x - xr)) <= 1e-8 * max(1.0, np.max(np.abs(xr)))


@pytest.mark.parametrize("gram", ["tensor", "cuda"])
def test_streaming_size_fp32_admm_matches_the_fp64_oracle(gram, monkeypatch):
    """m = 2560 > 2048: the blocked schedule is the only one.  fp32 data: the Gram matrix comes from the tensor cores
    (K5 fp16 pairs, 'tensor') or from fp64 FMAs ('cuda'); bar: objective within 1e-5 of the fp64 oracle on the
    fp32-rounded inputs (north_star, fp32 mode)."""
    if gram == "cuda":
        monkeypatch.setenv("GLPROX_ADMM_GRAM", "cuda")
    m, n, l = 2560, 8192, 2
    A, b, x0, u, mu = orc.gen_data(9, m, n, l, np.float32)
    opts = {"maxit": 15}
    x, k, out = _fn()(x0, A, b, mu, dict(opts))
    xr, kr, outr = orc.solve_admm_dual(x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, dict(opts))
    assert k == kr == 15
    fh, ref = np.asarray(out["f_hist"]), np.asarray(outr["f_hist"])
    assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-5
    assert abs(out["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])
    assert out["glprox"]["n_fwd"] == k + 1 and out["glprox"]["n_tpass"] == k


@pytest.mark.parametrize("blocked", ["0", "1"])
@pytest.mark.parametrize("l", [3, 5, 8])
def test_admm_dual_with_more_than_two_right_hand_sides(l, blocked, monkeypatch):
    """l > 2: the residual test needs the matrix 2-norms of the n x l / m x l residuals
    (gl_ADMM_dual.py:87-88, numpy LA.norm(., 2)) = sqrt of the largest eigenvalue of their l x l Gram matrices, taken
    by Jacobi rotations on the device; run to the reference's own stop rule on both schedules."""
    monkeypatch.setenv("GLPROX_ADMM_BLOCKED", blocked)
    A, b, x0, u, mu = orc.gen_data(13, 192, 448, l)
    opts = {"maxit": 80, "converge_len": 5, "thres": 2e-2}
    x, k, out = _fn()(x0, A, b, mu, dict(opts))
    xr, kr, outr = orc.solve_admm_dual(x0, A, b, mu, dict(opts))
    assert k == kr and 5 < k < 80          # the stop rule fired, at the same iteration
    fh, ref = np.asarray(out["f_hist"]), np.asarray(outr["f_hist"])
    assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10
    assert np.max(np.abs(x - xr)) <= 1e-9 * max(1.0, np.max(np.abs(xr)))
