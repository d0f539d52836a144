"""GPU parity tests of the drop-in solvers against the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py) and against the oracle on fresh seeded inputs.

Bars (BASELINE.json north_star): fp64 mode -- relative objective difference <= 1e-10 and identical
non-zero-group support (we also require the identical iteration count and f_hist <= 1e-10);
fp32 mode -- relative objective difference <= 1e-5 versus the fp64 reference on the fp32-rounded inputs.
"""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu

FP64_REL = 1e-10
FP32_REL = 1e-5

# Cases whose *trajectory* is chaotic under benign re-rounding of the products with A (a different BLAS
# summation order changes k by +-1..2 and f_hist by percents mid-run; tests/golden/sensitivity_probe.py
# reproduces this with NumPy alone).  For those the bar is BASELINE.json's: final objective <= 1e-10
# relative, identical support; k within +-3.
ROUNDING_CHAOTIC = {("thres1e-8", "ProxGD")}

FP64_CASES = ["fixture", "seed114514", "seed1", "seed2", "fixture_maxit3", "fixture_maxit40", "ragged", "l1",
              "l3", "l5", "harsh_ls", "mid", "diminishing", "diminishing2", "thres1e-8", "fixed_step"]


def _fn(algo):
    from gl_FProxGD_primal import gl_FProxGD_primal
    from gl_ProxGD_primal import gl_ProxGD_primal
    return gl_ProxGD_primal if algo == "ProxGD" else gl_FProxGD_primal


def _support(x):
    ax = np.abs(x)
    return np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1)), np.flatnonzero(np.sqrt((x * x).sum(1)) > 0)


# cases whose A fits in one cluster's shared memory also run on the persistent kernel (K4)
PERSISTENT_FITS = {"fixture", "seed114514", "seed1", "seed2", "fixture_maxit3", "fixture_maxit40", "ragged", "l1",
                   "l3", "harsh_ls", "diminishing", "diminishing2", "thres1e-8", "fixed_step"}


@pytest.mark.parametrize("backend", ["stream", "persistent"])
@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
@pytest.mark.parametrize("name", FP64_CASES)
def test_fp64_solver_matches_reference(name, algo, backend):
    if name == "fixed_step" and algo == "ProxGD" and backend == "stream":
        pytest.skip("7500 launch-bound iterations; the persistent backend covers it")
    if backend == "persistent" and name not in PERSISTENT_FITS:
        pytest.skip("A does not fit in one cluster (or l > 4)")
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    keep = (A.copy(), b.copy(), x0.copy())
    x, k, out = _fn(algo)(x0, A, b, mu, dict(rec["opts"], backend=backend))
    assert out["glprox"]["backend"] == backend
    assert np.array_equal(A, keep[0]) and np.array_equal(b, keep[1]) and np.array_equal(x0, keep[2])
    assert isinstance(k, int) and isinstance(out["tt"], float) and isinstance(x, np.ndarray)
    assert {"tt", "fval", "f_hist", "f_hist_best"} <= set(out)
    fh, ref = np.asarray(out["f_hist"]), rec[algo + "_f_hist"]
    if (name, algo) in ROUNDING_CHAOTIC:
        assert abs(k - int(rec[algo + "_k"])) <= 3 and len(fh) == k
        assert np.max(np.abs(fh[:200] - ref[:200]) / np.abs(ref[:200])) <= FP64_REL
    else:
        assert k == int(rec[algo + "_k"])
        assert fh.shape == ref.shape
        assert np.max(np.abs(fh - ref) / np.abs(ref)) <= FP64_REL
    assert abs(out["fval"] - rec[algo + "_fval"]) <= FP64_REL * abs(rec[algo + "_fval"])
    assert np.array_equal(np.asarray(out["f_hist_best"]), np.minimum.accumulate(fh))
    rule, nz = _support(x)
    assert np.array_equal(rule, rec[algo + "_support_rule"])
    assert np.array_equal(nz, rec[algo + "_support_nz"])
    assert np.max(np.abs(x - rec[algo + "_x"])) <= 1e-8
    # A was streamed ~twice per iteration (once transposed per iteration, once forward per trial)
    assert out["glprox"]["n_tpass"] <= k
    if backend == "persistent":
        assert out["glprox"]["n_launches"] == 1


def test_backend_selection():
    from glprox import _lib
    A, b, x0, u, mu = orc.gen_data(3, 64, 128, 2)
    o = {"maxit": 5, "alpha0": 1e-3}
    assert _fn("FProxGD")(x0, A, b, mu, o)[2]["glprox"]["backend"] == "persistent"   # auto: fits
    assert _fn("FProxGD")(x0, A, b, mu, dict(o, backend="stream"))[2]["glprox"]["backend"] == "stream"
    A, b, x0, u, mu = orc.gen_data(3, 512, 4096, 2)
    o = {"maxit": 2, "alpha0": 1e-4}
    assert _fn("FProxGD")(x0, A, b, mu, o)[2]["glprox"]["backend"] == "stream"       # auto: too large
    with pytest.raises(_lib.GlproxError):
        _fn("FProxGD")(x0, A, b, mu, dict(o, backend="persistent"))
    with pytest.raises(ValueError):
        _fn("FProxGD")(x0, A, b, mu, dict(o, backend="cpu"))


@pytest.mark.parametrize("backend", ["stream", "persistent"])
@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
@pytest.mark.parametrize("name", ["mid_f32in", "fixture_f32in"])
def test_fp32_solver_matches_fp64_reference_on_rounded_inputs(name, algo, backend):
    if backend == "persistent" and name == "mid_f32in":
        pytest.skip("A does not fit in one cluster")
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)  # fp32-representable values held in fp64
    x, k, out = _fn(algo)(x0.astype(np.float32), A.astype(np.float32), b.astype(np.float32), mu,
                          dict(rec["opts"], backend=backend))
    assert out["glprox"]["backend"] == backend
    assert x.dtype == np.float32
    ref_f = rec[algo + "_fval"]
    # (a) the objective the solver reports, (b) the objective of its x re-evaluated in fp64
    assert abs(out["fval"] - ref_f) <= FP32_REL * abs(ref_f)
    f_true = orc.objective(A, b, x.astype(np.float64), mu)
    assert abs(f_true - ref_f) <= FP32_REL * abs(ref_f)
    if name == "mid_f32in":  # bounded horizon: iteration count is fixed by maxit
        assert k == int(rec[algo + "_k"])
        fh, ref = np.asarray(out["f_hist"]), rec[algo + "_f_hist"]
        assert np.max(np.abs(fh - ref) / np.abs(ref)) <= FP32_REL
    else:
        rule, _ = _support(x.astype(np.float64))
        assert np.array_equal(rule, rec[algo + "_support_rule"])


def test_edge_cases_follow_the_reference():
    A, b, x0, u, mu = orc.gen_data(3, 32, 64, 2)
    x, k, out = _fn("ProxGD")(x0, A, b, mu, {"maxit": 0})
    assert k == 0 and np.array_equal(x, x0) and out["f_hist"] == []
    assert abs(out["fval"] - orc.objective(A, b, x0, mu)) <= 1e-12 * abs(out["fval"])
    with pytest.raises(UnboundLocalError):
        _fn("FProxGD")(x0, A, b, mu, {"maxit": 0})
    with pytest.raises(ValueError):
        _fn("ProxGD")(x0, A, b, mu, {"step_type": "no_such_rule"})
    # unknown keys are ignored, like {**default_opts, **opts}
    x2, k2, _ = _fn("FProxGD")(x0, A, b, mu, {"maxit": 5, "alpha0": 1e-3, "no_such_key": 1})
    xr, kr, _ = orc.solve("FProxGD", x0, A, b, mu, {"maxit": 5, "alpha0": 1e-3})
    assert k2 == kr and np.max(np.abs(x2 - xr)) <= 1e-10


@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
def test_device_tensor_path_and_resident_handle(algo):
    import torch
    from glprox import GroupLassoProblem
    A, b, x0, u, mu = orc.gen_data(21, 300, 700, 2)
    opts = {"maxit": 30, "alpha0": float(1.0 / (np.sqrt(300) + np.sqrt(700)) ** 2)}
    xr, kr, outr = orc.solve(algo, x0, A, b, mu, opts)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    for _ in range(2):  # the handle is reusable; state from the first solve must not leak
        x, k, out = prob.solve(algo, torch.from_numpy(x0).cuda(), mu, opts)
        assert k == kr and x.is_cuda
        assert abs(out["fval"] - outr["fval"]) <= FP64_REL * abs(outr["fval"])
        assert np.max(np.abs(x.cpu().numpy() - xr)) <= 1e-9
    prob.close()
    x, k, out = _fn(algo)(torch.from_numpy(x0).cuda(), torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda(), mu, opts)
    assert k == kr and abs(out["fval"] - outr["fval"]) <= FP64_REL * abs(outr["fval"])


def test_large_size_properties_fp32():
    """Size-independent checks at a bandwidth-bound size (no CPU oracle run): objective history is
    finite, the reported objective equals an independent evaluation through K1, and the forward /
    transposed streams are adjoint: <A x, r> == <x, A^T r>."""
    import ctypes as C
    import torch
    from glprox import GroupLassoProblem, _lib
    m, n, l = 4096, 16384, 2
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(5)
    A = torch.randn((m, n), generator=g, device="cuda", dtype=torch.float32)
    u = torch.zeros((n, l), device="cuda", dtype=torch.float32)
    idx = torch.randperm(n, generator=g, device="cuda")[: n // 10]
    u[idx] = torch.randn((n // 10, l), generator=g, device="cuda", dtype=torch.float32)
    prob = GroupLassoProblem(A, torch.zeros((m, l), device="cuda", dtype=torch.float32))
    Au = torch.empty((m, l), dtype=torch.float64, device="cuda")
    _lib.check(lib.glp_fwd(prob.handle, u.data_ptr(), l, Au.data_ptr(), None))
    prob.close()
    b = Au.float()
    prob = GroupLassoProblem(A, b)
    x0 = torch.randn((n, l), generator=g, device="cuda", dtype=torch.float32)
    mu = 1e-2
    opts = {"maxit": 20, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    x, k, out = prob.solve("FProxGD", x0, mu, opts)
    fh = np.asarray(out["f_hist"])
    assert k == 60 and np.all(np.isfinite(fh)) and fh[-1] < fh[0] * 1e-2
    # objective re-evaluated independently through K1 + torch norms
    Ax = torch.empty((m, l), dtype=torch.float64, device="cuda")
    rr = torch.zeros(1, dtype=torch.float64, device="cuda")
    _lib.check(lib.glp_fwd(prob.handle, x.data_ptr(), l, Ax.data_ptr(), rr.data_ptr()))
    f_chk = 0.5 * rr.item() + mu * x.double().norm(dim=1).sum().item()
    assert abs(f_chk - out["fval"]) <= 1e-6 * abs(f_chk)
    # adjointness
    r = torch.randn((m, l), generator=g, device="cuda", dtype=torch.float64)
    G = torch.empty((n, l), dtype=torch.float64, device="cuda")
    null = C.c_void_p()
    _lib.check(lib.glp_tpass_prox(prob.handle, None, r.data_ptr(), G.data_ptr(), null, null, null, null, null, null))
    lhs = (Ax * r.float().double()).sum().item()
    rhs = (x.double() * G).sum().item()
    assert abs(lhs - rhs) <= 1e-5 * (Ax.abs() * r.abs()).sum().item()
    prob.close()


def test_full_size_properties_c4_fp32():
    """BASELINE.json configs[3] at full size (m=32768 n=131072 l=2 fp32, 17 GB of A): no CPU oracle can run
    this, so parity rests on size-independent properties -- forward/transposed adjointness, linearity of the
    forward stream, the solver's reported objective re-evaluated independently through K1, a monotone
    objective under the 1/L step, and run-to-run bit-reproducibility."""
    import ctypes as C
    import torch
    from glprox import GroupLassoProblem, _lib
    if torch.cuda.mem_get_info()[0] < 40e9:
        pytest.skip("needs 40 GB of free HBM")
    lib = _lib.load()
    m, n, l = 32768, 131072, 2
    st = torch.cuda.current_stream().cuda_stream
    A = torch.empty((m, n), dtype=torch.float32, device="cuda")
    _lib.check(lib.glp_randn(A.data_ptr(), _lib.GLP_F32, m, n, n, 12345, 0, 0, st))
    g = torch.Generator(device="cuda").manual_seed(9)
    u = torch.zeros((n, l), device="cuda")
    idx = torch.randperm(n, generator=g, device="cuda")[: n // 10]
    u[idx] = torch.randn((n // 10, l), generator=g, device="cuda")
    prob = GroupLassoProblem(A, torch.zeros((m, l), device="cuda"))
    null = C.c_void_p()

    def fwd(x):
        out = torch.empty((m, x.shape[1]), dtype=torch.float64, device="cuda")
        _lib.check(lib.glp_fwd(prob.handle, x.contiguous().data_ptr(), x.shape[1], out.data_ptr(), None))
        return out

    def tps(r):
        out = torch.empty((n, l), dtype=torch.float64, device="cuda")
        _lib.check(lib.glp_tpass_prox(prob.handle, None, r.contiguous().data_ptr(), out.data_ptr(), null, null, null, null, null, null))
        return out

    Au = fwd(u)
    # linearity: A(2u - w) == 2Au - Aw (fp32 operands, fp64 outputs)
    w = torch.randn((n, l), generator=g, device="cuda")
    lin = fwd(2 * u - w) - (2 * Au - fwd(w))
    assert lin.abs().max().item() <= 2e-5 * (2 * Au).abs().max().item()
    # adjointness <A u, r> == <u, A^T r>
    r = torch.randn((m, l), generator=g, device="cuda", dtype=torch.float64)
    lhs = (Au * r.float().double()).sum().item()
    rhs = (u.double() * tps(r)).sum().item()
    assert abs(lhs - rhs) <= 1e-5 * (Au.abs() * r.abs()).sum().item()
    prob.close()
    # solver: 2 x 9 iterations, bit-identical reruns, monotone objective, independent re-evaluation
    b = Au.float()
    prob = GroupLassoProblem(A, b)
    x0 = torch.randn((n, l), generator=g, device="cuda")
    mu, opts = 1e-2, {"maxit": 3, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    x1, k1, o1 = prob.solve("FProxGD", x0, mu, opts)
    x2, k2, o2 = prob.solve("FProxGD", x0, mu, opts)
    assert k1 == k2 == 9 and torch.equal(x1, x2) and o1["f_hist"] == o2["f_hist"]
    xp, kp, op_ = prob.solve("ProxGD", x0, mu, opts)
    fp = np.asarray(op_["f_hist"])
    assert np.all(np.diff(fp) <= 0) and np.all(np.isfinite(o1["f_hist"]))
    assert kp + 1 <= op_["glprox"]["n_fwd"] <= kp + 3 and op_["glprox"]["n_tpass"] == kp   # step ~ 1/L: first trials accepted
    rr = torch.zeros(1, dtype=torch.float64, device="cuda")
    Ax = torch.empty((m, l), dtype=torch.float64, device="cuda")
    _lib.check(lib.glp_fwd(prob.handle, x1.data_ptr(), l, Ax.data_ptr(), rr.data_ptr()))
    f_chk = 0.5 * rr.item() + mu * x1.double().norm(dim=1).sum().item()
    assert abs(f_chk - o1["fval"]) <= 1e-6 * abs(f_chk)
    prob.close()
