"""GPU parity tests of the individual kernels, called through the C ABI (include/glprox.h):
K1 forward stream + residual epilogue, K2 transposed stream + fused prox epilogue."""
import ctypes as C

import numpy as np
import pytest

from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu

SHAPES = [(256, 512), (193, 391), (64, 128), (1, 130), (70, 4), (1000, 3000), (2048, 8192)]


def _problem(m, n, l, dtype, seed=0):
    import torch
    from glprox import GroupLassoProblem
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((m, n))
    b = rng.standard_normal((m, l))
    if dtype == "f4":
        A, b = A.astype(np.float32), b.astype(np.float32)
    tdt = torch.float32 if dtype == "f4" else torch.float64
    prob = GroupLassoProblem(torch.from_numpy(A).to("cuda", tdt), torch.from_numpy(b).to("cuda", tdt))
    return prob, A.astype(np.float64), b.astype(np.float64), rng


@pytest.mark.parametrize("dtype", ["f8", "f4"])
@pytest.mark.parametrize("ncols", [1, 2, 4, 6])
@pytest.mark.parametrize("shape", SHAPES)
def test_forward_stream_matches_numpy(shape, ncols, dtype):
    import torch
    from glprox import _lib
    m, n = shape
    l = 2
    prob, A, b, rng = _problem(m, n, l, dtype)
    lib = _lib.load()
    X = rng.standard_normal((n, ncols))
    X[rng.random(n) < 0.5] = 0.0
    tdt = torch.float32 if dtype == "f4" else torch.float64
    Xd = torch.from_numpy(X).to("cuda", tdt)
    Xr = Xd.double().cpu().numpy()  # the operand the GPU actually saw
    AX = torch.empty((m, ncols), dtype=torch.float64, device="cuda")
    rr = torch.zeros(1, dtype=torch.float64, device="cuda")
    for rep in range(2):  # twice: the arrival counters must have been reset by the first launch
        _lib.check(lib.glp_fwd(prob.handle, Xd.data_ptr(), ncols, AX.data_ptr(), rr.data_ptr()))
        torch.cuda.synchronize()
        ref = A @ Xr
        scale = np.abs(A) @ np.abs(Xr) + 1e-300
        tol = 1e-13 if dtype == "f8" else 3e-6
        assert np.max(np.abs(AX.cpu().numpy() - ref) / scale) <= tol
        nb = min(ncols, l)
        rr_ref = np.sum((AX.cpu().numpy()[:, :nb] - b[:, :nb]) ** 2)
        assert abs(rr.item() - rr_ref) <= 1e-12 * rr_ref
    prob.close()


@pytest.mark.parametrize("dtype", ["f8", "f4"])
@pytest.mark.parametrize("l", [1, 2, 3])
@pytest.mark.parametrize("shape", SHAPES)
def test_transposed_stream_matches_numpy(shape, l, dtype):
    import torch
    from glprox import _lib
    m, n = shape
    prob, A, b, rng = _problem(m, n, l, dtype)
    lib = _lib.load()
    R = rng.standard_normal((m, l))
    Rd = torch.from_numpy(R).cuda()
    G = torch.empty((n, l), dtype=torch.float64, device="cuda")
    null = C.c_void_p()
    for rep in range(2):
        _lib.check(lib.glp_tpass_prox(prob.handle, None, Rd.data_ptr(), G.data_ptr(), null, null, null, null, null, null))
        torch.cuda.synchronize()
        Rr = R.astype(np.float32).astype(np.float64) if dtype == "f4" else R
        ref = A.T @ Rr
        scale = np.abs(A).T @ np.abs(Rr) + 1e-300
        tol = 1e-13 if dtype == "f8" else 3e-6
        assert np.max(np.abs(G.cpu().numpy() - ref) / scale) <= tol
    prob.close()


@pytest.mark.parametrize("dtype", ["f8", "f4"])
@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
@pytest.mark.parametrize("l", [1, 2, 3])
@pytest.mark.parametrize("shape", [(256, 512), (193, 391), (1000, 3000)])
def test_fused_prox_epilogue_matches_oracle(shape, l, algo, dtype):
    import torch
    from glprox import _lib
    m, n = shape
    prob, A, b, rng = _problem(m, n, l, dtype)
    lib = _lib.load()
    thres, mu, t, theta = 1e-3, 0.3, 7e-4, 2.0 / 7.0
    xt = rng.standard_normal((n, l))
    xt[np.abs(xt) < 0.4] = 0.0
    v = rng.standard_normal((n, l))
    R = 0.05 * rng.standard_normal((m, l))
    sp = _lib.GlpStepParams(algo=1 if algo == "FProxGD" else 0, theta=theta, t=t, mu=mu, thres=thres)
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    Rd, xtd, vd = d(R), d(xt), d(v)
    G, xn, xnt, vn = (torch.empty((n, l), dtype=torch.float64, device="cuda") for _ in range(4))
    sc = torch.zeros(4, dtype=torch.float64, device="cuda")
    _lib.check(lib.glp_tpass_prox(prob.handle, C.byref(sp), Rd.data_ptr(), G.data_ptr(), xtd.data_ptr(), vd.data_ptr(),
                                  xn.data_ptr(), xnt.data_ptr(), vn.data_ptr(), sc.data_ptr()))
    torch.cuda.synchronize()
    grad = G.cpu().numpy()  # feed the oracle the GPU's own gradient: isolates the epilogue arithmetic
    exn, exnt, evn, gd, dd, reg, mx = orc.prox_epilogue(algo, grad, xt, v, theta, t, mu, thres)
    assert np.array_equal(xn.cpu().numpy(), exn)        # same operations, same order: identical bits
    assert np.array_equal(xnt.cpu().numpy(), exnt)
    if algo == "FProxGD":
        assert np.array_equal(vn.cpu().numpy(), evn)
    got = sc.cpu().numpy()
    base = (1 - theta) * xt + theta * v if algo == "FProxGD" else xt
    dvec = exn - base if algo == "FProxGD" else (base - exn) / t
    for g_, e_, s_ in zip(got[:3], (gd, dd, reg), (np.sum(np.abs(grad * dvec)), dd, reg)):
        assert abs(g_ - e_) <= 1e-12 * s_ + 1e-300
    assert got[3] == mx
    prob.close()
