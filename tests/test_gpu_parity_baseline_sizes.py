"""Parity with the oracle AT THE BASELINE.json SIZES (configs[2], [3], [4]).

The GPU solver and the oracle (`oracle.solve`, the restatement of code/gl_ProxGD_primal.py:49-146 and
code/gl_FProxGD_primal.py:50-161, fp64 arithmetic on the fp32-rounded inputs: SURVEY.md 8(c) "fp32 oracle") run
the same iterations on the same seeded problem; the problem is generated twice, on the device by the product's
counter-based generator and on the host by its CPU restatement (oracle/gen.py), whose bit-equality is itself
asserted here.  Bar (north_star, fp32 mode): every f_hist entry and fval within 1e-5 relative, identical
iteration count, identical rule-based support.
"""
import os

import numpy as np
import pytest

from oracle import gen
from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu
BAR = 1e-5     # north_star: fp32 mode, relative objective difference


def _host_gb():
    for line in open("/proc/meminfo"):
        if line.startswith("MemAvailable:"):
            return int(line.split()[1]) / 1e6
    return 0.0


def _support(x):
    ax = np.abs(x)
    return np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1))


def _compare(algo, x, k, out, xr, kr, outr, bar=BAR):
    assert k == kr, (algo, k, kr)
    fh, ref = np.asarray(out["f_hist"]), np.asarray(outr["f_hist"])
    worst = float(np.max(np.abs(fh - ref) / np.abs(ref)))
    rel = abs(out["fval"] - outr["fval"]) / abs(outr["fval"])
    print("%s: k=%d max rel f_hist diff %.2e, fval rel diff %.2e (gpu %.9e, oracle %.9e)" % (
        algo, k, worst, rel, out["fval"], outr["fval"]))
    assert worst <= bar, (algo, worst)
    assert rel <= bar, (algo, rel)
    assert np.array_equal(_support(x), _support(xr))


def test_device_generator_matches_cpu_restatement():
    """glp_randn (csrc/glp_misc.cuh) == oracle/glprox_gen.c == oracle/gen.py (NumPy), bit for bit, fp32 and fp64,
    at an offset block of the logical matrix."""
    import torch
    from glprox import _lib
    lib = _lib.load()
    st = torch.cuda.current_stream().cuda_stream
    for tdt, gid, ndt in ((torch.float32, _lib.GLP_F32, np.float32), (torch.float64, _lib.GLP_F64, np.float64)):
        d = torch.empty((1024, 4096), dtype=tdt, device="cuda")
        _lib.check(lib.glp_randn(d.data_ptr(), gid, 1024, 4096, 4096, 97006855, 31744, 126976, st))
        torch.cuda.synchronize()
        c = gen.randn_block(97006855, 1024, 4096, 31744, 126976, ndt)
        assert np.array_equal(d.cpu().numpy(), c)
        assert np.array_equal(c[:64, :512], gen.randn_block_numpy(97006855, 64, 512, 31744, 126976, ndt))
    assert abs(float(c.mean())) < 3e-3 and abs(float(c.std()) - 1.0) < 3e-3


def test_device_problem_matches_cpu_problem():
    """dataset.gen_data_device (A, b = A u through K1, x0) == oracle.gen.gen_problem on the host."""
    from glprox import dataset
    A, b, x0, u, mu, _ = dataset.gen_data_device(2048, 8192, 2, 5, "f32")
    Ah, bh, x0h, uh, muh = gen.gen_problem(2048, 8192, 2, 5, np.float32)
    assert np.array_equal(A.cpu().numpy(), Ah) and np.array_equal(x0.cpu().numpy(), x0h) and np.array_equal(u.cpu().numpy(), uh)
    # b: the same fp64 sum in the same order on both sides (glp_rhs_exact / glp_oracle_matmul), rounded to fp32
    assert np.array_equal(b.cpu().numpy(), bh)


@pytest.mark.parametrize("algo", ["FProxGD", "ProxGD"])
def test_c3_parity_with_oracle(algo):
    """BASELINE configs[2]: m=8192 n=32768 l=2 fp32, maxit=10 per stage (30 iterations)."""
    import torch
    from glprox import GroupLassoProblem
    m, n, l = 8192, 32768, 2
    A, b, x0, u, mu = gen.gen_problem(m, n, l, 97006855, np.float32)
    opts = {"maxit": 10, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x, k, out = prob.solve(algo, torch.from_numpy(x0).cuda(), mu, opts)
    prob.close()
    xr, kr, outr = orc.solve(algo, x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
    _compare(algo, x.cpu().numpy(), k, out, xr, kr, outr)


@pytest.mark.parametrize("algo", ["FProxGD", "ProxGD"])
def test_c4_parity_with_oracle(algo):
    """BASELINE configs[3], the headline workload: m=32768 n=131072 l=2 fp32, maxit=2 per stage (6 iterations).
    GPU inputs come from the device generator, oracle inputs from its CPU restatement (17 GB on the host; the
    fp64 arithmetic converts one row block at a time, oracle.gen.BlockedF64)."""
    import torch
    from glprox import GroupLassoProblem, dataset
    if _host_gb() < 40:
        pytest.skip("needs 40 GB of host RAM for the oracle's copy of A (have %.0f GB)" % _host_gb())
    if torch.cuda.mem_get_info()[0] < 40e9:
        pytest.skip("needs 40 GB of free HBM")
    m, n, l = 32768, 131072, 2
    opts = {"maxit": 2, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    A, b, x0, u, mu, _ = dataset.gen_data_device(m, n, l, 97006855, "f32")
    prob = GroupLassoProblem(A, b)
    x, k, out = prob.solve(algo, x0, mu, opts)
    prob.close()
    xh, bh, x0h = x.cpu().numpy(), b.cpu().numpy(), x0.cpu().numpy()
    probe = A[m - 256:, n - 1024:].cpu().numpy()
    del A, prob
    torch.cuda.empty_cache()
    Ah = gen.randn_block(97006855, m, n, 0, 0, np.float32)
    assert np.array_equal(Ah[m - 256:, n - 1024:], probe)     # same matrix on both sides
    xr, kr, outr = orc.solve(algo, x0h.astype(np.float64), gen.BlockedF64(Ah), bh.astype(np.float64), mu, opts)
    _compare(algo, xh, k, out, xr, kr, outr)


def test_c5_wide_parity_with_oracle():
    """BASELINE configs[4]: m=16384 n=65536 l=128 fp32, default tensor-core mode (fp16 pairs on tcgen05 + sparse
    threshold correction), maxit=1 per stage (3 iterations)."""
    import torch
    from glprox import GroupLassoProblem, dataset
    if _host_gb() < 30:
        pytest.skip("needs 30 GB of host RAM (have %.0f GB)" % _host_gb())
    m, n, l = 16384, 65536, 128
    opts = {"maxit": 1, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    A, b, x0, u, mu, _ = dataset.gen_data_device(m, n, l, 97006855, "f32")
    prob = GroupLassoProblem(A, b)
    x, k, out = prob.solve("FProxGD", x0, mu, opts)
    prob.close()
    Ah, bh, x0h, xh = A.cpu().numpy(), b.cpu().numpy(), x0.cpu().numpy(), x.cpu().numpy()
    del A, prob
    torch.cuda.empty_cache()
    xr, kr, outr = orc.solve("FProxGD", x0h.astype(np.float64), gen.BlockedF64(Ah, block_rows=512), bh.astype(np.float64), mu, opts)
    _compare("FProxGD wide l=128", xh, k, out, xr, kr, outr)


def test_c4_fp32_line_search_is_not_driven_by_rounding_noise():
    """BASELINE configs[3] in fp32, 50 iterations per stage: late in a stage |A x| ~ 114 while the residual entries are
    ~0.15, and the margin of ProxGD's sufficient-decrease test (gl_ProxGD_primal.py:91-92) falls to ~1e-6 of g.  With an
    fp32 second accumulation level in the forward stream the noise of ||A x - b||^2 exceeded it and the line search
    rejected at random (385 forward streams for 150 iterations, profiles/r02_ab_k1_acc64.log); the fp64 run, like the
    reference's arithmetic, accepts every first trial.  Pins: one forward stream per iteration for both solvers, and an
    fp32 objective within 1e-6 of the fp64 run's after 150 iterations."""
    import torch
    from glprox import GroupLassoProblem, dataset
    m, n, l = 32768, 131072, 2
    free = torch.cuda.mem_get_info()[0] / 1e9
    if free < 60:
        pytest.skip("needs ~52 GB of HBM for the fp32 and the fp64 copy of A (free: %.0f GB)" % free)
    a0 = float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)
    fvals = {}
    for dt in ("f32", "f64"):
        A, b, x0, u, mu, _ = dataset.gen_data_device(m, n, l, 7, dt)
        prob = GroupLassoProblem(A, b)
        for algo in ("ProxGD", "FProxGD"):
            x, k, out = prob.solve(algo, x0, mu, {"maxit": 50, "alpha0": a0, "backend": "stream"})
            assert k == 150
            assert out["glprox"]["n_fwd"] == k + 1, (dt, algo, out["glprox"]["n_fwd"])
            fvals[dt, algo] = float(out["fval"])
        prob.close()
        del A, b, x0, u, prob
        torch.cuda.empty_cache()
    for algo in ("ProxGD", "FProxGD"):
        rel = abs(fvals["f32", algo] - fvals["f64", algo]) / abs(fvals["f64", algo])
        print("%s after 150 iterations: fp32 %.9e fp64 %.9e rel %.2e" % (algo, fvals["f32", algo], fvals["f64", algo], rel))
        assert rel <= 1e-6, (algo, rel)
