"""CPU checks of the wide (tensor-core) mode's numerical design (tests/wide_emulator.py): the operand sets and the
sparse threshold correction, driven through the oracle's two-stream schedule, against the fp64 oracle.
The kernels themselves are tested on the GPU (tests/test_gpu_gemm.py, tests/test_gpu_wide.py)."""
import numpy as np
import pytest

import wide_emulator as we
from oracle import glprox_oracle as orc


def _case(m, n, l, seed, maxit):
    A, b, x0, u, mu = orc.gen_data(seed, m, n, l)
    A, b, x0 = (z.astype(np.float32) for z in (A, b, x0))
    opts = {"maxit": maxit, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    return A, b, x0, mu, opts


def test_fp16_scale_and_pairs():
    """max|a s| lands in [2^14, 2^15) (no fp16 overflow, v1 normal for all but tiny entries) and v0 + v1 carries
    22 significant bits of a s."""
    for mx in (1e-30, 3.3e-4, 0.5, 1.0, 4.7, 65504.0, 1e20):
        s = we.f16_scale_for(mx)
        assert 2.0 ** 14 <= mx * s < 2.0 ** 15 and np.log2(s) == np.round(np.log2(s))
    assert we.f16_scale_for(0.0) == 1.0 and we.f16_scale_for(float("inf")) == 1.0 and we.f16_scale_for(float("nan")) == 1.0
    g = np.random.default_rng(0)
    a = g.standard_normal(100000).astype(np.float32)
    s = we.f16_scale_for(np.max(np.abs(a)))
    v0, v1 = we.split_f16(a, s)
    assert np.all(np.isfinite(v0.astype(np.float32))) and np.all(np.isfinite(v1.astype(np.float32)))
    v = a.astype(np.float64) * s
    err = np.abs(v0.astype(np.float64) + v1.astype(np.float64) - v)
    big = np.abs(v) >= 2.0 ** -3          # v1 is a normal fp16 number there
    assert np.max(err[big] / np.abs(v[big])) <= 2.0 ** -21
    assert np.max(err[~big], initial=0.0) <= 2.0 ** -24


def test_tf32_split_is_exact_and_ten_bits():
    g = np.random.default_rng(1)
    a = (g.standard_normal(50000) * 10.0 ** g.integers(-6, 6, 50000)).astype(np.float32)
    hi, lo = we.split_tf32(a)
    assert np.array_equal(hi + lo, a)                                   # the split loses nothing
    assert np.all((hi.view(np.uint32) & 0x1FFF) == 0)                   # 13 low mantissa bits clear
    assert np.max(np.abs(lo) / np.abs(a)) <= 2.0 ** -11


@pytest.mark.parametrize("product", [we.product_f16x2, we.product_3xtf32])
def test_compensated_products_are_fp32_class(product):
    """Against the exact fp64 product of the same fp32 operands: errors of 2^-22-class relative to |P|.|Q|, i.e. what a
    careful fp32 GEMM delivers -- plain TF32 / fp16 operands alone sit at 2^-11."""
    g = np.random.default_rng(2)
    P = g.standard_normal((96, 2048)).astype(np.float32)
    Q = (g.standard_normal((2048, 24)) * 3.0).astype(np.float32)
    exact = P.astype(np.float64) @ Q.astype(np.float64)
    scale = np.abs(P).astype(np.float64) @ np.abs(Q).astype(np.float64)
    assert np.max(np.abs(product(P, Q) - exact) / scale) <= 2e-7
    p0, _ = we.split_f16(P, we.f16_scale_for(np.max(np.abs(P))))
    plain = (p0.astype(np.float64) / we.f16_scale_for(np.max(np.abs(P)))) @ Q.astype(np.float64)
    assert np.max(np.abs(plain - exact) / scale) > 1e-5                  # the corrections are what buys the accuracy


def test_threshold_correction_equals_the_direct_product():
    """A.thr(x) = A.x - A.(x - thr(x)) with the per-column ascending lists, to fp32 rounding of a term that is itself
    thres / max|x| of the product."""
    g = np.random.default_rng(3)
    m, n, l, thres = 64, 4096, 8, 2e-2
    A = g.standard_normal((m, n)).astype(np.float32)
    xn = g.standard_normal((n, l)).astype(np.float32)
    xnt = orc.hard_threshold(xn.astype(np.float64), thres).astype(np.float32)
    lists = we.threshold_lists(xn, xnt)
    assert sum(len(r) for r, _ in lists) > 0
    for rows, vals in lists:
        assert np.all(np.diff(rows) > 0) and np.all(np.abs(vals) < thres) and np.all(vals != 0)
    direct = A.astype(np.float64) @ xnt.astype(np.float64)
    fixed = A.astype(np.float64) @ xn.astype(np.float64) - we.threshold_correction(A, lists).astype(np.float64)
    assert np.max(np.abs(fixed - direct)) <= 1e-6 * np.max(np.abs(direct))


@pytest.mark.parametrize("algo", ["FProxGD", "ProxGD"])
@pytest.mark.parametrize("product", [we.product_f16x2, we.product_3xtf32])
def test_wide_schedule_meets_the_fp32_bar(product, algo):
    """The whole solve on the emulated wide arithmetic (fp32 operands, compensated products, sparse threshold
    correction) against the fp64 oracle on the fp32-rounded inputs: BASELINE.json's fp32 bar, 1e-5."""
    m, n, l = 128, 512, 16
    A, b, x0, mu, opts = _case(m, n, l, 7, 10)
    opts = dict(opts, thres=2e-2)        # makes the hard threshold bite in every iteration
    A64, b64, x064 = (z.astype(np.float64) for z in (A, b, x0))
    xr, kr, outr = orc.solve(algo, x064, A64, b64, mu, opts)
    fwd, bwd, stats = we.make_products(A, product)
    x, k, out = orc.solve_two_stream(algo, x064, A64, b64, mu, opts, matmul=fwd, rmatmul=bwd)
    assert k == kr == 30 and stats["entries"] > 0
    fh, fr = np.asarray(out["f_hist"]), np.asarray(outr["f_hist"])
    assert np.max(np.abs(fh - fr) / np.abs(fr)) <= 1e-5
    assert abs(out["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])
    # the correction route and the direct product over both operand halves agree far inside the bar
    fwd2, bwd2, _ = we.make_products(A, product, thrfix=False)
    x2, k2, out2 = orc.solve_two_stream(algo, x064, A64, b64, mu, opts, matmul=fwd2, rmatmul=bwd2)
    assert k2 == k and abs(out2["fval"] - out["fval"]) <= 2e-6 * abs(out["fval"])
