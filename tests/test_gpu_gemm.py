"""GPU parity tests of K5, the tcgen05/TMEM product kernel of the wide multi-RHS mode, through the C ABI
(glp_gemm_tf32 / glp_tf32_split).  Reference: fp64 torch matmul (the torch fp32 reference named in the
task is replaced by fp64 here because the compensated mode is *more* accurate than an fp32 GEMM)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _tf32_trunc(t):
    import torch
    return (t.view(torch.int32) & -8192).view(torch.float32)


def _run(M, K, N, passes, k_chunk=0, seed=0, p_scale=1.0, q_scale=1.0):
    import torch
    from glprox import _lib
    lib = _lib.load()
    g = torch.Generator(device="cuda").manual_seed(seed)
    ldp = (K + 7) // 8 * 8
    P = (torch.randn((M, ldp), generator=g, device="cuda", dtype=torch.float32) * p_scale)[:, :K]
    Q = (torch.randn((N, ldp), generator=g, device="cuda", dtype=torch.float32) * q_scale)[:, :K]
    Ph, Pl = torch.zeros((M, ldp), device="cuda"), torch.zeros((M, ldp), device="cuda")
    Qh, Ql = torch.zeros((N, ldp), device="cuda"), torch.zeros((N, ldp), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.glp_tf32_split(st, P.data_ptr(), M, K, ldp, Ph.data_ptr(), Pl.data_ptr(), ldp, 0))
    _lib.check(lib.glp_tf32_split(st, Q.data_ptr(), N, K, ldp, Qh.data_ptr(), Ql.data_ptr(), ldp, 0))
    ldd = (M + 7) // 8 * 8
    D = torch.full((N, ldd), float("nan"), device="cuda", dtype=torch.float64)
    if passes == 4:   # fp16 pairs are made from the fp32 operands themselves
        _lib.check(lib.glp_gemm_tf32(0, st, 4, M, K, N, P.data_ptr(), 0, ldp, Q.data_ptr(), 0, ldp, D.data_ptr(), ldd, k_chunk))
    else:
        _lib.check(lib.glp_gemm_tf32(0, st, passes, M, K, N, Ph.data_ptr(), Pl.data_ptr(), ldp, Qh.data_ptr(), Ql.data_ptr(),
                                     ldp, D.data_ptr(), ldd, k_chunk))
    torch.cuda.synchronize()
    return P, Q, Ph, Pl, Qh, Ql, D[:, :M].t()


def test_split_is_exact_and_hi_is_tf32():
    import torch
    P, Q, Ph, Pl, Qh, Ql, D = _run(128, 64, 16, 1)
    assert torch.equal(Ph[:, :64] + Pl[:, :64], P)                      # a = hi + lo exactly
    assert torch.equal(_tf32_trunc(Ph), Ph)                             # hi has 10 mantissa bits
    assert (Pl[:, :64].abs() <= P.abs() * 2.0 ** -11 + 1e-45).all()    # round to nearest


@pytest.mark.parametrize("shape", [(128, 32, 16), (128, 256, 64), (256, 1000, 128), (300, 4096, 256), (1000, 333, 200),
                                   (4096, 2048, 256), (70, 40, 4)])
def test_plain_tf32_matches_product_of_rounded_operands(shape):
    M, K, N = shape
    P, Q, Ph, Pl, Qh, Ql, D = _run(M, K, N, 1)
    ref = Ph[:, :K].double() @ Qh[:, :K].double().t()                   # what the tensor core is asked to compute
    scale = Ph[:, :K].double().abs() @ Qh[:, :K].double().abs().t() + 1e-300
    assert ((D - ref).abs() / scale).max().item() <= 2e-6               # fp32 accumulation error only
    full = P.double() @ Q.double().t()
    assert ((D - full).abs() / scale).max().item() <= 2e-3              # TF32 operand rounding


@pytest.mark.parametrize("k_chunk", [0, 8])
@pytest.mark.parametrize("shape", [(128, 256, 64), (300, 4096, 256), (1000, 333, 200), (2048, 8192, 128)])
def test_compensated_3xtf32_reaches_fp32_class_accuracy(shape, k_chunk):
    M, K, N = shape
    P, Q, Ph, Pl, Qh, Ql, D = _run(M, K, N, 3, k_chunk)
    full = P.double() @ Q.double().t()
    scale = P.double().abs() @ Q.double().abs().t() + 1e-300
    err = ((D - full).abs() / scale).max().item()
    assert err <= (5e-7 if k_chunk else 2e-5), err   # unchunked: the tensor core's fp32 accumulation spans all of K


@pytest.mark.parametrize("passes", [1, 3])
@pytest.mark.parametrize("shape", [(2048, 8192, 128), (16384, 4096, 256), (8192, 16384, 128), (300, 4096, 200), (128, 640, 16)])
def test_split_k_partials_sum_to_the_unsplit_product(shape, passes, monkeypatch):
    """k_chunk_blocks < 0 selects the split-K schedule (items = column groups x K splits x M tiles, partial
    products summed in split order).  The chunk sums are the same fp32 tensor-core sums either way, added
    in fp64, so the result matches the unsplit launch to fp64 rounding and is run-to-run identical."""
    import torch
    M, K, N = shape
    D0 = _run(M, K, N, passes, 8)[-1]
    monkeypatch.setenv("GLPROX_GEMM_SPLITS", "4")   # the heuristic only splits small shards; force it here
    D1 = _run(M, K, N, passes, -8)[-1]
    D2 = _run(M, K, N, passes, -8)[-1]
    assert torch.equal(D1, D2)
    assert not torch.isnan(D1).any()
    # passes = 3: the small correction terms (2^-11 of the result) are accumulated in fp32 over the whole K range
    # of an item, so their truncation bias ((K/8) 2^-24 of THEM) changes with the split: fp32-class differences
    err = ((D1 - D0).abs().max() / D0.abs().max()).item()
    assert err <= (1e-13 if passes == 1 else 2e-7), err


@pytest.mark.parametrize("shape", [(128, 256, 64), (300, 4096, 256), (1000, 333, 200), (2048, 8192, 128), (128, 64, 16)])
def test_bf16_corrections_stay_fp32_class(shape):
    """passes = 2: Ph.Qh on TF32 + (bf16(Ph).bf16(Ql) + bf16(Pl).bf16(Qh)) on kind::f16.  The corrections are 2^-11
    of the result and carry 8-bit mantissas: error ~2^-19/sqrt(K) of scale, still far below plain TF32 (5e-5)."""
    M, K, N = shape
    P, Q, Ph, Pl, Qh, Ql, D = _run(M, K, N, 2, 8)
    full = P.double() @ Q.double().t()
    scale = P.double().abs() @ Q.double().abs().t() + 1e-300
    err = ((D - full).abs() / scale).max().item()
    assert err <= 1e-6, err
    D3 = _run(M, K, N, 3, 8)[-1]
    assert ((D - D3).abs() / scale).max().item() <= 1e-6


def test_transposing_split():
    import torch
    from glprox import _lib
    lib = _lib.load()
    A = torch.randn((193, 391), device="cuda")
    hi, lo = torch.zeros((391, 200), device="cuda"), torch.zeros((391, 200), device="cuda")
    _lib.check(lib.glp_tf32_split(torch.cuda.current_stream().cuda_stream, A.data_ptr(), 193, 391, 391, hi.data_ptr(),
                                  lo.data_ptr(), 200, 1))
    torch.cuda.synchronize()
    assert torch.equal(hi[:, :193] + lo[:, :193], A.t())
    assert torch.equal(_tf32_trunc(hi), hi)


@pytest.mark.parametrize("scales", [(1.0, 1.0), (1e-4, 3e5), (7e6, 1e-9)])
@pytest.mark.parametrize("k_chunk", [4, -4])
@pytest.mark.parametrize("shape", [(128, 64, 16), (128, 256, 64), (300, 4096, 256), (1000, 333, 200), (2048, 8192, 128),
                                   (70, 40, 4)])
def test_fp16_pairs_reach_fp32_class_accuracy(shape, k_chunk, scales):
    """passes = 4: both operands as pairs of fp16 numbers of the power-of-two-scaled value (the same 11-bit
    significands as TF32 hi/lo at half the bytes), three kind::f16 products, the scale undone in the epilogue.
    Same bar as 3xTF32 with the fp32 accumulator flushed every 256 elements of K, whatever the operands' magnitude."""
    M, K, N = shape
    P, Q, Ph, Pl, Qh, Ql, D = _run(M, K, N, 4, k_chunk, p_scale=scales[0], q_scale=scales[1])
    full = P.double() @ Q.double().t()
    scale = P.double().abs() @ Q.double().abs().t() + 1e-300
    err = ((D - full).abs() / scale).max().item()
    assert err <= 5e-7, err


def test_fp16_pairs_keep_normwise_accuracy_over_a_wide_dynamic_range():
    """Entries 2^18 and more below the operand's maximum lose relative (not norm-wise) precision: the error stays
    at the 2^-22 level relative to |P||Q|."""
    import torch
    from glprox import _lib
    lib = _lib.load()
    M, K, N = 256, 2048, 32
    g = torch.Generator(device="cuda").manual_seed(5)
    P = torch.randn((M, K), generator=g, device="cuda") * torch.exp(12.0 * torch.randn((M, K), generator=g, device="cuda"))
    Q = torch.randn((N, K), generator=g, device="cuda") * torch.exp(12.0 * torch.randn((N, K), generator=g, device="cuda"))
    D = torch.full((N, M), float("nan"), device="cuda", dtype=torch.float64)
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.glp_gemm_tf32(0, st, 4, M, K, N, P.data_ptr(), 0, K, Q.data_ptr(), 0, K, D.data_ptr(), M, 4))
    torch.cuda.synchronize()
    full = P.double() @ Q.double().t()
    Pa, Qa = P.double().abs(), Q.double().abs()
    normwise = Qa.sum(dim=1)[:, None] * Pa.max() + Pa.sum(dim=1)[None, :] * Qa.max()      # [N][M]
    assert torch.isfinite(D).all()
    assert ((D - full.t()).abs() / normwise).max().item() <= 5e-7


@pytest.mark.parametrize("shape", [(16384, 8192, 128), (512, 65536, 128), (4096, 16384, 256), (1000, 3333, 200), (300, 700, 16),
                                   (65536, 2048, 64)])
def test_cta_pair_kernel_equals_the_one_cta_kernel(shape, monkeypatch):
    """K5 on CTA pairs (tcgen05.mma.cta_group::2, M = 256 over two SMs, stream-K over the 74 pairs; csrc/glp_gemm_pair.cuh)
    against the one-CTA kernel on the same fp16-pair operands: the tensor core sees the same 256-element chunks, the
    fp64 sums only differ in association (partial tiles of a stream-K range are added in pair order), and a rerun is
    bit-identical (fixed summation order).  Shapes: tiles shared by 2 pairs, by ~37 pairs, two column groups, ragged."""
    import torch
    M, K, N = shape
    monkeypatch.setenv("GLPROX_GEMM_PAIR", "0")
    D0 = _run(M, K, N, 4, 4, seed=3)[-1]
    monkeypatch.setenv("GLPROX_GEMM_PAIR", "1")
    P, Q, _, _, _, _, D1 = _run(M, K, N, 4, 4, seed=3)
    D2 = _run(M, K, N, 4, 4, seed=3)[-1]
    assert not torch.isnan(D1).any()
    assert torch.equal(D1, D2)
    scale = P.double().abs() @ Q.double().abs().t() + 1e-300
    assert ((D1 - D0).abs() / scale).max().item() <= 1e-7   # the fp32 correction accumulator spans a stream-K range (here up to 37 per tile), not a tile
    full = P.double() @ Q.double().t()
    assert ((D1 - full).abs() / scale).max().item() <= 5e-7
