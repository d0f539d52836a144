"""CPU emulation of the wide (tensor-core) mode's ARITHMETIC for host-side tests (no GPU, no CUDA).

What the K5 path of csrc/glp_gemm.cuh + csrc/glp_thrfix.cuh computes, restated in NumPy so that the
numerical design can be checked against the fp64 oracle on the CPU box:

* fp16-pair operand set (PASSES = 4): an fp32 operand a is scaled by a power of two s taken from its largest
  magnitude (max|a s| in [2^14, 2^15)) and held as v0 = fp16(a s), v1 = fp16(a s - v0); the product is
  P0.Q0 + (P0.Q1 + P1.Q0), the main term accumulated in fp32 over chunks of 256 elements of K and the
  chunks in fp64, the corrections in fp32 over the whole K, the result divided by the two scales;
* 3xTF32 operand set (PASSES = 3): hi = a rounded to 10 mantissa bits, lo = a - hi, Ph.Qh + (Ph.Ql + Pl.Qh);
* A.thr(x) from A.x by the sparse correction: E = x - thr(x) listed per column in ascending row order,
  (A.E)(:, c) accumulated with one fp32 FMA per entry, subtracted from the fp64 product.

Not emulated: the tensor core's truncating fp32 accumulation inside a chunk (BLAS fp32 summation stands in
for it) and the summation order of the MMA tree.  The emulator therefore bounds what the SCHEME can deliver;
the kernels themselves are held to the same bars on the GPU (tests/test_gpu_gemm.py, tests/test_gpu_wide.py).
Test infrastructure only, like oracle/.
"""
import math

import numpy as np

K_CHUNK = 256   # elements of K per fp32 accumulation (WIDE_K_CHUNK x 32 in csrc/glprox.cu)


def f16_scale_for(mx):
    """csrc/glp_gemm.cuh: f16_scale_for -- power of two that brings mx into [2^14, 2^15); 1 for 0 / non-finite."""
    mx = float(mx)
    if not (mx > 0.0) or not (mx < 3.0e38):
        return 1.0
    _, e = math.frexp(mx)            # mx = f * 2^e, f in [0.5, 1)
    k = min(126, max(-126, 15 - e))
    return math.ldexp(1.0, k)


def split_f16(a32, s):
    """csrc/glp_gemm.cuh: split_f16_one, element-wise."""
    v = (a32.astype(np.float32) * np.float32(s)).astype(np.float32)
    v0 = v.astype(np.float16)
    v1 = (v - v0.astype(np.float32)).astype(np.float16)
    return v0, v1


def split_tf32(a32):
    """csrc/glp_gemm.cuh: split_tf32_one -- round to nearest even on bit 13, lo = a - hi."""
    u = a32.astype(np.float32).view(np.uint32).astype(np.uint64)
    u = ((u + 0x00000FFF + ((u >> 13) & 1)) & 0xFFFFE000).astype(np.uint32)
    hi = u.view(np.float32)
    bad = ~(np.abs(hi) <= np.float32(3.0e38))
    hi = np.where(bad, a32.astype(np.float32), hi)
    return hi, (a32.astype(np.float32) - hi).astype(np.float32)


def _chunked(P, Q, chunk=K_CHUNK):
    """sum over chunks of K of fp32(P[:, chunk] @ Q[chunk, :]), the chunks summed in fp64."""
    out = np.zeros((P.shape[0], Q.shape[1]), dtype=np.float64)
    for k0 in range(0, P.shape[1], chunk):
        out += (P[:, k0:k0 + chunk].astype(np.float32) @ Q[k0:k0 + chunk].astype(np.float32)).astype(np.float64)
    return out


def product_f16x2(P32, Q32):
    """P [M x K] . Q [K x N] with both operands as fp16 pairs (the default wide operand set)."""
    sp, sq = f16_scale_for(np.max(np.abs(P32), initial=0.0)), f16_scale_for(np.max(np.abs(Q32), initial=0.0))
    p0, p1 = split_f16(P32, sp)
    q0, q1 = split_f16(Q32, sq)
    main = _chunked(p0, q0)
    corr = (p0.astype(np.float32) @ q1.astype(np.float32) + p1.astype(np.float32) @ q0.astype(np.float32)).astype(np.float64)
    return (main + corr) / (sp * sq)


def product_3xtf32(P32, Q32):
    ph, pl = split_tf32(P32)
    qh, ql = split_tf32(Q32)
    return _chunked(ph, qh) + (ph @ ql + pl @ qh).astype(np.float64)


def threshold_lists(xn32, xnt32):
    """csrc/glp_thrfix.cuh: glp_thr_list_kernel -- per column, the rows where the hard threshold changed the
    iterate (ascending) and xn - xnt there (exact: xnt is xn or 0)."""
    d = xn32.astype(np.float32) - xnt32.astype(np.float32)
    return [(np.flatnonzero(d[:, c]), d[np.flatnonzero(d[:, c]), c]) for c in range(d.shape[1])]


def threshold_correction(A32, lists):
    """csrc/glp_thrfix.cuh: glp_thr_fix_kernel -- (A.E)(:, c) with one fp32 FMA per listed entry, in list order.
    (fp32 multiply-add of exactly representable factors; NumPy has no fused fp32 FMA, the double rounding stays far
    below the 2^-22 of the product it corrects.)"""
    out = np.zeros((A32.shape[0], len(lists)), dtype=np.float32)
    for c, (rows, vals) in enumerate(lists):
        acc = np.zeros(A32.shape[0], dtype=np.float32)
        for j, v in zip(rows, vals):
            acc = (acc + A32[:, j] * np.float32(v)).astype(np.float32)
        out[:, c] = acc
    return out


def make_products(A32, product=product_f16x2, thrfix=True):
    """(matmul, rmatmul) for oracle.solve_two_stream that follow the wide mode: the forward operand [xn | xnt]
    and the residual are rounded to fp32 (they are fp32 device arrays), the products run on the emulated operand
    set, and A.xnt comes from A.xn by the sparse correction."""
    A32 = np.ascontiguousarray(A32, dtype=np.float32)
    At32 = np.ascontiguousarray(A32.T)
    stats = {"entries": 0, "fwd": 0}

    def matmul(X):
        l = X.shape[1] // 2
        xn, xnt = X[:, :l].astype(np.float32), X[:, l:].astype(np.float32)
        stats["fwd"] += 1
        if not thrfix:
            return np.concatenate([product(A32, xn), product(A32, xnt)], axis=1)
        axn = product(A32, xn)
        lists = threshold_lists(xn, xnt)
        stats["entries"] += sum(len(r) for r, _ in lists)
        return np.concatenate([axn, axn - threshold_correction(A32, lists).astype(np.float64)], axis=1)

    def rmatmul(R):
        return product(At32, R.astype(np.float32))

    return matmul, rmatmul, stats
