"""glp_metrics_sums_noalloc (include/glprox.h): the allocation-free variant of the driver-metric reductions, called
through the C ABI and checked against the existing entry point glp_metrics_eval.  No counterpart in the reference.
Named test_zz_* so that pytest collects it last:
a failure here must not hide any other test.

The records of the reduction live in ONE file-scope device array shared by all calls, which keeps the previous call's
contents -- hence two calls in a row (the second on a smaller input, so stale records of the first would show), both
dtypes, and a call with no reference array.  torch is used by the TEST for device memory; the entry point itself
allocates nothing."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FIELDS = ("sumsq_x", "sumsq_ref", "sumsq_diff", "reg", "max_abs", "count")


def _existing(torch, _lib, lib, A, b, x, ref):
    """glp_metrics_eval through a handle over (A, b): the six raw fields + the finished ratios."""
    from glprox import GroupLassoProblem
    prob = GroupLassoProblem(A, b)
    met = _lib.GlpMetrics()
    with torch.cuda.device(A.device):
        _lib.check(lib.glp_metrics_eval(prob.handle, x.data_ptr(), ref.data_ptr() if ref is not None else None, 0.01, -1.0,
                                        C.byref(met)))
    prob.close()
    return np.array([getattr(met, f) for f in FIELDS]), met.sparsity, met.err_exact


def _noalloc(torch, _lib, lib, x, ref, out6, max_override=-1.0):
    n, l = x.shape
    dt = _lib.GLP_F32 if x.dtype == torch.float32 else _lib.GLP_F64
    st = torch.cuda.current_stream(x.device).cuda_stream
    _lib.check(lib.glp_metrics_sums_noalloc(x.device.index, C.c_void_p(st), dt, x.data_ptr(),
                                            ref.data_ptr() if ref is not None else None, n, l, float(max_override),
                                            out6.data_ptr()))
    torch.cuda.synchronize(x.device)
    return out6.cpu().numpy().copy()


@pytest.mark.parametrize("dtype", ["f64", "f32"])
def test_noalloc_metrics_match_the_existing_entry_point_on_two_calls_in_a_row(dtype):
    import torch
    from glprox import _lib
    lib = _lib.load()
    tdt = torch.float64 if dtype == "f64" else torch.float32
    g = torch.Generator(device="cpu").manual_seed(20261105)
    out6 = torch.full((6,), float("nan"), dtype=torch.float64, device="cuda")
    # first call: 200,000 row groups (592 records, the full scratch); second: 700 (3 records) -- stale records of the
    # first call would be folded into the second if the clear were missing
    for call, (m, n, l, with_ref) in enumerate([(64, 200_000, 2, True), (64, 700, 3, True), (32, 5_000, 1, False)]):
        A = torch.randn((m, n), generator=g, dtype=torch.float64).to(tdt).cuda()
        b = torch.randn((m, l), generator=g, dtype=torch.float64).to(tdt).cuda()
        x = torch.randn((n, l), generator=g, dtype=torch.float64)
        x[torch.rand((n,), generator=g) < 0.8] = 0.0          # a sparse iterate: the count is not trivially n l
        x = x.to(tdt).cuda()
        ref = (x + 1e-3 * torch.randn((n, l), generator=g, dtype=torch.float64).to(tdt).cuda()) if with_ref else None
        want, sparsity, err_exact = _existing(torch, _lib, lib, A, b, x, ref)
        free_before = torch.cuda.mem_get_info()[0]
        got = _noalloc(torch, _lib, lib, x, ref, out6)
        free_after = torch.cuda.mem_get_info()[0]
        # same kernels, same grid, same summation order: the bar of tests/test_gpu_metrics.py (1e-12) is met exactly
        assert np.array_equal(got, want), (call, got, want)
        assert got[5] / (n * l) == sparsity
        if with_ref:
            assert abs(np.sqrt(got[2]) / (1.0 + np.sqrt(got[0])) - err_exact) <= 1e-12 * err_exact
        else:
            assert got[1] == 0.0 and got[2] == 0.0
        if call > 0:   # kernels already loaded by the first call: device memory in use must not move
            assert free_after == free_before, (free_before, free_after)
        # against NumPy as well (fp64 accumulation of the stored values), the tolerance of the metrics tests
        xh = x.cpu().double().numpy()
        assert abs(got[0] - np.sum(xh * xh)) <= 1e-12 * np.sum(xh * xh)
        assert got[4] == np.max(np.abs(xh)) and got[5] == np.sum(np.abs(xh) > 1e-6 * np.max(np.abs(xh)))
    # the same input again after the others: the result does not depend on what the array held
    again = _noalloc(torch, _lib, lib, x, ref, out6)
    assert np.array_equal(again, got)
    # a cross-rank max handed in by the caller moves only the count
    big = _noalloc(torch, _lib, lib, x, ref, out6, max_override=1e9)
    assert big[5] == 0.0 and np.array_equal(big[:5], got[:5])


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch_and_bad_arguments():
    import torch
    from glprox import _lib
    lib = _lib.load()
    x = torch.zeros((8, 2), dtype=torch.float64, device="cuda")
    out6 = torch.zeros((6,), dtype=torch.float64, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    call = lambda n, l, xp=x.data_ptr(), op=out6.data_ptr(), dt=_lib.GLP_F64: lib.glp_metrics_sums_noalloc(
        0, st, dt, xp, None, n, l, -1.0, op)
    for n, l in (((1 << 24) + 1, 2), (8, 129)):          # GLP_NOALLOC_MAX_ROWS, GLP_NOALLOC_MAX_L
        assert call(n, l) == _lib.GLP_ERR_ARG
        msg = lib.glp_last_error().decode()
        assert "exceeds the fixed-scratch limit" in msg and "glp_metrics_eval" in msg
    assert call(0, 2) == _lib.GLP_ERR_ARG and call(8, 0) == _lib.GLP_ERR_ARG
    assert call(8, 2, xp=None) == _lib.GLP_ERR_ARG and call(8, 2, op=None) == _lib.GLP_ERR_ARG
    assert call(8, 2, dt=7) == _lib.GLP_ERR_ARG
    assert call(8, 2) == _lib.GLP_OK
    torch.cuda.synchronize()
    assert np.array_equal(out6.cpu().numpy(), np.zeros(6))
