"""GPU parity tests of the wide multi-RHS mode (l >= 16, fp32): both products on tcgen05 tensor cores
(K5).  Bars: fp16 pairs (default), 3xTF32 and TF32 + bf16 corrections -- the fp32 bar of BASELINE.json, relative
objective difference <= 1e-5 versus the fp64 oracle on the fp32-rounded inputs; plain TF32 (opt-in, approximate) -- 10-bit operand
mantissas perturb every product by ~1e-3, which the 45 iterations amplify: the trajectory is held to 1e-1
and the final objective to 5e-2 (stated here because the north star gives no TF32 tolerance)."""
import numpy as np
import pytest

from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu


def _case(m, n, l, seed, maxit):
    A, b, x0, u, mu = orc.gen_data(seed, m, n, l)
    A, b, x0 = (z.astype(np.float32) for z in (A, b, x0))
    opts = {"maxit": maxit, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    return A, b, x0, mu, opts


@pytest.mark.parametrize("algo", ["FProxGD", "ProxGD"])
@pytest.mark.parametrize("shape", [(256, 1024, 16), (300, 1100, 24), (512, 2048, 128), (1024, 4096, 64)])
def test_wide_mode_matches_fp64_oracle(shape, algo):
    from gl_FProxGD_primal import gl_FProxGD_primal
    from gl_ProxGD_primal import gl_ProxGD_primal
    fn = gl_FProxGD_primal if algo == "FProxGD" else gl_ProxGD_primal
    m, n, l = shape
    A, b, x0, mu, opts = _case(m, n, l, 7, 15)
    xr, kr, outr = orc.solve(algo, x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
    ref_h = np.asarray(outr["f_hist"])
    for mode, tol, tol_hist in (("3xtf32", 1e-5, 1e-5), ("tf32+bf16", 1e-5, 1e-5), ("f16x2", 1e-5, 1e-5), ("tf32", 5e-2, 1e-1),
                                ("off", 1e-5, 1e-5)):
        x, k, out = fn(x0, A, b, mu, dict(opts, tensor_cores=mode))
        assert k == kr == 45 and x.dtype == np.float32
        fh = np.asarray(out["f_hist"])
        assert np.max(np.abs(fh - ref_h) / np.abs(ref_h)) <= tol_hist, (mode, np.max(np.abs(fh - ref_h) / np.abs(ref_h)))
        assert abs(out["fval"] - outr["fval"]) <= tol * abs(outr["fval"])
        f_true = orc.objective(A.astype(np.float64), b.astype(np.float64), x.astype(np.float64), mu)
        assert abs(f_true - outr["fval"]) <= tol * abs(outr["fval"])


def test_wide_mode_streams_A_twice_per_iteration():
    import torch
    from glprox import GroupLassoProblem
    m, n, l = 512, 2048, 32
    A, b, x0, mu, opts = _case(m, n, l, 3, 10)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x, k, out = prob.solve("FProxGD", torch.from_numpy(x0).cuda(), mu, opts)
    x2, k2, out2 = prob.solve("FProxGD", torch.from_numpy(x0).cuda(), mu, dict(opts, tensor_cores="off"))
    assert k == k2 == 30
    # CUDA-core path: 2l/4 forward + l/2 transposed launches per iteration; tensor path: one of each
    assert out["glprox"]["n_launches"] < out2["glprox"]["n_launches"] / 2
    assert abs(out["fval"] - out2["fval"]) <= 1e-5 * abs(out2["fval"])
    prob.close()


@pytest.mark.parametrize("algo", ["FProxGD", "ProxGD"])
@pytest.mark.parametrize("G,shape", [(2, (256, 1024, 16)), (4, (300, 2048, 24)), (8, (512, 4096, 128)), (3, (1024, 4096, 64))])
def test_wide_mode_sharded_matches_single_handle_and_oracle(G, shape, algo):
    """Wide mode over G column shards on ONE device (phase-ordered launches): tcgen05 products per shard,
    reduce-scatter + all-gather of the m x 2l partial products through the peer mailboxes
    (csrc/glp_wide_comm.cuh).  Same bar as the single-GPU wide mode (1e-5 vs the fp64 oracle)."""
    import torch
    from glprox import GroupLassoProblem, shard_bounds, solve_group
    m, n, l = shape
    A, b, x0, mu, opts = _case(m, n, l, 7, 15)
    Ad, bd = torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda()
    bounds = shard_bounds(n, G, 128)
    probs = [GroupLassoProblem(Ad[:, c0:c1], bd) for c0, c1 in bounds]
    ptrs = [p.export_mailbox(G)[1] for p in probs]
    for r, p in enumerate(probs):
        p.attach_peers(r, G, n, ptrs, kind=1)
    x0s = [torch.from_numpy(x0[c0:c1].copy()).cuda() for c0, c1 in bounds]
    single = GroupLassoProblem(Ad, bd)
    x1, k1, out1 = single.solve(algo, torch.from_numpy(x0).cuda(), mu, opts)
    xr, kr, outr = orc.solve(algo, x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
    for rep in range(2):   # the mailboxes and slab parities are reused by the second solve
        xs, k, out = solve_group(probs, algo, x0s, mu, opts)
        assert k == k1 == kr == 45
        fh, f1, fr = (np.asarray(o["f_hist"]) for o in (out, out1, outr))
        assert np.max(np.abs(fh - fr) / np.abs(fr)) <= 1e-5
        assert np.max(np.abs(fh - f1) / np.abs(f1)) <= 1e-5
        assert abs(out["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])
        x = torch.cat(xs, 0).cpu().numpy().astype(np.float64)
        f_true = orc.objective(A.astype(np.float64), b.astype(np.float64), x, mu)
        assert abs(f_true - outr["fval"]) <= 1e-5 * abs(outr["fval"])
        pr = out["glprox"]["per_rank"]   # every rank took bit-identical decisions
        assert all(r["iters"] == k and r["fval"] == pr[0]["fval"] and r["n_fwd"] == pr[0]["n_fwd"] for r in pr)
        # 2 product launches (one gated off on the device) + entry lists / correction + 3 exchange kernels per
        # forward instead of 2l/4 CUDA-core streams
        assert out["glprox"]["n_launches"] < 45 * 40
    single.close()
    for p in probs:
        p.close()


def _solve_with_env(monkeypatch, env, algo, A, b, x0, mu, opts):
    """One solve on a fresh handle (the wide operand set and the threshold-correction switches are read when
    the handle first goes wide)."""
    import torch
    from glprox import GroupLassoProblem
    for k_, v_ in env.items():
        monkeypatch.setenv(k_, v_)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x, k, out = prob.solve(algo, torch.from_numpy(x0).cuda(), mu, opts)
    prob.close()
    for k_ in env:
        monkeypatch.delenv(k_)
    return x.cpu().numpy(), k, out


@pytest.mark.parametrize("algo", ["FProxGD", "ProxGD"])
@pytest.mark.parametrize("mode", ["3xtf32", "tf32+bf16", "f16x2"])
@pytest.mark.parametrize("shape", [(256, 1024, 16), (300, 1100, 24), (512, 2048, 128)])
def test_wide_threshold_correction_equals_direct_product(monkeypatch, shape, algo, mode):
    """A.thr(x) = A.x - A.(x - thr(x)) (csrc/glp_thrfix.cuh) against the direct product over both operand halves.
    thres = 2e-2 makes the hard threshold bite in every iteration (x0 ~ N(0,1): ~1.6 % of the entries), so the
    per-column entry lists, their overflow fallback (cap = 0: device-gated second product, bit-identical to the
    direct product) and the default capacity (which some columns exceed at this thres) are all exercised."""
    m, n, l = shape
    A, b, x0, mu, opts = _case(m, n, l, 11, 8)
    x0[5, 3] = 3e-3   # a row that is certainly hit
    opts = dict(opts, thres=2e-2, tensor_cores=mode)
    xd, kd, outd = _solve_with_env(monkeypatch, {"GLPROX_WIDE_THRFIX": "0"}, algo, A, b, x0, mu, opts)
    xg, kg, outg = _solve_with_env(monkeypatch, {"GLPROX_WIDE_THRFIX_CAP": "0"}, algo, A, b, x0, mu, opts)
    xl, kl, outl = _solve_with_env(monkeypatch, {"GLPROX_WIDE_THRFIX_CAP": str(n)}, algo, A, b, x0, mu, opts)
    xa, ka, outa = _solve_with_env(monkeypatch, {}, algo, A, b, x0, mu, opts)   # default cap: mixes both
    assert kd == kg == kl == ka
    fd, fg, fl, fa = (np.asarray(o["f_hist"]) for o in (outd, outg, outl, outa))
    assert np.array_equal(fd, fg) and np.array_equal(xd, xg)          # gated product = direct product, bit for bit
    # the correction is exact in fp64 where the direct product rounds A.E through the TF32 splits; the objective
    # falls by 5 orders of magnitude over the run (cancellation in r = Ax - b), so the two agree to ~1e-7
    for f, x in ((fl, xl), (fa, xa)):
        assert np.max(np.abs(f - fd) / np.abs(fd)) <= 2e-6, np.max(np.abs(f - fd) / np.abs(fd))
        assert np.max(np.abs(x - xd)) <= 1e-5 * max(1.0, np.max(np.abs(xd)))
    xr, kr, outr = orc.solve(algo, x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu,
                             {k_: v_ for k_, v_ in opts.items() if k_ != "tensor_cores"})
    assert kl == kr
    assert abs(outl["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])
    # the correction replaces half of the forward product: fewer tensor-core column groups, a few small launches
    assert outl["glprox"]["n_launches"] >= outd["glprox"]["n_launches"]


@pytest.mark.parametrize("shape", [(256, 1024, 16), (512, 2048, 128)])
def test_split_operand_rings_compute_the_same_bits(monkeypatch, shape):
    """GLPROX_GEMM_RINGS=1 (opt-in staging of the fp16-pair product: separate P / Q rings, csrc/glp_gemm.cuh) changes
    how the operand tiles reach shared memory, not the sequence of tensor-core instructions: same history, bit for bit."""
    m, n, l = shape
    A, b, x0, mu, opts = _case(m, n, l, 5, 8)
    x0_, k0, out0 = _solve_with_env(monkeypatch, {"GLPROX_GEMM_RINGS": "0"}, "FProxGD", A, b, x0, mu, opts)
    x1_, k1, out1 = _solve_with_env(monkeypatch, {"GLPROX_GEMM_RINGS": "1"}, "FProxGD", A, b, x0, mu, opts)
    assert k0 == k1 == 24
    assert np.array_equal(np.asarray(out0["f_hist"]), np.asarray(out1["f_hist"])) and np.array_equal(x0_, x1_)
    xr, kr, outr = orc.solve("FProxGD", x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
    assert abs(out1["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])



def test_auto_mode_falls_back_to_3xtf32_on_badly_scaled_columns():
    """the fp16-pair set keeps ~2^-40 of max|A| in absolute terms; when whole columns of A are scaled
    down by more than 2^18 (un-normalised features) and meet large x entries, that is below fp32 accuracy.  'auto'
    measures the range of the column / row maxima and takes the 3xTF32 set instead; an explicit 'f16x2' is refused."""
    import torch
    from glprox import GroupLassoProblem, _lib
    m, n, l = 256, 1024, 16
    A, b, x0, mu, opts = _case(m, n, l, 11, 5)
    scale = np.ones(n, dtype=np.float32)
    scale[::7] = 1e-7                                   # every 7th feature column tiny
    A = A * scale[None, :]
    x0 = (x0 / np.maximum(scale, 1e-3)[:, None]).astype(np.float32)   # and its coefficients large
    xr, kr, outr = orc.solve("FProxGD", x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x, k, out = prob.solve("FProxGD", torch.from_numpy(x0).cuda(), mu, opts)          # auto
    assert k == kr and abs(out["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])
    fh, ref = np.asarray(out["f_hist"]), np.asarray(outr["f_hist"])
    assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-5
    prob.close()
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    with pytest.raises(_lib.GlproxError):
        prob.solve("FProxGD", torch.from_numpy(x0).cuda(), mu, dict(opts, tensor_cores="f16x2"))
    prob.close()


def test_wide_default_mode_over_a_long_run_with_the_default_stop_rule_armed():
    """a wide solve with the reference's default stop rule armed (ftol 1e-6, stable_len 70) and a long
    budget, against the fp64 oracle.  The late FProxGD iterates oscillate at the 1e-5 level even in fp64, so an fp32
    trajectory decorrelates there (as NumPy fp32 does from fp64): the first 100 iterations are held
    to the fp32 bar, the whole history and the stage ends to the decorrelation level."""
    import torch
    from glprox import GroupLassoProblem
    m, n, l = 128, 512, 16
    A, b, x0, mu, _ = _case(m, n, l, 5, 0)
    opts = {"alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2), "maxit": 400}
    xr, kr, outr = orc.solve("FProxGD", x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x, k, out = prob.solve("FProxGD", torch.from_numpy(x0).cuda(), mu, opts)
    prob.close()
    assert abs(k - kr) <= max(3, kr // 50)
    kk = min(k, kr)
    fh, ref = np.asarray(out["f_hist"])[:kk], np.asarray(outr["f_hist"])[:kk]
    rel = np.abs(fh - ref) / np.abs(ref)
    assert np.max(rel[:100]) <= 1e-5 and np.max(rel) <= 2e-3
    assert abs(out["fval"] - outr["fval"]) <= 2e-3 * abs(outr["fval"])
