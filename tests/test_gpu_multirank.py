"""GPU tests of the sharded (multi-rank) path on ONE device: G column shards of A, each behind its
own handle and peer mailbox, driven phase by phase on a single stream (glp_group_solve).  The
kernels, the push/flag exchange and the gather are the ones a G-GPU job runs; only the transport
(NVLink vs. the local HBM) differs.  Compared against the golden vectors of the unmodified
reference and against the single-handle solve."""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu


def _shards(A, b, G, dtype):
    import torch
    from glprox import GroupLassoProblem, shard_bounds
    tdt = torch.float32 if dtype == "f4" else torch.float64
    Ad = torch.from_numpy(np.ascontiguousarray(A)).to("cuda", tdt)
    bd = torch.from_numpy(np.ascontiguousarray(b)).to("cuda", tdt)
    bounds = shard_bounds(A.shape[1], G, align=128 if dtype == "f4" else 64)
    probs = [GroupLassoProblem(Ad[:, c0:c1], bd) for c0, c1 in bounds]
    ptrs = [p.export_mailbox(G)[1] for p in probs]
    for r, p in enumerate(probs):
        p.attach_peers(r, G, A.shape[1], ptrs, kind=1)
    return probs, bounds, Ad


@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
@pytest.mark.parametrize("G", [2, 3, 8])
@pytest.mark.parametrize("name", ["fixture_maxit40", "harsh_ls", "l3", "fixture"])
def test_sharded_fp64_matches_reference(name, G, algo):
    import torch
    from glprox import solve_group
    if name == "fixture" and (G != 8 or algo != "FProxGD"):
        pytest.skip("full-length solve once")
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    if A.shape[1] < 64 * G:
        pytest.skip("too few column blocks")
    probs, bounds, keep = _shards(A, b, G, "f8")
    x0s = [torch.from_numpy(x0[c0:c1].copy()).cuda() for c0, c1 in bounds]
    for rep in range(2):  # the mailboxes are reused by the second solve (epoch barrier)
        xs, k, out = solve_group(probs, algo, x0s, mu, dict(rec["opts"]))
        x = np.concatenate([t.cpu().numpy() for t in xs], axis=0)
        assert k == int(rec[algo + "_k"])
        fh, ref = np.asarray(out["f_hist"]), rec[algo + "_f_hist"]
        assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10
        assert abs(out["fval"] - rec[algo + "_fval"]) <= 1e-10 * abs(rec[algo + "_fval"])
        assert np.max(np.abs(x - rec[algo + "_x"])) <= 1e-8
        ax = np.abs(x)
        assert np.array_equal(np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1)), rec[algo + "_support_rule"])
        # every rank reached bit-identical scalars (same additions in the same order everywhere)
        pr = out["glprox"]["per_rank"]
        assert all(r["iters"] == k and r["fval"] == pr[0]["fval"] and r["n_fwd"] == pr[0]["n_fwd"] for r in pr)
    for p in probs:
        p.close()


@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
def test_sharded_fp32_matches_single_handle(algo):
    import torch
    from glprox import GroupLassoProblem, solve_group
    m, n, l, G = 1024, 4096, 2, 4
    A, b, x0, u, mu = orc.gen_data(11, m, n, l, np.float32)
    opts = {"maxit": 30, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    probs, bounds, Ad = _shards(A, b, G, "f4")
    x0s = [torch.from_numpy(x0[c0:c1].copy()).cuda() for c0, c1 in bounds]
    xs, k, out = solve_group(probs, algo, x0s, mu, opts)
    single = GroupLassoProblem(Ad, torch.from_numpy(b).cuda())
    x1, k1, out1 = single.solve(algo, torch.from_numpy(x0).cuda(), mu, opts)
    assert k == k1 == 90
    assert np.max(np.abs(np.asarray(out["f_hist"]) - np.asarray(out1["f_hist"])) / np.abs(out1["f_hist"])) <= 1e-5
    x = torch.cat(xs, 0)
    assert (x - x1).abs().max().item() <= 1e-4 * x1.abs().max().item()
    # and against the fp64 oracle on the same fp32-rounded inputs
    xr, kr, outr = orc.solve(algo, x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
    assert abs(out["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])
    single.close()
    for p in probs:
        p.close()


@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
def test_fused_gather_decide_kernel_gives_identical_bits(algo, monkeypatch):
    """GLPROX_FUSED_GATHER=1 (glp_gather_decide_kernel: gather, grid barrier, decide in one launch) against the
    default two launches: same partial layout and reduction order, hence the same history bit for bit."""
    import torch
    from glprox import solve_group
    rec = load_golden("harsh_ls")
    A, b, x0, u, mu = golden_inputs(rec)
    hist = []
    for fused in ("0", "1"):
        monkeypatch.setenv("GLPROX_FUSED_GATHER", fused)
        probs, bounds, keep = _shards(A, b, 2, "f8")
        x0s = [torch.from_numpy(x0[c0:c1].copy()).cuda() for c0, c1 in bounds]
        xs, k, out = solve_group(probs, algo, x0s, mu, dict(rec["opts"]))
        hist.append((k, list(out["f_hist"]), torch.cat(xs, 0).cpu().numpy()))
        for p in probs:
            p.close()
    assert hist[0][0] == hist[1][0] == int(rec[algo + "_k"])
    assert hist[0][1] == hist[1][1] and np.array_equal(hist[0][2], hist[1][2])
