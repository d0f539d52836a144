"""GPU tests of the opt-in Barzilai-Borwein step of gl_ProxGD_primal (opts['step_type'] = 'bb'; the
reference has no such rule -- parity unpinned, see tests/test_bb_cpu.py).  The device takes the BB value
from the line-search scalars it has already reduced; the CPU schedule emulator runs the same state
machine (csrc/glp_decide.h) on NumPy products, the oracle computes the step from its definition."""
import numpy as np
import pytest

import gpu_emulator as emu
from oracle import glprox_oracle as orc

pytestmark = pytest.mark.gpu
F_STAR = 6.1037724766016532e-01


def _solve(x0, A, b, mu, opts):
    import gl_ProxGD_primal
    return gl_ProxGD_primal.gl_ProxGD_primal(x0, A, b, mu, opts)


@pytest.mark.parametrize("backend", ["stream", "persistent"])
def test_bb_first_iterations_match_the_cpu_schedule(backend):
    A, b, x0, u, mu = orc.gen_data(97006855)
    opts = {"step_type": "bb", "maxit": 12}
    x, k, out = _solve(x0, A, b, mu, dict(opts, backend=backend))
    xe, ke, oe, info = emu.emulate("ProxGD", x0.copy(), A, b, mu, opts)
    xo, ko, oo = orc.solve("ProxGD", x0.copy(), A, b, mu, opts)
    assert k == ke == ko == 36 and out["glprox"]["backend"] == backend
    fh = np.asarray(out["f_hist"])
    assert np.max(np.abs(fh - np.asarray(oe["f_hist"])) / np.abs(fh)) <= 1e-10      # same state machine
    assert np.max(np.abs(fh[:12] - np.asarray(oo["f_hist"])[:12]) / np.abs(fh[:12])) <= 1e-10  # definition of BB1
    assert abs(out["glprox"]["n_fwd"] - info["n_fwd"]) <= 2   # a borderline trial may round the other way


@pytest.mark.parametrize("backend", ["stream", "persistent"])
def test_bb_converges_to_the_optimum_with_fewer_streams(backend):
    A, b, x0, u, mu = orc.gen_data(97006855)
    x, k, out = _solve(x0, A, b, mu, {"step_type": "bb", "thres": 1e-8, "backend": backend})
    assert k < 1000 and out["glprox"]["n_fwd"] < 0.5 * 5720    # reference rule with the same thres: 1916 / 5720
    assert abs(out["fval"] - F_STAR) <= 1e-10 * F_STAR
    assert abs(orc.objective(A, b, x, mu) - F_STAR) <= 1e-10 * F_STAR


def test_bb_default_threshold_reaches_the_reference_objective_and_support():
    A, b, x0, u, mu = orc.gen_data(97006855)
    ref = 6.103773850792892e-01
    x, k, out = _solve(x0, A, b, mu, {"step_type": "bb"})
    assert abs(out["fval"] - ref) <= 1e-9 * ref
    ax = np.abs(x)
    assert np.count_nonzero((ax > 1e-6 * ax.max()).any(axis=1)) == 51


def test_bb_fp32_streaming_problem_beats_the_fixed_restart():
    """m=2048 n=8192 fp32 (streaming backend): with the safe step 1/L as alpha0 the reference rule can only
    shrink the step; BB lengthens it and reaches a lower objective within the same iteration budget."""
    import torch
    from glprox import GroupLassoProblem
    m, n, l = 2048, 8192, 2
    A, b, x0, u, mu = orc.gen_data(5, m, n, l, np.float32)
    alpha0 = float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x0d = torch.from_numpy(x0).cuda()
    res = {}
    for st in ("line_search", "bb"):
        x, k, out = prob.solve("ProxGD", x0d, mu, {"maxit": 150, "alpha0": alpha0, "step_type": st, "backend": "stream"})
        f_true = orc.objective(A.astype(np.float64), b.astype(np.float64), x.cpu().numpy().astype(np.float64), mu)
        assert abs(out["fval"] - f_true) <= 1e-5 * f_true          # the reported objective is the true one
        res[st] = (k, out["fval"], out["glprox"]["n_fwd"])
    assert res["bb"][1] < res["line_search"][1]
    prob.close()


def test_bb_sharded_matches_single_handle():
    import torch
    from glprox import GroupLassoProblem, shard_bounds, solve_group
    A, b, x0, u, mu = orc.gen_data(97006855)
    opts = {"step_type": "bb", "maxit": 10}   # BB dynamics amplify rounding differences: compare a short run
    Ad, bd = torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda()
    bounds = shard_bounds(512, 4, 64)
    probs = [GroupLassoProblem(Ad[:, c0:c1], bd) for c0, c1 in bounds]
    ptrs = [p.export_mailbox(4)[1] for p in probs]
    for r, p in enumerate(probs):
        p.attach_peers(r, 4, 512, ptrs, kind=1)
    xs, k, out = solve_group(probs, "ProxGD", [torch.from_numpy(x0[c0:c1].copy()).cuda() for c0, c1 in bounds], mu, opts)
    single = GroupLassoProblem(Ad, bd)
    x1, k1, out1 = single.solve("ProxGD", torch.from_numpy(x0).cuda(), mu, dict(opts, backend="stream"))
    assert k == k1
    f, f1 = np.asarray(out["f_hist"]), np.asarray(out1["f_hist"])
    assert np.max(np.abs(f - f1) / np.abs(f1)) <= 1e-9
    assert abs(out["fval"] - out1["fval"]) <= 1e-8 * abs(out1["fval"])
    single.close()
    for p in probs:
        p.close()
