"""Opt-in Barzilai-Borwein first trial step of gl_ProxGD_primal (opts['step_type'] = 'bb').

The reference has no such rule, so there is no golden vector for it: PARITY UNPINNED.
What is checked instead: the state machine the GPU runs (csrc/glp_decide.h, compiled for the host) takes
the BB value from line-search scalars by the quadratic identity ||A d||^2 = rr_new - rr_base + 2 t <grad, gt>;
the oracle computes ||d||^2 / ||A d||^2 from its definition with an extra product.  The two agree to
rounding until the (non-monotone, rounding-chaotic) BB dynamics amplify the difference, and both
converge to the optimum the reference's rule reaches."""
import numpy as np
import pytest

import gpu_emulator as emu
from oracle import glprox_oracle as orc

F_STAR = 6.1037724766016532e-01


def test_bb_identity_matches_definition_on_the_first_iterations():
    A, b, x0, u, mu = orc.gen_data(97006855)
    opts = {"step_type": "bb", "maxit": 12}
    tr = {}
    x, k, out = orc.solve("ProxGD", x0.copy(), A, b, mu, opts, trace=tr)
    xe, ke, oe, info = emu.emulate("ProxGD", x0.copy(), A, b, mu, opts)
    assert k == ke == 36
    fh, fe = np.asarray(out["f_hist"]), np.asarray(oe["f_hist"])
    assert np.max(np.abs(fh[:12] - fe[:12]) / np.abs(fh[:12])) <= 1e-12
    steps = np.asarray(tr["steps"])
    assert steps[0] <= 2e-3 and len(set(np.round(steps[:12], 12))) > 6      # the step adapts every iteration
    assert np.all(steps >= 2e-3 * 1e-3) and np.all(steps <= 2e-3 * 1e3)     # safeguard interval


@pytest.mark.parametrize("impl", ["oracle", "schedule"])
def test_bb_reaches_the_optimum_in_fewer_streams(impl):
    A, b, x0, u, mu = orc.gen_data(97006855)
    opts = {"step_type": "bb", "thres": 1e-8}
    if impl == "oracle":
        x, k, out = orc.solve("ProxGD", x0.copy(), A, b, mu, opts)
    else:
        x, k, out, info = emu.emulate("ProxGD", x0.copy(), A, b, mu, opts)
        assert info["n_fwd"] < 0.5 * 5720        # reference rule, same thres: 1916 iterations, 5720 forward streams
    assert k < 1000                               # reference rule: 1916
    assert abs(out["fval"] - F_STAR) <= 1e-10 * F_STAR


def test_bb_with_default_threshold_reaches_the_reference_objective():
    A, b, x0, u, mu = orc.gen_data(97006855)
    ref = 6.103773850792892e-01                   # gl_ProxGD_primal defaults
    x, k, out, info = emu.emulate("ProxGD", x0.copy(), A, b, mu, {"step_type": "bb"})
    assert abs(out["fval"] - ref) <= 1e-9 * ref
    ax = np.abs(x)
    assert np.count_nonzero((ax > 1e-6 * ax.max()).any(axis=1)) == 51   # the rows the reference keeps


def test_bb_is_rejected_for_the_other_solvers():
    import sys, os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "convex-optimization_b200"))
    from glprox.solver import merge_opts, pack_opts
    assert pack_opts("ProxGD", merge_opts("ProxGD", {"step_type": "bb"})).step_type == 4
    for algo in ("FProxGD", "FGD", "GD", "SGD"):
        with pytest.raises(ValueError, match="gl_ProxGD_primal only"):
            pack_opts(algo, merge_opts(algo, {"step_type": "bb"}))
    with pytest.raises(ValueError):
        orc.solve("FProxGD", np.zeros((4, 2)), np.zeros((3, 4)), np.zeros((3, 2)), 1e-2, {"step_type": "bb"})
