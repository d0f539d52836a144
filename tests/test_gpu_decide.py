"""K3 at step level through the C ABI (glp_decide): the DEVICE build of csrc/glp_decide.h is fed the very
states the g++ build sees when tests/gpu_emulator.py drives the GPU schedule on the CPU, and must return the
same control block, action and recorded history entries bit for bit."""
import ctypes as C

import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from gpu_emulator import Ctrl, DecideOut, Reduced, SolveConst, emulate

pytestmark = pytest.mark.gpu


def _bytes(obj):
    return bytes(memoryview(obj).cast("B"))


@pytest.mark.parametrize("algo,name", [("FProxGD", "fixture_maxit40"), ("ProxGD", "harsh_ls"), ("ProxGD", "fixture_maxit40"),
                                       ("FProxGD", "seed2"), ("ProxGD", "steps_diminishing")])
def test_device_decide_equals_host_build(algo, name):
    from glprox import _lib
    lib = _lib.load()
    try:
        rec = load_golden(name)
    except FileNotFoundError:
        pytest.skip("golden case %s not present" % name)
    A, b, x0, u, mu = golden_inputs(rec)
    log = []
    x, k, out, info = emulate(algo, x0, A, b, mu, dict(rec["opts"]), decide_log=log)
    assert k == int(rec[algo + "_k"]) and len(log) >= k
    n = len(log)
    sc = log[0][0]
    cin = (Ctrl * n)(*[e[1] for e in log])
    red = (Reduced * n)(*[e[2] for e in log])
    cout = (Ctrl * n)()
    act = (DecideOut * n)()
    hist = np.empty((n, 3))
    _lib.check(lib.glp_decide(0, C.byref(sc), cin, red, log[0][3], cout, act, hist.ctypes.data, n))
    actions = set()
    for i, e in enumerate(log):
        want = Ctrl.from_buffer_copy(e[4])
        if e[5].action == 2:        # the emulator flips gy_sel after the step, as the decide kernel does
            want.gy_sel = e[1].gy_sel
        assert _bytes(cout[i]) == _bytes(want), (i, "control block differs")
        assert _bytes(act[i]) == _bytes(e[5]), (i, "action differs")
        actions.add(e[5].action)
        if e[6] is None:
            assert np.all(np.isnan(hist[i]))
        else:
            assert hist[i, 0] == e[6][0] and hist[i, 1] == e[6][1] and hist[i, 2] == want.s_cur
    assert 2 in actions and 3 in actions            # iterate and finish were both exercised
    if name == "harsh_ls":
        assert 1 in actions                           # and rejected trials


def test_sparsity_history_and_debug_log_lines(caplog):
    """The reference's DEBUG lines (gl_FProxGD_primal.py:45,56,149-151 / gl_ProxGD_primal.py:45,54,134-136):
    alpha0=, new mu= per stage, and every 100th iteration objective + sparsity of the new iterate -- compared
    with the lines the staged unmodified reference emits on the same problem."""
    import logging
    from oracle import glprox_oracle as orc
    from oracle import make_ref
    import gl_FProxGD_primal as ours_f
    import gl_ProxGD_primal as ours_p
    staged = make_ref.load()
    if staged is None:
        pytest.skip("oracle/_ref not staged (python oracle/make_ref.py in the build container)")
    A, b, x0, u, mu = orc.gen_data(97006855, 256, 512, 2)
    opts = {"maxit": 230}
    for ref_fn, our_fn in ((staged[1], ours_f.gl_FProxGD_primal), (staged[0], ours_p.gl_ProxGD_primal)):
        lines = []
        for fn in (ref_fn, our_fn):
            caplog.clear()
            with caplog.at_level(logging.DEBUG, logger="opt"):
                x, k, out = fn(np.copy(x0), A, b, mu, dict(opts))
            lines.append([r.getMessage() for r in caplog.records if r.name == "opt"])
        ref, got = lines
        assert len(ref) == len(got) and len(ref) >= 1 + 3 + 4, (len(ref), len(got))
        for a, c in zip(ref, got):
            if a.startswith("iter="):
                pa = [float(t.split("=")[1]) for t in a.split(",")]
                pc = [float(t.split("=")[1]) for t in c.split(",")]
                assert pa[0] == pc[0]
                assert abs(pa[1] - pc[1]) <= 2e-6 * abs(pa[1]) and abs(pa[2] - pc[2]) <= 1.5e-6, (a, c)
            else:
                assert a == c
