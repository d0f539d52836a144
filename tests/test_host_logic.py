"""CPU tests of the host-compilable pieces of the CUDA path: the control state machine
(csrc/glp_decide.h) driven through the GPU schedule, and the stream-K tile partition
(csrc/glp_partition.h)."""
import numpy as np
import pytest

from conftest import golden_inputs, load_golden
from gpu_emulator import emulate, shim

CASES = ["fixture", "seed114514", "fixture_maxit3", "fixture_maxit40", "ragged", "l1", "l3", "harsh_ls",
         "diminishing", "diminishing2", "thres1e-8"]


@pytest.mark.parametrize("algo", ["ProxGD", "FProxGD"])
@pytest.mark.parametrize("name", CASES)
def test_device_state_machine_reproduces_reference(name, algo):
    rec = load_golden(name)
    A, b, x0, u, mu = golden_inputs(rec)
    x, k, out, info = emulate(algo, x0, A, b, mu, rec["opts"])
    assert k == int(rec[algo + "_k"])
    ref = rec[algo + "_f_hist"]
    fh = np.asarray(out["f_hist"])
    assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 2e-11
    assert abs(out["fval"] - rec[algo + "_fval"]) <= 2e-11 * abs(rec[algo + "_fval"])
    assert np.array_equal(np.asarray(out["f_hist_best"]), rec[algo + "_f_hist_best"][:k] * 0 + np.minimum.accumulate(fh))
    assert np.max(np.abs(x - rec[algo + "_x"])) <= 1e-9
    ax = np.abs(x)
    assert np.array_equal(np.flatnonzero((ax > 1e-6 * ax.max()).any(axis=1)), rec[algo + "_support_rule"])
    assert info["n_tpass"] <= k and info["stage_end"][2] == k


def test_state_machine_edge_cases():
    from oracle import glprox_oracle as orc
    A, b, x0, u, mu = orc.gen_data(3, 32, 64, 2)
    x, k, out, info = emulate("ProxGD", x0, A, b, mu, {"maxit": 0})
    assert k == 0 and np.array_equal(x, x0) and out["f_hist"] == []
    assert abs(out["fval"] - orc.objective(A, b, x0, mu)) <= 1e-12 * abs(out["fval"])
    x, k, out, info = emulate("FProxGD", x0, A, b, mu, {"maxit": 0})
    assert k == 0 and info["ran_any"] == 0  # host maps this to UnboundLocalError like the reference
    # maxit_line_search_iter = 0: the carried / initial step is used untested
    a0 = float(1.0 / (np.sqrt(32) + np.sqrt(64)) ** 2)
    for algo in ("ProxGD", "FProxGD"):
        opts = {"maxit": 20, "maxit_line_search_iter": 0, "alpha0": a0}
        xr, kr, outr = orc.solve(algo, x0, A, b, mu, opts)
        x, k, out, info = emulate(algo, x0, A, b, mu, opts)
        assert k == kr and np.allclose(out["f_hist"], outr["f_hist"], rtol=1e-11, atol=0)


def test_partition_covers_every_tile_once_and_slots_are_unique():
    lib = shim()
    rng = np.random.default_rng(0)
    cases = [(148, 512, 1024), (148, 128, 256), (148, 4, 4), (7, 3, 5), (148, 1, 1), (148, 1, 3), (5, 9, 1),
             (148, 3, 100)]
    cases += [tuple(int(v) for v in rng.integers(1, 60, size=3)) for _ in range(40)]
    for G0, nblk, per in cases:
        N = nblk * per
        G = min(G0, N)
        lo = [lib.shim_tile_lo(c, G, N) for c in range(G + 1)]
        assert lo[0] == 0 and lo[-1] == N and all(lo[i] <= lo[i + 1] for i in range(G))
        assert max(lo[i + 1] - lo[i] for i in range(G)) - min(lo[i + 1] - lo[i] for i in range(G)) <= 1
        owner = np.empty(N, dtype=np.int64)
        for c in range(G):
            owner[lo[c]:lo[c + 1]] = c
        for t in list(range(min(N, 50))) + [N - 1] + list(rng.integers(0, N, size=20)):
            assert lib.shim_tile_owner(int(t), G, N) == owner[t]
        slots = set()
        for blk in range(nblk):
            first = lib.shim_contrib_first(blk, per, G, N)
            last = lib.shim_contrib_last(blk, per, G, N)
            touching = sorted(set(owner[blk * per:(blk + 1) * per].tolist()))
            assert touching == list(range(first, last + 1))
            for c in touching:
                assert (c + blk) not in slots and c + blk < G + nblk
                slots.add(c + blk)


def test_handle_path_refuses_wrong_shapes():
    """the C ABI takes raw pointers, so b / x0 / x / ref shapes are validated on the host before any
    library call (the reference raises NumPy's shape error at its first product)."""
    import torch
    from glprox.solver import GroupLassoProblem
    p = object.__new__(GroupLassoProblem)
    p.n, p.l, p.device, p.dtype = 12, 2, torch.device("cpu"), torch.float64
    good = torch.zeros(12, 2)
    assert p._iterate(good).shape == (12, 2) and p._iterate(good).dtype == torch.float64
    for bad in (torch.zeros(12, 3), torch.zeros(11, 2), torch.zeros(24), torch.zeros(12, 2, 1), np.zeros((12, 2))):
        with pytest.raises(ValueError):
            p._iterate(bad)
    # constructor: A must be a CUDA tensor (no CPU fallback), checked before b
    with pytest.raises(ValueError):
        GroupLassoProblem(torch.zeros(4, 8), torch.zeros(4, 2))


def test_truncated_solve_is_an_error():
    from glprox import _lib
    from glprox.solver import _raise_for
    res = _lib.GlpResult()
    res.status = _lib.GLP_OK
    _raise_for(_lib.GLP_OK, res)
    res.status, res.iters = _lib.GLP_ERR_STATE, 7
    with pytest.raises(RuntimeError):
        _raise_for(_lib.GLP_OK, res)


def test_debug_log_lines_follow_the_reference_format(caplog):
    """_log_history (host side of SURVEY 8a-14): alpha0=, new mu= per stage, every 100th iteration with the
    sparsity of the iterate produced by that iteration, nothing for the iteration a stage broke on."""
    import logging
    from glprox import _lib
    from glprox.solver import _log_history
    res = _lib.GlpResult()
    res.iters, res.sparsity = 450, 0.125
    res.stage_end[0], res.stage_end[1], res.stage_end[2] = 200, 300, 450      # stage 2 broke early at k=300
    f = np.arange(1, 451, dtype=np.float64)
    s = np.linspace(0.9, 0.1, 450)
    with caplog.at_level(logging.DEBUG, logger="opt"):
        _log_history(f, res, {"alpha0": 1e-3, "maxit": 200}, 1e-2, s)
    msgs = [r.getMessage() for r in caplog.records]
    assert msgs[0] == "alpha0= {:10E}".format(1e-3)
    assert msgs[1] == "new mu= {:10E}".format(1.0) and "new mu= {:10E}".format(0.1) in msgs and "new mu= {:10E}".format(1e-2) in msgs
    iters = [m for m in msgs if m.startswith("iter=")]
    assert iters[0] == "iter= {:5}, objective= {:10E}, sparsity= {:3f}".format(100, 100.0, s[100])
    assert [int(m.split(",")[0].split("=")[1]) for m in iters] == [100, 200, 400]   # 300: the stage broke there
