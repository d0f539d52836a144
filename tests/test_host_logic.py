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
