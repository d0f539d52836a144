"""CPU tests (gloo, world_size 2) of the host side of the sharded path: column partition, the handle
all-gather, and the sharding algebra itself -- two processes each own a column block of A, run the
real device state machine (csrc/glp_decide.h via the host shim) on rank-ordered sums of partial
products and scalars, and must reproduce the golden vectors of the unmodified reference while
taking bit-identical decisions on both ranks."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_bounds_partition_columns():
    from glprox import shard_bounds
    for n, w, a in [(131072, 8, 128), (512, 8, 64), (391, 2, 64), (4096, 3, 128), (130, 1, 128), (200, 2, 128)]:
        bnd = shard_bounds(n, w, a)
        assert bnd[0][0] == 0 and bnd[-1][1] == n and len(bnd) == w
        assert all(bnd[i][1] == bnd[i + 1][0] for i in range(w - 1))
        assert all(c0 % a == 0 and c1 > c0 for c0, c1 in bnd)
        sizes = [c1 - c0 for c0, c1 in bnd]
        assert max(sizes) - min(sizes) <= a
    with pytest.raises(ValueError):
        shard_bounds(100, 4, 128)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    from conftest import golden_inputs, load_golden
    from glprox import exchange_handles, shard_bounds
    from gpu_emulator import emulate
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # (1) handle exchange: fixed-size byte strings, concatenated in rank order on every rank
        mine = bytes([rank * 16 + (i % 16) for i in range(64)])
        blob = exchange_handles(mine)
        assert len(blob) == 64 * world
        for r in range(world):
            assert blob[64 * r:64 * (r + 1)] == bytes([r * 16 + (i % 16) for i in range(64)])
        # (2) sharded solve: partial products summed over ranks in rank order
        rec = load_golden("fixture_maxit40")
        A, b, x0, u, mu = golden_inputs(rec)
        n = A.shape[1]
        c0, c1 = shard_bounds(n, world, 64)[rank]
        Aloc = A[:, c0:c1]

        def gather_sum(part):
            t = torch.from_numpy(np.ascontiguousarray(part))
            lst = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(lst, t)
            tot = lst[0].clone()
            for z in lst[1:]:
                tot += z  # rank order, identical on both ranks
            return tot.numpy()

        def fwd(X):       # X is the full n x cols operand in the emulator; a rank only uses its rows
            return gather_sum(Aloc @ X[c0:c1])

        def bwd(R):       # every rank owns grad[J_rank]; assemble the full vector for the emulator
            g = np.zeros((n, R.shape[1]))
            g[c0:c1] = Aloc.T @ R
            return gather_sum(g)

        for algo in ("FProxGD", "ProxGD"):
            x, k, out, info = emulate(algo, x0, A, b, mu, rec["opts"], matmul=fwd, rmatmul=bwd)
            assert k == int(rec[algo + "_k"])
            fh, ref = np.asarray(out["f_hist"]), rec[algo + "_f_hist"]
            assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10
            t = torch.tensor([out["fval"], float(k), float(info["n_fwd"])], dtype=torch.float64)
            lst = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(lst, t)
            assert all(torch.equal(lst[0], z) for z in lst)  # bit-identical decisions
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        q.put((rank, "FAIL %r" % (e,)))
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_sharded_solve_matches_reference():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29650 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
