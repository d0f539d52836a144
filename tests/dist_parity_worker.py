"""Worker of tests/test_gpu_multigpu.py (launched with torch.distributed.run, one process per GPU):
sharded drop-in solves through opts['dist'], checked against the golden vectors of the reference.
Prints one line 'DIST_PARITY_OK <world>' from rank 0 on success."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "convex-optimization_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    import torch
    import torch.distributed as dist
    from conftest import golden_inputs, load_golden
    from gl_FProxGD_primal import gl_FProxGD_primal
    from gl_ProxGD_primal import gl_ProxGD_primal
    from glprox import shard_bounds
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    for name in ("fixture_maxit40", "harsh_ls", "fixture"):
        rec = load_golden(name)
        A, b, x0, u, mu = golden_inputs(rec)
        c0, c1 = shard_bounds(A.shape[1], world, 64)[rank]
        for algo, fn in (("FProxGD", gl_FProxGD_primal), ("ProxGD", gl_ProxGD_primal)):
            if name == "fixture" and algo == "ProxGD":
                continue
            opts = dict(rec["opts"])
            opts["dist"] = True
            x, k, out = fn(x0[c0:c1].copy(), np.ascontiguousarray(A[:, c0:c1]), b, mu, opts)
            assert k == int(rec[algo + "_k"]), (name, algo, k)
            fh, ref = np.asarray(out["f_hist"]), rec[algo + "_f_hist"]
            assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10, (name, algo)
            assert abs(out["fval"] - rec[algo + "_fval"]) <= 1e-10 * abs(rec[algo + "_fval"])
            assert np.max(np.abs(x - rec[algo + "_x"][c0:c1])) <= 1e-8
            # all ranks took bit-identical decisions
            t = torch.tensor([out["fval"], float(k)], device="cuda", dtype=torch.float64)
            lst = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(lst, t)
            assert all(torch.equal(lst[0], z) for z in lst)
    # baseline exchange (two ncclAllReduce calls per forward stream instead of the peer-mailbox push): same bar
    from glprox import GroupLassoProblem, attach_distributed
    for name in ("fixture_maxit40", "harsh_ls"):
        rec = load_golden(name)
        A, b, x0, u, mu = golden_inputs(rec)
        c0, c1 = shard_bounds(A.shape[1], world, 64)[rank]
        prob = GroupLassoProblem(torch.from_numpy(np.ascontiguousarray(A[:, c0:c1])).cuda(), torch.from_numpy(b).cuda())
        attach_distributed(prob, A.shape[1], exchange="nccl")
        for algo in ("FProxGD", "ProxGD"):
            for rep in range(2):
                x, k, out = prob.solve(algo, torch.from_numpy(x0[c0:c1].copy()).cuda(), mu, dict(rec["opts"]))
                assert k == int(rec[algo + "_k"]), (name, algo, k)
                fh, ref = np.asarray(out["f_hist"]), rec[algo + "_f_hist"]
                assert np.max(np.abs(fh - ref) / np.abs(ref)) <= 1e-10, (name, algo)
                assert np.max(np.abs(x.cpu().numpy() - rec[algo + "_x"][c0:c1])) <= 1e-8
                t = torch.tensor([out["fval"], float(k)], device="cuda", dtype=torch.float64)
                lst = [torch.empty_like(t) for _ in range(world)]
                dist.all_gather(lst, t)
                assert all(torch.equal(lst[0], z) for z in lst)
        prob.close()
    # wide multi-RHS mode (l = 32, fp32): tcgen05 products per shard, reduce-scatter + all-gather over NVLink
    from oracle import glprox_oracle as orc
    m, n, l = 512, 1024 * world, 32
    A, b, x0, u, mu = orc.gen_data(7, m, n, l)
    A, b, x0 = (z.astype(np.float32) for z in (A, b, x0))
    wopts = {"maxit": 15, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    xr, kr, outr = orc.solve("FProxGD", x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, wopts)
    c0, c1 = shard_bounds(n, world, 128)[rank]
    x, k, out = gl_FProxGD_primal(x0[c0:c1].copy(), np.ascontiguousarray(A[:, c0:c1]), b, mu, dict(wopts, dist=True))
    assert k == kr == 45, k
    fh, fr = np.asarray(out["f_hist"]), np.asarray(outr["f_hist"])
    assert np.max(np.abs(fh - fr) / np.abs(fr)) <= 1e-5
    assert abs(out["fval"] - outr["fval"]) <= 1e-5 * abs(outr["fval"])
    t = torch.tensor([out["fval"], float(k)], device="cuda", dtype=torch.float64)
    lst = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(lst, t)
    assert all(torch.equal(lst[0], z) for z in lst)
    dist.barrier()
    if rank == 0:
        print("DIST_PARITY_OK %d" % world, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
