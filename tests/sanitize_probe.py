"""Small solves on every backend for compute-sanitizer runs (not a pytest file):
compute-sanitizer --tool memcheck python tests/sanitize_probe.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))


def main():
    import torch
    from glprox import GroupLassoProblem, shard_bounds, solve_group
    from oracle import glprox_oracle as orc
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    A, b, x0, u, mu = orc.gen_data(3, 96, 300, 2)
    if which in ("all", "stream", "persistent"):
        for backend in (("stream", "persistent") if which == "all" else (which,)):
            for algo in ("FProxGD", "ProxGD", "FGD", "SGD"):
                for dt in (np.float64, np.float32):
                    prob = GroupLassoProblem(torch.from_numpy(A.astype(dt)).cuda(), torch.from_numpy(b.astype(dt)).cuda())
                    x, k, out = prob.solve(algo, torch.from_numpy(x0.astype(dt)).cuda(), mu, {"maxit": 6, "backend": backend})
                    met = prob.metrics(x, torch.from_numpy(u.astype(dt)).cuda(), mu)
                    prob.close()
                    print(backend, algo, dt.__name__, k, float(out["fval"]), met["sparsity"], flush=True)
    if which in ("all", "wide"):
        m, n, l = 160, 640, 16
        A, b, x0, u, mu = orc.gen_data(5, m, n, l)
        A, b, x0 = (z.astype(np.float32) for z in (A, b, x0))
        opts = {"maxit": 4, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
        Ad, bd = torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda()
        for mode in ("auto", "3xtf32", "tf32+bf16", "tf32"):   # auto = fp16 pairs on the CTA-pair kernel (cta_group::2)
            prob = GroupLassoProblem(Ad, bd)
            x, k, out = prob.solve("FProxGD", torch.from_numpy(x0).cuda(), mu, dict(opts, tensor_cores=mode))
            prob.close()
            print("wide", mode, k, float(out["fval"]), flush=True)
    if which in ("all", "sharded"):
        m, n, l = 160, 768, 16
        A, b, x0, u, mu = orc.gen_data(5, m, n, l)
        A, b, x0 = (z.astype(np.float32) for z in (A, b, x0))
        opts = {"maxit": 4, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
        Ad, bd = torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda()
        for G, ll in ((3, 16), (2, 2)):
            bounds = shard_bounds(n, G, 128)
            bb, xx = bd[:, :ll].contiguous(), x0[:, :ll]
            probs = [GroupLassoProblem(Ad[:, c0:c1], bb) for c0, c1 in bounds]
            ptrs = [p.export_mailbox(G)[1] for p in probs]
            for r, p in enumerate(probs):
                p.attach_peers(r, G, n, ptrs, kind=1)
            xs, k, out = solve_group(probs, "FProxGD", [torch.from_numpy(xx[c0:c1].copy()).cuda() for c0, c1 in bounds], mu, opts)
            for p in probs:
                p.close()
            print("sharded", G, ll, k, float(out["fval"]), flush=True)
    if which in ("all", "admm"):
        A, b, x0, u, mu = orc.gen_data(3, 96, 300, 2)
        for blocked in ("0", "1"):   # single-CTA factor / blocked Cholesky + streams of Y = L^-T
            os.environ["GLPROX_ADMM_BLOCKED"] = blocked
            for dt in (np.float64, np.float32):
                prob = GroupLassoProblem(torch.from_numpy(A.astype(dt)).cuda(), torch.from_numpy(b.astype(dt)).cuda())
                x, k, out = prob.solve_admm_dual(torch.from_numpy(x0.astype(dt)).cuda(), mu, {"maxit": 5})
                prob.close()
                print("admm blocked=" + blocked, dt.__name__, k, float(out["fval"]), flush=True)
        os.environ.pop("GLPROX_ADMM_BLOCKED", None)
        prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
        x, k, out = prob.solve_admm_primal(torch.from_numpy(x0).cuda(), mu, {"maxit": 5})
        print("admm primal", k, float(out["fval"]), flush=True)
        x, k, out = prob.solve_alm_dual(torch.from_numpy(x0).cuda(), mu, {"maxit": 1})
        print("alm dual", k, float(out["fval"]), flush=True)
        prob.close()
    torch.cuda.synchronize()
    print("SANITIZE_PROBE_DONE", flush=True)


if __name__ == "__main__":
    main()
