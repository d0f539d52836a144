"""Probe (not a test): objective vs device time on a streaming-size problem for the reference's rules
and the opt-in BB rule.  usage: python tests/bench_bb_probe.py [m n]  -> one JSON line per run."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))


def main():
    import torch
    from glprox import GroupLassoProblem, dataset
    m, n = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (8192, 32768)
    A, b, x0, u, mu, _ = dataset.gen_data_device(m, n, 2, 97006855, "f32")
    prob = GroupLassoProblem(A, b)
    alpha0 = float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)
    runs = [("FProxGD", {"alpha0": alpha0}), ("ProxGD", {"alpha0": alpha0, "maxit": 1500}),
            ("ProxGD", {"alpha0": alpha0, "maxit": 1500, "step_type": "bb"})]
    best = None
    rows = []
    for algo, opts in runs:
        prob.solve(algo, x0, mu, dict(opts, maxit=5))  # warm
        x, k, out = prob.solve(algo, x0, mu, opts)
        met = prob.metrics(x, u, mu)
        fh = np.asarray(out["f_hist"])
        rows.append((algo, opts.get("step_type", "line_search"), k, out, fh, met))
        best = fh.min() if best is None else min(best, fh.min())
    for algo, st, k, out, fh, met in rows:
        secs = out["glprox"]["device_ms"] * 1e-3
        hit = {}
        for tol in (1e-2, 1e-4, 1e-6):
            idx = np.flatnonzero(fh <= best * (1 + tol))
            hit["k_to_%g" % tol] = int(idx[0]) + 1 if idx.size else None
            hit["s_to_%g" % tol] = secs * (int(idx[0]) + 1) / k if idx.size else None
        print(json.dumps({"m": m, "n": n, "algo": algo, "step_type": st, "iters": k, "seconds": secs,
                          "fval": float(out["fval"]), "best_over_runs": float(best), "n_fwd": out["glprox"]["n_fwd"],
                          "sparsity": met["sparsity"], "err_exact": met["err_exact"], **hit}))
    prob.close()


if __name__ == "__main__":
    main()
