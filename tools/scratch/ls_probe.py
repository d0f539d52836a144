"""Probe (not a test): how many line-search trials gl_ProxGD_primal / gl_FProxGD_primal take per iteration on the
stream backend, by size and dtype.  python tools/scratch/ls_probe.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))
import numpy as np
import torch
from glprox import GroupLassoProblem, dataset

for m, n in ((2048, 8192), (8192, 32768), (32768, 131072)):
    for dt in ("f32", "f64"):
        A, b, x0, u, mu, _ = dataset.gen_data_device(m, n, 2, 7, dt)
        prob = GroupLassoProblem(A, b)
        a0 = float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)
        for algo in ("ProxGD", "FProxGD"):
            for maxit in (10, 50):
                x, k, out = prob.solve(algo, x0, mu, {"maxit": maxit, "alpha0": a0, "backend": "stream"})
                g = out["glprox"]
                print("%6d x %6d %s %-7s maxit %3d: k=%3d forward streams %4d (%.2f per iteration) fval %.8e  %.1f ms"
                      % (m, n, dt, algo, maxit, k, g["n_fwd"], (g["n_fwd"] - 1) / max(k, 1), out["fval"], g["device_ms"]), flush=True)
        prob.close()
        del A, b, x0, u, prob
        torch.cuda.empty_cache()
