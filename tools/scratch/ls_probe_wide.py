"""Probe: forward products per iteration in the wide (l = 128, tensor-core) mode at the C5 size."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))
import numpy as np
import torch
from glprox import GroupLassoProblem, dataset
m, n, l = 16384, 65536, 128
A, b, x0, u, mu, _ = dataset.gen_data_device(m, n, l, 7, "f32")
prob = GroupLassoProblem(A, b)
a0 = float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)
for algo in ("ProxGD", "FProxGD"):
    for maxit in (10, 50, 150):
        x, k, out = prob.solve(algo, x0, mu, {"maxit": maxit, "alpha0": a0})
        g = out["glprox"]
        print("%d x %d l=%d f32 %-7s maxit %3d: k=%3d forward products %4d (%.2f per iteration) fval %.8e  %.1f ms"
              % (m, n, l, algo, maxit, k, g["n_fwd"], (g["n_fwd"] - 1) / max(k, 1), out["fval"], g["device_ms"]), flush=True)
