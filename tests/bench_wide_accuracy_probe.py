"""Probe (not a test): error of the wide-mode operand sets against the fp64 oracle, same inputs."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))
from oracle import glprox_oracle as orc
from gl_FProxGD_primal import gl_FProxGD_primal
from gl_ProxGD_primal import gl_ProxGD_primal

for (m, n, l, maxit) in [(512, 2048, 128, 15), (1024, 4096, 64, 15), (2048, 8192, 32, 40)]:
    A, b, x0, u, mu = orc.gen_data(7, m, n, l)
    A, b, x0 = (z.astype(np.float32) for z in (A, b, x0))
    opts = {"maxit": maxit, "alpha0": float(1.0 / (np.sqrt(m) + np.sqrt(n)) ** 2)}
    for algo, fn in (("FProxGD", gl_FProxGD_primal), ("ProxGD", gl_ProxGD_primal)):
        xr, kr, outr = orc.solve(algo, x0.astype(np.float64), A.astype(np.float64), b.astype(np.float64), mu, opts)
        fr = np.asarray(outr["f_hist"])
        for mode in ("3xtf32", "f16x2", "tf32+bf16", "off"):
            x, k, out = fn(x0, A, b, mu, dict(opts, tensor_cores=mode))
            fh = np.asarray(out["f_hist"])
            print("m=%d n=%d l=%d %-8s %-10s k=%d  max rel f_hist err %.2e  final %.2e  max|x-x_ref| %.2e" % (
                m, n, l, algo, mode, k, np.max(np.abs(fh - fr) / np.abs(fr)), abs(out["fval"] - outr["fval"]) / abs(outr["fval"]),
                np.max(np.abs(x - xr))), flush=True)
