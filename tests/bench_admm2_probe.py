"""Timing probe (not a test): gl_Admm_primal and gl_Alm_dual on the homework fixture and on a streaming-size problem.
    python tests/bench_admm2_probe.py"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))

import torch  # noqa: E402
from glprox import GroupLassoProblem  # noqa: E402
from glprox import dataset  # noqa: E402
from oracle import glprox_oracle as orc  # noqa: E402


def fixture():
    A, b, x0, u, mu = orc.gen_data(97006855, 256, 512, 2)
    prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
    x0d = torch.from_numpy(x0).cuda()
    for name, fn, ofn in (("ADMM_primal", prob.solve_admm_primal, orc.solve_admm_primal),
                          ("ALM_dual", prob.solve_alm_dual, orc.solve_alm_dual)):
        fn(x0d, mu, {"maxit": 2})
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        x, k, out = fn(x0d, mu, {})
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        xr, kr, outr = ofn(x0, A, b, mu, {})
        t2 = time.perf_counter()
        print("fixture %-12s k=%3d fval=%.10e  GPU %.1f ms (device %.1f ms, %d launches)   NumPy port %.1f ms (k=%d)   published %s s"
              % (name, k, out["fval"], 1e3 * (t1 - t0), out["glprox"]["device_ms"], out["glprox"]["n_launches"], 1e3 * (t2 - t1), kr,
                 {"ADMM_primal": "6.23", "ALM_dual": "10.40"}[name]), flush=True)
    prob.close()


def streaming(m, n, maxit_p, maxit_a):
    A, b, x0, u, mu, _ = dataset.gen_data_device(m, n, 2, 7, "f64")
    prob = GroupLassoProblem(A, b)
    for name, fn, mi in (("ADMM_primal", prob.solve_admm_primal, maxit_p), ("ALM_dual", prob.solve_alm_dual, maxit_a)):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        x, k, out = fn(x0, mu, {"maxit": mi})
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        g = out["glprox"]
        print("%d x %d fp64 %-12s k=%3d fval=%.6e  %.1f ms (device %.1f ms), %d forward + %d transposed streams of A, %d launches"
              % (m, n, name, k, out["fval"], 1e3 * (t1 - t0), g["device_ms"], g["n_fwd"], g["n_tpass"], g["n_launches"]), flush=True)
    prob.close()


if __name__ == "__main__":
    fixture()
    streaming(4096, 16384, 20, 1)
    streaming(8192, 32768, 20, 1)
