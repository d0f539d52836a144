"""Probe (not a test): time gl_Admm_dual on the homework fixture; run under ncu for the launch list."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))
from oracle import glprox_oracle as orc
from gl_ADMM_dual import gl_Admm_dual

m, n, l = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (256, 512, 2)))
A, b, x0, u, mu = orc.gen_data(97006855, m, n, l)
reps = int(os.environ.get("REPS", "3"))
for r in range(reps):
    t0 = time.perf_counter()
    x, k, out = gl_Admm_dual(x0, A, b, mu, {})
    dt = time.perf_counter() - t0
    g = out.get("glprox", {})
    print("rep %d  k=%d fval=%.12e host_s=%.4f device_ms=%s launches=%s" % (
        r, k, out["fval"], dt, g.get("device_ms"), g.get("n_launches")))
