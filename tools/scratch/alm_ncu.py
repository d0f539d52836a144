import os, sys
ROOT="/root/repo"
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "convex-optimization_b200"))
import torch
from glprox import GroupLassoProblem, dataset
A, b, x0, u, mu, _ = dataset.gen_data_device(8192, 32768, 2, 7, "f64")
prob = GroupLassoProblem(A, b)
x, k, out = prob.solve_alm_dual(x0, mu, {"maxit": 1})
print(out["glprox"])
x, k, out = prob.solve("FProxGD", x0, mu, {"maxit": 20, "backend": "stream", "alpha0": 1.0/(8192**.5+32768**.5)**2})
print("FProxGD fp64", k, out["glprox"], "ms/iter", out["glprox"]["device_ms"]/k)
