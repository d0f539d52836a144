import sys, time, os, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/convex-optimization_b200')
import torch
from glprox import GroupLassoProblem
from oracle import glprox_oracle as orc
A,b,x0,u,mu = orc.gen_data(97006855,256,512,2)
prob = GroupLassoProblem(torch.from_numpy(A).cuda(), torch.from_numpy(b).cuda())
x0d = torch.from_numpy(x0).cuda()
for algo in ("FProxGD","ProxGD"):
    for be in ("stream","persistent"):
        for rep in range(3):
            x,k,out = prob.solve(algo, x0d, mu, {"backend":be})
        print(algo, be, "k=%d fval=%.12e device_ms=%.3f tt=%.4f it/s=%.0f launches=%d"%(k,out["fval"],out["glprox"]["device_ms"],out["tt"],k/(out["glprox"]["device_ms"]/1e3),out["glprox"]["n_launches"]))
t0=time.time(); xr,kr,outr = orc.solve("FProxGD",x0,A,b,mu,{}); print("cpu oracle FProxGD", kr, outr["tt"], kr/outr["tt"])
