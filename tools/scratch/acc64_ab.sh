#!/bin/bash
# A/B of K1's second accumulation level (run on a GPU box: gpurun -- tools/scratch/acc64_ab.sh).
#   default build : fp32 FMAs flushed every 2 tiles into an fp64 second level, fp64 cross-lane reduction
#   round-1 build : -DGLP_FWD_ACC64=0 -DGLP_FWD_FLUSH=8 (fp32 second level, flushed every 8 tiles)
# Prints, per build, the forward streams per iteration of ProxGD / FProxGD at C3 and C4 (tools/scratch/ls_probe.py) and the
# headline bench line.  profiles/r02_ab_k1_acc64.log is the output of the run that decided the default (there the two
# builds were named "default" = round 1 and "acc64" = today's default).
L=convex-optimization_b200/glprox/libglprox.so
cp $L /tmp/default.so
(cd convex-optimization_b200/csrc && nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC \
   --expt-relaxed-constexpr -DGLP_FWD_ACC64=0 -DGLP_FWD_FLUSH=8 -shared -cudart static -o /tmp/round1.so glprox.cu -ldl -lpthread) || exit 1
for v in default round1; do
  cp /tmp/$v.so $L
  echo "=== $v"
  python tools/scratch/ls_probe.py 2>&1 | grep "32768 x 131072 f32\|8192 x  32768 f32"
  python bench.py --steps 3 --warmup 3 --no-e2e --no-converge --no-extra 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); k=d['roofline']['kernels']
        print('bench it/s %.2f  K1 %.0f GB/s  K2 %.0f GB/s  parity rel_diff %.3e f_hist %.3e'%(d['value'],k['glp_fwd_kernel']['GBps'],k['glp_tpass_kernel']['GBps'],d['parity']['rel_diff'],d['parity']['f_hist_max_rel_diff']))
"
done
cp /tmp/default.so $L
