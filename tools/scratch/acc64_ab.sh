#!/bin/bash
# A/B of the fp64 second-level accumulation in K1 (GLP_FWD_ACC64=1, GLP_FWD_FLUSH=2) against the default build.
L=convex-optimization_b200/glprox/libglprox.so
cp $L /tmp/default.so
for v in default acc64; do
  if [ $v = acc64 ]; then cp tools/scratch/libglprox_acc64.so $L; else cp /tmp/default.so $L; fi
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
