#!/bin/bash
# A/B of the multi-GPU launch-latency switches on one box with N GPUs (run under gpurun --gpus N).
#   tools/multigpu_ab.sh N            -> gpurun_out/multigpu_ab_nN.log (one line per variant) + the full default line
# Variants: default (PDL + fused gather/decide + one system fence per CTA), each switch off alone, all off (round 1).
N=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
run() {  # name, env..., extra flags via FLAGS
  local name=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
      bench.py --gpus $N --steps 10 --warmup 3 $FLAGS > $OUT/_ab_$name.log 2>&1
  python - "$name" $OUT/_ab_$name.log <<'PY' >> $OUT/multigpu_ab_n$N.log
import json, sys
name, path = sys.argv[1], sys.argv[2]
lines = [l for l in open(path) if l.startswith("{")]
if not lines:
    print(name, "FAILED", open(path).read()[-400:].replace("\n", " | "))
else:
    d = json.loads(lines[-1])
    k = d["roofline"]["kernels"]
    print("%-22s it/s %9.2f  ms/iter %.4f  fwd %.1f GB/s  tpass %.1f GB/s  obj %.10e" % (
        name, d["value"], 1e3 / d["value"], k["glp_fwd_kernel"]["GBps"], k["glp_tpass_kernel"]["GBps"], d["final_objective"]))
PY
}
: > $OUT/multigpu_ab_n$N.log
FLAGS="--no-cpu-baseline --no-e2e --no-extra --no-converge"
run default_on       GLPROX_PDL=1
run pdl_off          GLPROX_PDL=0
run fused_gather_off GLPROX_FUSED_GATHER=0
run fence_per_block  GLPROX_PUSH_FENCE=block
run all_off_round1   GLPROX_PDL=0 GLPROX_FUSED_GATHER=0 GLPROX_PUSH_FENCE=block
cat $OUT/multigpu_ab_n$N.log
