#!/bin/bash
# A/B of where the fused epilogues' partial sums are finalised (RN_BN_FINALIZE = kernel / apply / conv)
run() {  # name, env assignments..., workload
  local name=$1; shift; local wl=${@: -1}; set -- "${@:1:$(($#-1))}"
  env "$@" timeout 200 python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --sustained-steps 0 > gpurun_out/exp_${name}_$wl.json 2>gpurun_out/exp_${name}_$wl.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/exp_${name}_$wl.json").read().strip().split("\n")[-1]); print("$name $wl", d["value"], d["ms_per_step"], d["gpu_launches"]//d["steps"])
except Exception as e: print("$name $wl FAILED", e)
PY
}
for wl in imagenet50 cifar50 cifar20 wrn28_10; do
  run kernel RN_BN_FINALIZE=kernel $wl
  run apply RN_BN_FINALIZE=apply $wl
  run conv256 RN_BN_FINALIZE=conv $wl
done
