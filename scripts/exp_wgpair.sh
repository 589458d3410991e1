#!/bin/bash
# A/B of CTA pairs in the weight-gradient kernel (RN_WG_PAIR = 0 / default heuristic / 1)
for wl in imagenet50 wrn28_10; do
for v in 0 -1 1; do
  RN_WG_PAIR=$v timeout 200 python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --sustained-steps 0 --dump-layers gpurun_out/layers_wgpair${v}_$wl.txt > gpurun_out/wgpair${v}_$wl.json 2>gpurun_out/wgpair${v}_$wl.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/wgpair${v}_$wl.json").read().strip().split("\n")[-1]); print("RN_WG_PAIR=$v $wl", d["value"], d["ms_per_step"], [k["ms"] for k in d["kernels"] if k["kernel"]=="conv_wgrad_tc"])
except Exception as e: print("RN_WG_PAIR=$v $wl FAILED", e)
PY
done; done
