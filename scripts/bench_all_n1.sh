#!/bin/bash
# Every single-GPU bench line kept under profiles/: the default run (C3), the CIFAR workloads, BASELINE config 2 at
# global batch 128, and the reference arm.  Writes gpurun_out/<tag>_bench_*.json.
TAG=${1:-r02}
timeout 600 python bench.py --dump-layers gpurun_out/${TAG}_per_layer_n1.txt > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err
tail -c 300 gpurun_out/${TAG}_bench_n1.err
for wl in cifar50 cifar20 wrn28_10; do
  timeout 300 python bench.py --workload $wl > gpurun_out/${TAG}_bench_${wl}_n1.json 2> gpurun_out/${TAG}_bench_${wl}_n1.err
done
timeout 300 python bench.py --workload cifar50 --global-batch 128 --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/${TAG}_bench_cifar50_gb128_n1.json 2> gpurun_out/${TAG}_bench_cifar50_gb128_n1.err
timeout 600 python bench.py --impl reference > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
python - <<PY
import json
for f in ["bench_n1", "bench_cifar50_n1", "bench_cifar20_n1", "bench_wrn28_10_n1", "bench_cifar50_gb128_n1", "bench_reference"]:
    try:
        d = json.loads(open("gpurun_out/${TAG}_%s.json" % f).read().strip().split("\n")[-1])
        print(f, d.get("value"), d.get("ms_per_step"), "e2e", d.get("e2e", {}).get("value"), "launches/step",
              (d.get("gpu_launches") or 0) // max(d.get("steps", 1), 1), "roofline", d.get("roofline", {}).get("frac"),
              d.get("roofline", {}).get("traffic"), "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(f, "ERR", e)
PY
