#!/bin/bash
# bench.py under torchrun on N GPUs of this box: the C3 weak-scaling line and the C2 (global batch 128) strong-scaling line
N=$1; TAG=${2:-r02}
P=$((29500 + N))
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P bench.py --gpus $N "$@"; }
run --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
run --workload cifar50 --global-batch 128 --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/${TAG}_bench_cifar50_gb128_n$N.json 2> gpurun_out/${TAG}_bench_cifar50_gb128_n$N.err
python - <<PY
import json
for f in ("gpurun_out/${TAG}_bench_n$N.json", "gpurun_out/${TAG}_bench_cifar50_gb128_n$N.json"):
    try:
        d = json.loads(open(f).read().strip().split("\n")[-1])
        print(f, d["value"], d["ms_per_step"], d["scaling"], "e2e", d["e2e"]["value"], "ddp", d.get("ddp_check"), "ar", d.get("allreduce"))
    except Exception as e:
        print(f, "ERR", e)
PY
tail -n 2 gpurun_out/${TAG}_bench_n$N.err; tail -n 2 gpurun_out/${TAG}_bench_cifar50_gb128_n$N.err
