#!/bin/bash
# One training step of ResNet-50 / 224 / batch 128 under ncu: per-launch duration, DRAM bytes, tensor-pipe activity,
# L2 -> SM bytes and L2 throughput for EVERY kernel launch of the step (un-captured launch list).
set -e
TAG=${1:-r02}
python scripts/one_step.py imagenet50 3 > gpurun_out/${TAG}_plain.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,l1tex__m_xbar2l1tex_read_bytes.sum,lts__throughput.avg.pct_of_peak_sustained_elapsed,sm__inst_executed.avg.per_cycle_elapsed \
    --profile-from-start off --clock-control none --csv --log-file gpurun_out/${TAG}_ncu_step.csv python scripts/one_step.py imagenet50 2 > gpurun_out/${TAG}_ncu.log 2>&1
tail -2 gpurun_out/${TAG}_plain.log
