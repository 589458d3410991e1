#!/bin/bash
# step-time sweep of the BN apply kernels' launch-shape knobs (ResNet-50 / 224 / batch 128)
run() {
  env "$@" timeout 200 python bench.py --workload imagenet50 --steps 30 --warmup 5 --no-cpu-baseline --sustained-steps 0 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k={r['kernel']:r['ms'] for r in d['kernels']}
print('$*', d['ms_per_step'], 'bn_apply', k.get('bn_apply'), 'bn_bwd_apply', k.get('bn_bwd_apply'))"
}
run X=0
for v in 2 3 6 8; do run RN_BN_APPLY_WAVES=$v; done
for v in 2 4 6; do run RN_BN_BWD_APPLY_WAVES=$v; done
for v in 8 32 64; do run RN_BN_BWD_APPLY_KB=$v; done
for v in 8 32; do run RN_BN_APPLY_CW=$v; done
run X=0
