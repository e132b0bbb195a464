#!/bin/bash
# Round-end measurement in one gpurun call: piece-split A/B, workload table, then the full check (tests, smoke, bench,
# ncu launch list and --set full capture, reference arm).
TAG=${1:-final}
O=gpurun_out/$TAG; mkdir -p $O
for v in even balance; do
  if [ $v = balance ]; then export ASSET_B200_FUSED_BALANCE=1; fi
  timeout 300 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 > $O/split_$v.json 2> $O/split_$v.err
  python -c "import json,sys; j=json.loads(open('$O/split_$v.json').read().strip().splitlines()[-1]); print('split $v step', j['ms_per_step'], 'fused', j['kernels_ms'].get('fused'))"
done
unset ASSET_B200_FUSED_BALANCE
bash tools/workload_table.sh $TAG/wl
bash tools/gpu_check.sh $TAG/check "ab_fused"
