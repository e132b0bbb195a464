#!/bin/bash
O=gpurun_out/${1:-slots}; mkdir -p $O
run() {
  tag=$1; shift
  ASSET_B200_JIT_VERBOSE=1 timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 10 --warmup 3 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:34s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms  eval_frac {j['roofline']['eval_frac']:.3f}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for sl in 160 192 224 256 288; do ASSET_B200_FUSED_SLOTS=$sl run lgl7_100k_slots$sl --mode LGL7 --segments 100000; done
ASSET_B200_FUSED_PIECES=8 run lgl7_100k_even8 --mode LGL7 --segments 100000
ASSET_B200_FUSED_PIECES=6 ASSET_B200_FUSED_BALANCE=1 run lgl7_100k_balance6 --mode LGL7 --segments 100000
for sl in 192 224 256; do ASSET_B200_FUSED_SLOTS=$sl run lgl5_200k_slots$sl; done
ASSET_B200_FUSED_SLOTS=256 run lgl7_block_100k --mode LGL7 --segments 100000
grep "ab_fused_kkt" $O/lgl7_100k_slots224.err | sed 's/.*ab_fused_kkt/  kkt/' | cut -c1-200
