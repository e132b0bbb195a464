#!/bin/bash
O=gpurun_out/${1:-direct3}; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_fused.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/pytest.log | cut -c1-200
run() {
  tag=$1; shift
  ASSET_B200_JIT_VERBOSE=1 timeout 600 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:28s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms  eval_frac {j['roofline']['eval_frac']:.3f}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
export ASSET_B200_FUSED_PIECES=2
ASSET_B200_JIT_VLOAD=0 run p2_novload
run p2_vload
ASSET_B200_JIT_STCS=1 run p2_vload_cs
ASSET_B200_FUSED_BALANCE=1 run p2_vload_balance
ASSET_B200_FUSED_PIECES=3 ASSET_B200_FUSED_BALANCE=1 run p3_vload_balance
ASSET_B200_FUSED_PIECES=3 run p3_vload
ASSET_B200_FUSED_PIECES=4 ASSET_B200_FUSED_BALANCE=1 run p4_vload_balance
