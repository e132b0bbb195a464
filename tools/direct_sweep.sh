#!/bin/bash
# Fused kernels with direct 32-byte stores: parity first, then piece count x register budget (device-resident evalKKT).
O=gpurun_out/${1:-direct}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 8 $O/pytest.log | cut -c1-200
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
  grep "ab_fused_kkt" $O/$tag.err | sed 's/.*ab_fused_kkt/  kkt/' | cut -c1-200 | head -8
}
ASSET_B200_FUSED_DIRECT=0 run tile_p4
for np in 1 2 3 4; do
  for regs in 112 100; do
    ASSET_B200_FUSED_PIECES=$np ASSET_B200_FUSED_REGS=$regs run direct_p${np}_r$regs
  done
done
