#!/bin/bash
# Early gathers on the side stream, 24-bit offset groups, spill retry: parity, then the step time.
O=gpurun_out/${1:-sweep4}; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_fused.py tests/test_gpu_parity.py tests/test_gpu_multi.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest.log | cut -c1-200
run() {
  tag=$1; shift
  ASSET_B200_JIT_VERBOSE=1 timeout 600 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:28s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms  eval_frac {j['roofline']['eval_frac']:.3f}")
    print("   ", {k: round(v, 3) for k, v in j['kernels_ms'].items()})
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
  grep "ab_fused_kkt" $O/$tag.err | sed 's/.*ab_fused_kkt/  kkt/' | cut -c1-220 | head -8
}
export ASSET_B200_FUSED_PIECES=2
run p2
ASSET_B200_FUSED_KEEP_SPILLS=1 run p2_keep_spills
ASSET_B200_FUSED_PIECES=4 run p4
run p2_10k --segments 10000
run p2_50k --segments 50000
