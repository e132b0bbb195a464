#!/bin/bash
# Residual-only fused modes: parity, then step times of evalRHS / OGC / OCC / OBJ (fused and per-function); block-constant
# control; guard-band test.
O=gpurun_out/${1:-modes}; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_fused.py tests/test_gpu_guards.py tests/test_gpu_parity.py -m gpu -x -q -k "residual or guards or every_eval_mode or no_kernel" > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 12 $O/pytest.log | cut -c1-250
run() {
  tag=$1; shift
  timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 20 --warmup 5 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:34s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  eval_frac {j['roofline']['eval_frac']:.3f}  ", {k: round(v, 3) for k, v in j['kernels_ms'].items() if v > 0.02})
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for m in RHS OGC OCC OBJ; do
  run fused_$m --eval $m
  ASSET_B200_NO_FUSED_RESIDUAL=1 run perfunction_$m --eval $m
done
for m in KKTNO SOE AUG; do run perfunction_$m --eval $m; done
run block_constant_kkt --control-mode BlockConstant
