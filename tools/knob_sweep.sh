#!/bin/bash
# Sweep of the kernel specialiser's knobs on one GPU: prints ms_per_step and the bulk kernels' device times per variant.
O=gpurun_out/${1:-sweep}; mkdir -p $O
run() {
  tag=$1; shift
  env "$@" timeout 300 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    k = j["kernels_ms"]
    print(f"{sys.argv[1]:28s} step {j['ms_per_step']:.3f} ms  defect {k.get('defect',0):.3f}  spline {k.get('control_spline',0):.3f}  gatherKKT {k.get('gather:KKT',0):.3f}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run zrun4_default   ASSET_B200_X=0
run zrun1           ASSET_B200_JIT_ZRUN=1
run zrun2           ASSET_B200_JIT_ZRUN=2
run zrun8           ASSET_B200_JIT_ZRUN=8
run zrun16          ASSET_B200_JIT_ZRUN=16
run zrun1000        ASSET_B200_JIT_ZRUN=1000
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
