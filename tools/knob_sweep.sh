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
    print(f"{sys.argv[1]:28s} step {j['ms_per_step']:.3f} ms  fused {k.get('fused',0):.3f}  defect {k.get('defect',0):.3f}  spline {k.get('control_spline',0):.3f}  gatherKKT {k.get('gather:KKT',0):.3f}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run base             ASSET_B200_JIT_VERBOSE=1
run light96          ASSET_B200_FUSED_REGS_BY_PIECE=112,112,96,96
run light80          ASSET_B200_FUSED_REGS_BY_PIECE=112,112,80,80 ASSET_B200_JIT_VERBOSE=1
run light64          ASSET_B200_FUSED_REGS_BY_PIECE=112,112,64,64 ASSET_B200_JIT_VERBOSE=1
run light48          ASSET_B200_FUSED_REGS_BY_PIECE=112,112,48,48 ASSET_B200_JIT_VERBOSE=1
run heavy120         ASSET_B200_FUSED_REGS_BY_PIECE=120,120,112,112
run heavy104         ASSET_B200_FUSED_REGS_BY_PIECE=104,104,112,112
grep -h "ab_fused" $O/light80.err $O/light64.err $O/light48.err | cut -c1-260
