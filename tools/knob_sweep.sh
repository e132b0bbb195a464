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
run fz_default       ASSET_B200_JIT_VERBOSE=1
run fz_zrun1000      ASSET_B200_FUSED_ZRUN=1000
run fz_zrun4         ASSET_B200_FUSED_ZRUN=4
run fz_t32           ASSET_B200_FUSED_THREADS=32 ASSET_B200_JIT_VERBOSE=1
run fz_regs96        ASSET_B200_FUSED_REGS=96
run fz_regs120       ASSET_B200_FUSED_REGS=120
run fz_t32_zrun1000  ASSET_B200_FUSED_THREADS=32 ASSET_B200_FUSED_ZRUN=1000
grep -h "ab_fused" $O/fz_default.err $O/fz_t32.err | cut -c1-260
