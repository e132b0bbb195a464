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
run base            ASSET_B200_JIT_PREFETCH=0 ASSET_B200_JIT_HOIST=0
run prefetch        ASSET_B200_JIT_PREFETCH=1 ASSET_B200_JIT_HOIST=0
run hoist           ASSET_B200_JIT_PREFETCH=0 ASSET_B200_JIT_HOIST=1
run both            ASSET_B200_JIT_PREFETCH=1 ASSET_B200_JIT_HOIST=1
run both_t128       ASSET_B200_JIT_THREADS=128
run both_t128_lock  ASSET_B200_JIT_THREADS=128 ASSET_B200_JIT_LOCKSTEP=1
run both_t64_lock   ASSET_B200_JIT_THREADS=64 ASSET_B200_JIT_LOCKSTEP=1
run regs64          ASSET_B200_JIT_REGS=64
run regs80          ASSET_B200_JIT_REGS=80
run regs112         ASSET_B200_JIT_REGS=112
run t32             ASSET_B200_JIT_THREADS=32
run t32_regs80      ASSET_B200_JIT_THREADS=32 ASSET_B200_JIT_REGS=80
