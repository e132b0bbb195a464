#!/bin/bash
# Where does the two-piece direct-store kernel's time go: stores never executed / L2-resident output stream / lanes of
# one alignment phase per warp (measurement knobs of jit.cpp, JitSpec::debug bits 1, 16, 32).
O=gpurun_out/${1:-probe2}; mkdir -p $O
run() {
  tag=$1; shift
  timeout 600 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:28s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms  eval_frac {j['roofline']['eval_frac']:.3f}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for np in 2 4; do
export ASSET_B200_FUSED_PIECES=$np
run p${np}
ASSET_B200_JIT_DEBUG=1 run p${np}_stores_never_execute
ASSET_B200_JIT_DEBUG=16 run p${np}_l2_resident_outputs
ASSET_B200_JIT_DEBUG=32 run p${np}_phase_lanes
ASSET_B200_JIT_DEBUG=48 run p${np}_phase_lanes_l2_resident
done
