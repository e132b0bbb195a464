#!/bin/bash
# Round-2 probe: DFMA pipe microbenchmark, two-piece direct-store kernels with/without volatile parks (measurement), ncu
# --set full capture of the two-piece kernels.
O=gpurun_out/${1:-probe}; mkdir -p $O
tools/build/dfma_bench > $O/dfma_bench.txt 2>&1; cat $O/dfma_bench.txt
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
run p2
ASSET_B200_JIT_VOLATILE=0 run p2_novolatile
ASSET_B200_JIT_DEBUG=1 run p2_nostores
ASSET_B200_FUSED_REGS=120 run p2_r120
timeout 600 ncu --set full --clock-control none --import-source on -k regex:ab_fused -c 4 -o $O/fused_p2 -f \
    python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e > $O/ncu_full.log 2>&1; echo "ncu full rc=$?"
