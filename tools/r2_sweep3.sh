#!/bin/bash
# Sector base registers, per-piece register budgets and output pacing for the two-piece direct-store kernels.
O=gpurun_out/${1:-sweep3}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_fused.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest.log | cut -c1-200
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
export ASSET_B200_FUSED_PIECES=2
run p2_sector_base
ASSET_B200_JIT_PHASE_BASES=0 run p2_no_sector_base
ASSET_B200_JIT_DEBUG=1 run p2_stores_never_execute
ASSET_B200_FUSED_REGS_BY_PIECE=112,96 run p2_r112_96
ASSET_B200_FUSED_REGS_BY_PIECE=112,80 run p2_r112_80
ASSET_B200_FUSED_REGS_BY_PIECE=112,64 run p2_r112_64
ASSET_B200_FUSED_REGS_BY_PIECE=104,80 run p2_r104_80
ASSET_B200_FUSED_REGS_BY_PIECE=96,80 run p2_r96_80
ASSET_B200_JIT_PACE=1 run p2_pace
ASSET_B200_FUSED_PIECES=3 run p3_sector_base
ASSET_B200_FUSED_PIECES=4 run p4_sector_base
