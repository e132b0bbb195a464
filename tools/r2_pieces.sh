#!/bin/bash
# Piece count of the fused kernels with direct stores, over the workloads of BASELINE.json's configurations.
O=gpurun_out/${1:-pieces}; mkdir -p $O
run() {
  tag=$1; shift
  timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 10 --warmup 3 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:34s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms  eval_frac {j['roofline']['eval_frac']:.3f}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for np in 2 3 4 6; do ASSET_B200_FUSED_PIECES=$np run lowthrust_lgl7_100k_p$np --mode LGL7 --segments 100000; done
for np in 1 2 3; do ASSET_B200_FUSED_PIECES=$np run lowthrust_lgl3_200k_p$np --mode LGL3; done
for np in 1 2; do ASSET_B200_FUSED_PIECES=$np run lowthrust_trap_200k_p$np --mode Trapezoidal; done
for np in 1 2; do ASSET_B200_FUSED_PIECES=$np run vanderpol_lgl5_200k_p$np --ode vanderpol; done
for np in 1 2; do ASSET_B200_FUSED_PIECES=$np run hypersens_lgl5_200k_p$np --ode hypersens; done
for np in 1 2 3; do ASSET_B200_FUSED_PIECES=$np run lowthrust_lgl5_200k_p$np; done
