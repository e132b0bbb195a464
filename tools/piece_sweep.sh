#!/bin/bash
# Piece count of the fused kernels on the other workloads (device-resident evalKKT, one GPU).
O=gpurun_out/${1:-pieces}; mkdir -p $O
run() {
  tag=$1; np=$2; shift 2
  ASSET_B200_FUSED_PIECES=$np timeout 600 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 "$@" > $O/${tag}_p$np.json 2> $O/${tag}_p$np.err
  python - "$tag pieces=$np" "$O/${tag}_p$np.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:34s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for np in 4 6 8; do run lowthrust_lgl7_100k $np --ode lowthrust --mode LGL7 --segments 100000; done
for np in 1 2 3; do run lowthrust_lgl3_200k $np --ode lowthrust --mode LGL3 --segments 200000; done
for np in 1 2; do run lowthrust_trap_200k $np --ode lowthrust --mode Trapezoidal --segments 200000; done
for np in 1 2; do run vanderpol_lgl5_1m $np --ode vanderpol --mode LGL5 --segments 1000000; done
for np in 1 2; do run hypersens_lgl5_1m $np --ode hypersens --mode LGL5 --segments 1000000; done
