#!/bin/bash
# Device-resident evalKKT rate of the other BASELINE.json configurations on one GPU (no e2e / CPU legs).
O=gpurun_out/${1:-wl}; mkdir -p $O
run() {
  tag=$1; shift
  timeout 600 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    r = j["roofline"]
    print(f"{sys.argv[1]:32s} {j['config']['segments_total']:8d} seg  {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  eval_frac {r['eval_frac']:.3f}  top {r['kernel']} {r['kernel_ms']:.3f} ms")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run lowthrust_lgl7_100k   --ode lowthrust --mode LGL7 --segments 100000
run lowthrust_lgl3_200k   --ode lowthrust --mode LGL3 --segments 200000
run lowthrust_trap_200k   --ode lowthrust --mode Trapezoidal --segments 200000
run lowthrust_lgl5_10k    --ode lowthrust --mode LGL5 --segments 10000
run lowthrust_lgl5_30k    --ode lowthrust --mode LGL5 --segments 30000
run lowthrust_lgl5_100k   --ode lowthrust --mode LGL5 --segments 100000
run vanderpol_lgl5_1m     --ode vanderpol --mode LGL5 --segments 1000000
run hypersens_lgl5_1m     --ode hypersens --mode LGL5 --segments 1000000
