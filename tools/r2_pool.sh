#!/bin/bash
# Redirected end maps + side stream pool (no edge launches): parity, timeline, step time with and without graphs.
O=gpurun_out/${1:-pool}; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest.log | cut -c1-200
export ASSET_B200_FUSED_PIECES=2
ASSET_B200_TIMELINE=1 timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 3 --warmup 2 > $O/tl.json 2> $O/tl.err; grep timeline $O/tl.err | tail -5
run() {
  tag=$1; shift
  timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 20 --warmup 5 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:28s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms  eval_frac {j['roofline']['eval_frac']:.3f}")
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
for n in 200000 50000 10000; do
  run graph_$n --segments $n
  ASSET_B200_GRAPH=0 run eager_$n --segments $n
done
ASSET_B200_NO_PIECE_OVERLAP=1 run graph_200000_pieces_one_stream
ASSET_B200_FUSED_PIECES=4 run graph_200000_p4
ASSET_B200_FUSED_PIECES=3 run graph_200000_p3
run graph_lgl7_100k --mode LGL7 --segments 100000
