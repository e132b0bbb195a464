#!/bin/bash
# CUDA graph replay of the evaluation: parity, then step time with and without graphs at several mesh sizes; store
# microbenchmark with the pieces' partial regions.
O=gpurun_out/${1:-graph}; mkdir -p $O
(tools/build/direct_store_bench 200000 586 910 7 | grep -E "nseg|I v4 ilp4"; tools/build/direct_store_bench 200000 324 910 5 | grep -E "nseg|I v4 ilp4") > $O/direct_store_partial.txt 2>&1; cat $O/direct_store_partial.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest.log | cut -c1-200
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
export ASSET_B200_FUSED_PIECES=2
for n in 200000 50000 10000; do
  run graph_$n --segments $n
  ASSET_B200_GRAPH=0 run eager_$n --segments $n
done
ASSET_B200_NO_PIECE_OVERLAP=1 run graph_200000_pieces_one_stream
