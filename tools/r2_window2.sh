#!/bin/bash
O=gpurun_out/${1:-window2}; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_window.py -m gpu -x -q > $O/pytest_window.log 2>&1; echo "window pytest rc=$?"; tail -n 15 $O/pytest_window.log | cut -c1-220
timeout 1500 python -m pytest tests/test_gpu_fused.py tests/test_gpu_parity.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest.log | cut -c1-200
export ASSET_B200_FUSED_PIECES=2
export ASSET_B200_NO_PIECE_OVERLAP=1
run() {
  tag=$1; shift
  timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 20 --warmup 5 "$@" > $O/$tag.json 2> $O/$tag.err
  python - "$tag" "$O/$tag.json" <<'PY'
import json, sys
try:
    j = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(f"{sys.argv[1]:28s} {j['ms_per_step']:.3f} ms  {j['value']/1e6:8.1f} Mseg/s  fused {j['kernels_ms'].get('fused', 0):.3f} ms  eval_frac {j['roofline']['eval_frac']:.3f}")
    print("   ", {k: round(v, 3) for k, v in j['kernels_ms'].items()})
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
}
run graph_200000
ASSET_B200_GRAPH=0 run eager_200000
run graph_10000 --segments 10000
ASSET_B200_TIMELINE=1 timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 3 --warmup 2 > $O/tl.json 2> $O/tl.err; grep timeline $O/tl.err | tail -9
