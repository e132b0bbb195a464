#!/bin/bash
O=gpurun_out/${1:-suite}; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 30 $O/pytest.log | cut -c1-220
ASSET_B200_JIT_VERBOSE=1 timeout 600 python bench.py --no-cpu --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
grep fused $O/bench.err | tail -3
python - $O/bench.json <<'PY'
import json, sys
j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("step", j["ms_per_step"], "value", j["value"], "e2e", j["e2e"]["value"]); print({k: round(v, 4) for k, v in j["kernels_ms"].items()})
PY
