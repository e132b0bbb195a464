#!/bin/bash
O=gpurun_out/${1:-fu}; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_fused.py -m gpu -x -q > $O/pytest_fused.log 2>&1; echo "fused pytest rc=$?"; tail -n 25 $O/pytest_fused.log
ASSET_B200_JIT_VERBOSE=1 timeout 600 python bench.py --no-cpu --no-e2e --steps 10 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
grep fused $O/bench.err | tail -2
python - $O/bench.json <<'PY'
import json, sys
j = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("step", j["ms_per_step"], "value", j["value"]); print({k: round(v, 4) for k, v in j["kernels_ms"].items()}); print(j["roofline"])
PY
