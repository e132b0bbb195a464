#!/bin/bash
O=gpurun_out/${1:-timeline}; mkdir -p $O
export ASSET_B200_FUSED_PIECES=2
for n in 200000 10000; do
  ASSET_B200_TIMELINE=1 timeout 600 python bench.py --no-cpu --no-e2e --no-parity --steps 3 --warmup 2 --segments $n > $O/tl_$n.json 2> $O/tl_$n.err
  echo "segments $n"; grep timeline $O/tl_$n.err | tail -16
done
free -g | head -2; nproc
( time timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -m gpu -x -q -k "full_size or world8" ) > $O/pytest_new.log 2>&1; echo "pytest rc=$?"; tail -n 12 $O/pytest_new.log | cut -c1-250
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -c 1500 $O/bench.json; tail -3 $O/bench.err
