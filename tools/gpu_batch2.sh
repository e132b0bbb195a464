#!/bin/bash
O=gpurun_out/${1:-b2}; mkdir -p $O
./tools/icache_bench > $O/icache.txt 2>&1; cat $O/icache.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/pytest.log
timeout 600 python bench.py --no-cpu > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -c 1500 $O/bench.json
