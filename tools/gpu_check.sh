#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, ncu launch list, one --set full capture of the dominant kernel.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_check.sh TAG [KERNEL_REGEX]'
TAG=${1:-run}; KRE=${2:-ab_seg_f0_m0}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
tail -c 3000 $O/bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $O/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$KRE -c 16 -o $O/top_kernel -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > $O/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -q -n 3 $O/pytest.log $O/smoke.log
timeout 600 python bench.py --impl reference > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference arm rc=$?"; cat $O/bench_reference.json
