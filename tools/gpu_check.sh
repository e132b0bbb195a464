#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench, ncu launch list, one --set full capture of the dominant kernels (with
# the fp64-pipe counters), reference arm.
# usage: gpurun --timeout 2400 -- 'bash tools/gpu_check.sh TAG [KERNEL_REGEX]'
TAG=${1:-run}; KRE=${2:-ab_fused}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/smoke.log
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
tail -c 3500 $O/bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $O/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-parity > $O/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --metrics smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active \
    --clock-control none --import-source on -k regex:$KRE -c 8 -o $O/top_kernel -f \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-parity > $O/ncu_full.log 2>&1; echo "ncu full rc=$?"
tail -q -n 3 $O/pytest.log $O/smoke.log
timeout 900 python bench.py --impl reference --steps 5 --warmup 2 > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference arm rc=$?"; cat $O/bench_reference.json | cut -c1-1500
