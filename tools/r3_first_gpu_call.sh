#!/bin/bash
# First GPU call after round 2: everything that was built after the round's GPU minutes were spent, in the order of
# least to most plan-side risk, then the measurements that are missing.  Run under gpurun from the repository root:
#   gpurun --timeout 2400 -- 'bash tools/r3_first_gpu_call.sh > gpurun_out/r3_first.log 2>&1'
set -x
python -m pytest tests/test_zz1_gpu_sliced_inputs.py tests/test_zz2_gpu_tables.py -q -m gpu
python -m pytest tests/test_zz3_gpu_partial_modes.py tests/test_zz4_gpu_blocked_fused.py -q -m gpu
# the default line, then the opt-in paths next to their per-function twins
python bench.py --steps 20 --warmup 5 > gpurun_out/r3_bench_default.json
for ev in KKTNO SOE AUG; do
  python bench.py --steps 20 --warmup 5 --eval $ev --no-cpu > gpurun_out/r3_bench_${ev}_per_function.json
  python bench.py --steps 20 --warmup 5 --eval $ev --fused-partial --no-cpu > gpurun_out/r3_bench_${ev}_fused.json
done
python bench.py --steps 20 --warmup 5 --control-mode BlockConstant --no-cpu > gpurun_out/r3_bench_blocked_per_function.json
python bench.py --steps 20 --warmup 5 --control-mode BlockConstant --fused-blocked --no-cpu > gpurun_out/r3_bench_blocked_fused.json
python bench.py --steps 10 --warmup 3 --ode minclimb --no-cpu > gpurun_out/r3_bench_minclimb.json
python bench.py --steps 10 --warmup 3 --segments 1000000 --no-cpu > gpurun_out/r3_bench_one_million_segments.json
# In a call of its own (gpurun --gpus 8): the north-star line -- one 1.6 M-segment phase over eight GPUs, rank-local analysis,
# sliced upload -- which has never completed (host analysis size, then the per-rank exit of the warm loop):
#   gpurun --gpus 8 --timeout 1500 -- 'python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
#       --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r3_bench_eight_gpus.json 2> gpurun_out/r3_bench_eight_gpus.err'
