#!/bin/bash
# usage: gpurun --gpus N -- 'bash tools/gpu_multi.sh TAG N'
O=gpurun_out/${1:-multi}; N=${2:-2}; mkdir -p $O
free -g | head -2 > $O/host.txt; nproc >> $O/host.txt; cat $O/host.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench N=$N rc=$?"; tail -c 2500 $O/bench_n$N.json; tail -n 5 $O/bench_n$N.err
[ "${3:-phases}" = "phases" ] && timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --shard phases --no-e2e > $O/bench_phases_n$N.json 2> $O/bench_phases_n$N.err; echo "bench phases rc=$?"; tail -c 600 $O/bench_phases_n$N.json
true
