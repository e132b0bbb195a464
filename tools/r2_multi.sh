#!/bin/bash
# N-GPU bench of one phase with rank-local analysis (windows): usage r2_multi.sh TAG NGPU [SEGMENTS]
O=gpurun_out/${1:-multi}; N=${2:-2}; SEG=${3:-200000}; mkdir -p $O

( time timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --segments $SEG ) > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?"
tail -n 1 $O/bench_n$N.json | cut -c1-2500; tail -n 6 $O/bench_n$N.err | cut -c1-300
