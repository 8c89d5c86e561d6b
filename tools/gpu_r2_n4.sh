#!/bin/bash
# N = 4 bench line only
mkdir -p gpurun_out; O=gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29654 bench.py --gpus 4 --steps 5 --warmup 3 --cpu-seconds 4 > $O/bench_n4.json 2> $O/bench_n4.err; echo "rc=$?" >> $O/bench_n4.err; tail -2 $O/bench_n4.err
