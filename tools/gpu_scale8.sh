#!/bin/bash
# 8-GPU box: multi-GPU parity tests, then bench.py at N=8 and N=4 (C4: 1M x 1M x 128, train sharded)
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L > $O/gpus8.txt
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > $O/pytest_mgpu.log 2>&1; echo "rc=$?" >> $O/pytest_mgpu.log
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus $n --steps 5 --warmup 3 --cpu-seconds 5 > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "rc=$?" >> $O/bench_n$n.err
done
tail -3 $O/pytest_mgpu.log; cat $O/bench_n8.json $O/bench_n4.json; tail -2 $O/bench_n8.err
