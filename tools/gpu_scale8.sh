#!/bin/bash
# 8-GPU box: multi-GPU parity tests, bench.py at N=8/4/2 (C4: 1M x 1M x 128, train sharded), C5 cascade sharded
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L > $O/gpus8.txt
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > $O/pytest_mgpu.log 2>&1; echo "rc=$?" >> $O/pytest_mgpu.log
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2965$n bench.py --gpus $n --steps 5 --warmup 3 --cpu-seconds 4 > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "rc=$?" >> $O/bench_n$n.err
done
for n in 8 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2966$n tools/bench_cascade_mgpu.py 1000000 128 5 > $O/cascade_n$n.json 2> $O/cascade_n$n.err; echo "rc=$?" >> $O/cascade_n$n.err
done
tail -3 $O/pytest_mgpu.log; for n in 8 4 2; do tail -1 $O/bench_n$n.json | cut -c1-200; done; cat $O/cascade_n8.json $O/cascade_n2.json | grep workload
