#!/bin/bash
# 8-GPU closing pass: the multi-GPU tests, then the N = 8 bench line only (N = 4 is not repeated: GPU-minutes)
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L > $O/gpus_n8.txt
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > $O/pytest_mgpu_n8.log 2>&1; echo "rc=$?" >> $O/pytest_mgpu_n8.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29658 bench.py --gpus 8 --steps 5 --warmup 3 --cpu-seconds 4 > $O/bench_n8.json 2> $O/bench_n8.err; echo "rc=$?" >> $O/bench_n8.err
tail -3 $O/pytest_mgpu_n8.log; tail -2 $O/bench_n8.err
