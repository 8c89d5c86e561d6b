#!/bin/bash
# after the K4 list re-deal and the K1f changes: multi-GPU parity tests again + the sharded cascade, on N GPUs (arg 1)
N=${1:-2}
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L > $O/gpus_n$N.txt
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > $O/pytest_mgpu_n$N.log 2>&1; echo "rc=$?" >> $O/pytest_mgpu_n$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2966$N tools/bench_cascade_mgpu.py 1000000 128 5 > $O/cascade_n$N.json 2> $O/cascade_n$N.err; echo "rc=$?" >> $O/cascade_n$N.err
tail -3 $O/pytest_mgpu_n$N.log; cat $O/cascade_n$N.json; tail -2 $O/cascade_n$N.err
