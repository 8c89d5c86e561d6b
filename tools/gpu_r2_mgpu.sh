#!/bin/bash
# round 2, multi-GPU pass: usage  gpu_r2_mgpu.sh N   (N = 2, 4 or 8 GPUs visible)
N=${1:-2}
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L > $O/gpus_n$N.txt
timeout 1200 python -m pytest tests/test_multigpu.py -m gpu -x -q > $O/pytest_mgpu_n$N.log 2>&1; echo "rc=$?" >> $O/pytest_mgpu_n$N.log
for n in $N $(( N / 2 )); do
  [ $n -lt 2 ] && continue
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2965$n bench.py --gpus $n --steps 5 --warmup 3 --cpu-seconds 4 > $O/bench_n$n.json 2> $O/bench_n$n.err; echo "rc=$?" >> $O/bench_n$n.err
done
tail -3 $O/pytest_mgpu_n$N.log; for n in 8 4 2; do [ -f $O/bench_n$n.json ] && { tail -1 $O/bench_n$n.json | cut -c1-2500; tail -2 $O/bench_n$n.err; }; done
