#!/bin/bash
# 8-GPU box: C5 cascade sharded 8/4/2-way
mkdir -p gpurun_out; O=gpurun_out
for n in 8 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2966$n tools/bench_cascade_mgpu.py 1000000 128 5 > $O/cascade_n$n.json 2> $O/cascade_n$n.err; echo "rc=$?" >> $O/cascade_n$n.err
done
cat $O/cascade_n8.json $O/cascade_n2.json; tail -3 $O/cascade_n8.err
