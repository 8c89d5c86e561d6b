#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_cascade_gpu.py -m gpu -x -q 2>&1 | tail -3
python tools/bench_cascade.py 1000000 128 5 2000 > gpurun_out/cascade_c5.json 2> gpurun_out/cascade_c5.err; echo "rc=$?" >> gpurun_out/cascade_c5.err
python tools/bench_cascade.py 1000000 128 1 0 > gpurun_out/plain_c.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:probe_rerank -c 1 -f -o gpurun_out/k4_full python tools/bench_cascade.py 1000000 128 1 0 > gpurun_out/ncu_c.log 2>&1
cat gpurun_out/cascade_c5.json; tail -3 gpurun_out/cascade_c5.err
