#!/bin/bash
mkdir -p gpurun_out
python tools/bench_cascade.py 1000000 128 5 2000 > gpurun_out/cascade_c5.json 2> gpurun_out/cascade_c5.err; echo "rc=$?" >> gpurun_out/cascade_c5.err
python tools/bench_cascade.py 200000 128 2 0 > gpurun_out/plain_c.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active --clock-control none -k regex:"hash_project|bucket|scan|probe_rerank|dict_image" -c 40 --csv --log-file gpurun_out/cascade_launches.csv python tools/bench_cascade.py 1000000 128 1 0 > gpurun_out/ncu_c.log 2>&1
cat gpurun_out/cascade_c5.json; tail -3 gpurun_out/cascade_c5.err
