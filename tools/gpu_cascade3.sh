#!/bin/bash
# K4/K6 iteration: parity tests, C5 bench, per-launch ncu metrics + one full capture of K4
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_cascade_gpu.py tests/test_ratio_gpu.py -m gpu -x -q 2>&1 | tail -15 > $O/pytest_c.log
timeout 600 python tools/bench_cascade.py 1000000 128 5 2000 > $O/cascade_c5.json 2> $O/cascade_c5.err; echo "rc=$?" >> $O/cascade_c5.err
C="python tools/bench_cascade.py 1000000 128 1 0"
timeout 300 $C > $O/plain_c.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,lts__t_bytes.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"hash_project|bucket|scan|probe_rerank|dict_image" -c 24 --csv --log-file $O/cascade_launches.csv $C > $O/ncu_c.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"probe_rerank" -c 1 -f -o $O/k4_full $C > $O/ncu_c_full.log 2>&1
cat $O/pytest_c.log; cat $O/cascade_c5.json; tail -3 $O/cascade_c5.err
