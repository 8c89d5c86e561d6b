#!/bin/bash
# 8f rank 3 iteration: parity tests (new + the generic matcher's, which shares the kernel), bench, launch list
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_kmedians_gpu.py tests/test_bruteforce_gpu.py -m gpu -x -q 2>&1 | tail -15 > $O/pytest_km.log
timeout 600 python tools/bench_kmedians.py 10000 132 10 > $O/kmedians_c2.json 2> $O/kmedians_c2.err; echo "rc=$?" >> $O/kmedians_c2.err
C="python tools/bench_kmedians.py 10000 132 1"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"lp_topk|lp_merge" -c 12 --csv --log-file $O/kmedians_launches.csv $C > $O/ncu_km.log 2>&1
cat $O/pytest_km.log; cat $O/kmedians_c2.json; tail -3 $O/kmedians_c2.err; grep -c lp_ $O/kmedians_launches.csv
