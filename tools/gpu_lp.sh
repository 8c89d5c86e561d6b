#!/bin/bash
# K1f iteration: parity (generic matcher + filter overload), timing, launch list
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_kmedians_gpu.py tests/test_bruteforce_gpu.py -m gpu -x -q 2>&1 | tail -8 > $O/pytest_lp.log
timeout 600 python tools/bench_lp.py 20000 5 > $O/lp_20k.json 2> $O/lp_20k.err
timeout 600 python tools/bench_kmedians.py 10000 132 10 > $O/kmedians_c2.json 2> $O/kmedians_c2.err
timeout 600 ncu --metrics gpu__time_duration.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k regex:"lp_topk" -c 3 --csv --log-file $O/lp_launches.csv python tools/bench_lp.py 20000 1 > $O/ncu_lp.log 2>&1
cat $O/pytest_lp.log; cat $O/lp_20k.json; cat $O/kmedians_c2.json; grep lp_topk $O/lp_launches.csv | cut -d, -f5,13- | head -20
