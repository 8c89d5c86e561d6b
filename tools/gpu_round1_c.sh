#!/bin/bash
# closing pass after the K1f / 8f-rank-3 work: all parity tests, smoke, both bench arms, K1f + k-medians launch lists
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?" >> $O/bench_n1.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 600 python tools/bench_kmedians.py 10000 132 10 > $O/kmedians_c2.json 2> $O/kmedians_c2.err
timeout 600 python tools/bench_lp.py 20000 5 > $O/lp_20k.json 2> $O/lp_20k.err
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active
timeout 600 ncu --metrics $M --clock-control none -k regex:"lp_topk|lp_merge" -c 8 --csv --log-file $O/kmedians_launches.csv python tools/bench_kmedians.py 10000 132 1 > $O/ncu_km.log 2>&1
timeout 600 ncu --metrics $M --clock-control none -k regex:"lp_topk|lp_merge" -c 6 --csv --log-file $O/lp_launches.csv python tools/bench_lp.py 20000 1 > $O/ncu_lp.log 2>&1
tail -3 $O/pytest_gpu.log; tail -3 $O/smoke.log; cat $O/bench_n1.json; tail -2 $O/bench_n1.err; cat $O/bench_ref.json; cat $O/kmedians_c2.json
