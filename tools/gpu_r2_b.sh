#!/bin/bash
# round 2, profiling pass (1 GPU): each ncu command runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out; O=gpurun_out
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active
# 1. launch list of the headline command
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-extras"
timeout 600 $B > $O/plain_bench.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_c3.csv $B > $O/ncu_b.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:l1_top2_kernel -s 3 -c 1 -f -o $O/k1_full $B > $O/ncu_k1_full.log 2>&1
# 2. K1f: TMA-fed kernel, per-launch metrics + one full capture
L="python tools/bench_lp.py 20000 1"
timeout 600 $L > $O/lp_20k.json 2> $O/lp.err && {
  timeout 900 ncu --metrics $M --clock-control none -k regex:"lp_topk|lp_merge|transpose" -c 12 --csv --log-file $O/r02_lp_launches.csv $L > $O/ncu_lp.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"lp_topk_tma" -c 1 -f -o $O/k1f_tma_full $L > $O/ncu_lp_full.log 2>&1
}
# 3. cascade C5: per-launch metrics + full captures of K2 (train side) and K4
C="python tools/bench_cascade.py 1000000 128 1 0"
timeout 600 $C > $O/plain_c.log 2>&1 && {
  timeout 900 ncu --metrics $M --clock-control none -k regex:"hash_project|bucket|scan|probe_rerank|dict_image" -c 24 --csv --log-file $O/r02_cascade_launches_c5.csv $C > $O/ncu_c.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"hash_project_i8" -c 2 -f -o $O/k2_full $C > $O/ncu_k2_full.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"probe_rerank" -c 1 -f -o $O/k4_full $C > $O/ncu_k4_full.log 2>&1
}
# 4. K8
R="python tools/bench_ransac.py 2000 4000 32"
timeout 600 $R > $O/ransac.json 2> $O/ransac.err && {
  timeout 900 ncu --metrics $M --clock-control none -k regex:"ransac" -c 9 --csv --log-file $O/r02_ransac_launches.csv $R > $O/ncu_r.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"ransac_score" -s 1 -c 1 -f -o $O/k8_full $R > $O/ncu_k8_full.log 2>&1
}
for f in k1_full k1f_tma_full k2_full k4_full k8_full; do
  [ -f $O/$f.ncu-rep ] && ncu -i $O/$f.ncu-rep --page details --print-units base 2>/dev/null | grep -E "^  [a-z_0-9]+.*\(|Duration|Throughput|Pipe|Issue|Registers|Stall|dram__bytes|Hit Rate|Achieved Occupancy|Theoretical Occupancy|Executed Ipc|L2 Cache|DRAM|Shared Memory|Bank" > $O/r02_${f}_summary.txt
done
cat $O/lp_20k.json | cut -c1-1500; tail -2 $O/lp.err; cat $O/ransac.json | cut -c1-1200; tail -1 $O/plain_c.log | cut -c1-1500
