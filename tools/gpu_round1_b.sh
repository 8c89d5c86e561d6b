#!/bin/bash
# full GPU pass: parity tests, smoke, bench (both arms), ncu launch list + full captures of K1, K2, K4
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/gpu_info.txt 2>&1
nproc > $O/host.txt; lscpu | grep -E 'Model name|^CPU\(s\)|Thread' | cut -c1-200 >> $O/host.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?" >> $O/bench_n1.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
B="python bench.py --steps 2 --warmup 3 --no-cpu --no-peak --no-e2e"
timeout 300 $B > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv $B > $O/ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:l1_top2 -s 3 -c 1 -f -o $O/k1_full $B > $O/ncu_full.log 2>&1
timeout 600 python tools/bench_cascade.py 1000000 128 5 2000 > $O/cascade_c5.json 2> $O/cascade_c5.err; echo "rc=$?" >> $O/cascade_c5.err
C="python tools/bench_cascade.py 1000000 128 1 0"
timeout 300 $C > $O/plain_c.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,lts__t_bytes.sum --clock-control none -k regex:"hash_project|bucket|scan|probe_rerank|dict_image|f32_to_u8|keys" -c 40 --csv --log-file $O/cascade_launches.csv $C > $O/ncu_c.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"hash_project|probe_rerank" -c 3 -f -o $O/cascade_full $C > $O/ncu_c_full.log 2>&1
tail -3 $O/pytest_gpu.log; tail -2 $O/smoke.log; cat $O/bench_n1.json; tail -2 $O/bench_n1.err; cat $O/bench_ref.json; cat $O/cascade_c5.json
