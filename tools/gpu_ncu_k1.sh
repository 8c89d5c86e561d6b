#!/bin/bash
# ncu full capture of K1 (one launch) on the c3 workload; run the plain command first
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu --no-peak --no-e2e > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:l1_top2 -s 3 -c 1 -f -o gpurun_out/k1_full \
  python bench.py --steps 2 --warmup 3 --no-cpu --no-peak --no-e2e > gpurun_out/ncu_full.log 2>&1
cat gpurun_out/plain.log | tail -1 | cut -c1-300
