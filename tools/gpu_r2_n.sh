#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 120 python tools/k1_cold_vs_seeded.py > $O/n_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:l1_top2_kernel -c 2 -f -o $O/k1_cold_seeded_new python tools/k1_cold_vs_seeded.py > $O/n_ncu.log 2>&1
tail -2 $O/n_ncu.log
