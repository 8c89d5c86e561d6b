#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
: > $O/r_k4_variants.txt
for rep in 1 2; do
for v in PF1 PF2 PF4 PF8; do
  SPECTAVI_B200_LIB=$PWD/build/libspectavi_k4_$v.so timeout 300 python tools/bench_cascade.py 1000000 128 5 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', 'cascade_ms %.4f' % d['cascade_ms'], 'K4 %.4f' % d['stages']['K4(probe + gather + re-rank)']['ms'])" >> $O/r_k4_variants.txt
done
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.mem,temperature.gpu,power.draw --format=csv >> $O/r_k4_variants.txt
cat $O/r_k4_variants.txt
