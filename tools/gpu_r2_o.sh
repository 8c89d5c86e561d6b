#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
: > $O/o_k8_variants.txt
for v in base cs smemp smemp5 b3 smemp3 b5; do
  for rep in 1 2; do
    SPECTAVI_B200_LIB=$PWD/build/libspectavi_k8_$v.so timeout 300 python tools/bench_ransac.py 2000 4000 16 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v', 'device_ms %.4f' % d['gpu_device_ms'], 'parity', d.get('parity_bit_exact'))" >> $O/o_k8_variants.txt
  done
done
cat $O/o_k8_variants.txt
