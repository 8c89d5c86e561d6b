#!/bin/bash
# same-box A/B of cascade variants: build/libspectavi_k2_<v>.so
mkdir -p gpurun_out; O=gpurun_out
: > $O/r_k2_variants.txt
for rep in 1 2; do for v in occ8 occ9 occ10; do
  SPECTAVI_B200_LIB=$PWD/build/libspectavi_k2_$v.so timeout 300 python tools/bench_cascade.py 1000000 128 5 0 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); st=d['stages']; print('$v', 'cascade_ms %.4f' % d['cascade_ms'], 'K2train %.4f' % st['K2_train(projection + codes)']['ms'], 'K2query %.4f' % st['K2_query(+ least-confident bits)']['ms'], 'K4 %.4f' % st['K4(probe + gather + re-rank)']['ms'])" >> $O/r_k2_variants.txt
done; done
cat $O/r_k2_variants.txt
