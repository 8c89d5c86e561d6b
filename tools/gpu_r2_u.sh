#!/bin/bash
# K1 slow path A/B: old = 64 unrolled offer sites, new = row loop through a per-lane scratch line
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_l1k2_gpu.py tests/test_castle_c2.py tests/test_ratio_gpu.py -m gpu -q -x > $O/u_pytest.log 2>&1; echo "rc=$?" >> $O/u_pytest.log; tail -3 $O/u_pytest.log
: > $O/u_k1_slowpath_ab.txt
for rep in 1 2; do for v in new2 new3; do
  echo "== $v" >> $O/u_k1_slowpath_ab.txt
  SPECTAVI_B200_LIB=$PWD/build/libspectavi_k1_$v.so timeout 300 python tools/castle_k1.py 2>&1 | grep -v "^ *[0-9]* *[0-9.]* *[0-9.]* *[0-9.]*$\|sweep" >> $O/u_k1_slowpath_ab.txt
  SPECTAVI_B200_LIB=$PWD/build/libspectavi_k1_$v.so timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-extras 2>/dev/null | python -c "
import json,sys
l=json.loads(sys.stdin.read().strip().split(chr(10))[-1]); print('C3 value %.2f ms %.3f frac %.4f e2e %.2f' % (l['value'], l['ms_per_step'], l['roofline']['frac'] or 0, l['e2e']['value']))" >> $O/u_k1_slowpath_ab.txt
done; done
cat $O/u_k1_slowpath_ab.txt
