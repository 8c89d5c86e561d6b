#!/bin/bash
# round 2: full GPU test suite (no -x), host timeline of the drop-in call, K1 A/B (seeded thresholds on / off), K1f A/B
mkdir -p gpurun_out; O=gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python tools/trace_e2e.py > $O/trace_e2e.log 2>&1
for v in "" build/libspectavi_noseed.so build/libspectavi_r1thr.so ""; do
  SPECTAVI_B200_LIB=${v:+$PWD/$v} timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-extras --no-e2e 2>/dev/null | python -c "import sys,json; l=json.loads(sys.stdin.readline()); print('K1 A/B lib=${v:-default}', l['value'], l['roofline']['k1_ms'], l['roofline']['frac'])" >> $O/k1_ab.log 2>&1
done
timeout 600 python tools/bench_lp.py 20000 5 > $O/lp_20k.json 2> $O/lp.err
timeout 600 python tools/bench_cascade.py 1000000 128 5 1000 > $O/cascade_c5.json 2> $O/cascade.err
tail -12 $O/pytest_gpu.log; cat $O/trace_e2e.log; cat $O/k1_ab.log; python -c "
import json; l=json.load(open('$O/lp_20k.json')); print({k:l[k] for k in ('p=1','p=2','roofline')})
c=json.load(open('$O/cascade_c5.json')); print(c['cascade_ms'], json.dumps(c['stages'])[:1800]); print(c.get('e2e')); print(c.get('cpu_baseline'))"
tail -3 $O/lp.err $O/cascade.err
