#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 300 python tools/trace_e2e.py > $O/trace_e2e.log 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?" >> $O/bench_n1.err
bash tools/gpu_r2_b.sh > $O/r2b.log 2>&1
cat $O/trace_e2e.log | grep -v "piece"; python -c "
import json; l=json.loads(open('$O/bench_n1.json').read().strip().split(chr(10))[-1]); print(l['value'], l['roofline']['frac'], l['e2e']['ms_per_step'], l['ms_per_step'], l['extras']['c2_castle']['nn_bruteforcel1k2_ms_e2e'])"
tail -30 $O/r2b.log | cut -c1-400; ls -la $O
