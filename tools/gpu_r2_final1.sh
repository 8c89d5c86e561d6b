#!/bin/bash
# round 2, closing 1-GPU pass: all parity tests, smoke, both bench arms, then the profiling pass (tools/gpu_r2_b.sh)
mkdir -p gpurun_out; O=gpurun_out
nvidia-smi -L > $O/gpus_n1.txt; lscpu | grep -E "Model name|^CPU\(s\)" > $O/host.txt
timeout 1800 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?" >> $O/bench_n1.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
bash tools/gpu_r2_b.sh > $O/r2b.log 2>&1
tail -4 $O/pytest_gpu.log; tail -3 $O/smoke.log; python -c "
import json; l=json.loads(open('$O/bench_n1.json').read().strip().split(chr(10))[-1]); print(l['value'], l['roofline']['frac'], l['e2e']['value']/l['value'], l['parity_bit_exact']); print({k:(v.get('error') or v.get('wall_s')) for k,v in l['extras'].items()})"
cat $O/bench_ref.json | cut -c1-300; ls $O
