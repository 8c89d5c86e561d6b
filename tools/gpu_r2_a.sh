#!/bin/bash
# round 2, first pass: parity of the new host path (streamed uploads, seeded thresholds), wide rows, k > 128, then a bench line
mkdir -p gpurun_out; O=gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?" >> $O/bench_n1.err
tail -15 $O/pytest_gpu.log; tail -3 $O/smoke.log; cat $O/bench_n1.json; tail -2 $O/bench_n1.err
