#!/bin/bash
# last sanity of the shipped tree: the GPU suite with -x (as the round's end runs it), smoke, and the default bench line
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > $O/last_pytest.log 2>&1; echo "rc=$?" >> $O/last_pytest.log; tail -3 $O/last_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/last_smoke.log 2>&1; echo "smoke rc=$?" >> $O/last_smoke.log; tail -2 $O/last_smoke.log
timeout 300 python bench.py --no-extras > $O/last_bench.json 2> $O/last_bench.err; echo "bench rc=$?"; cut -c1-330 $O/last_bench.json
