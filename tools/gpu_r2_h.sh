#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_ransac_gpu.py -m gpu -q > $O/pytest_g.log 2>&1; echo "rc=$?" >> $O/pytest_g.log
timeout 600 python tools/bench_ransac.py 2000 4000 64 > $O/ransac.json 2> $O/ransac.err
tail -3 $O/pytest_g.log; cat $O/ransac.json | cut -c1-400
