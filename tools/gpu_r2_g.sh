#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_ransac_gpu.py tests/test_castle_c2.py tests/test_l1k2_gpu.py -m gpu -q > $O/pytest_g.log 2>&1; echo "rc=$?" >> $O/pytest_g.log
timeout 600 python tools/bench_ransac.py 2000 4000 64 > $O/ransac.json 2> $O/ransac.err
timeout 600 python tools/bench_ransac.py 2000 60000 32 > $O/ransac_big.json 2>> $O/ransac.err
tail -4 $O/pytest_g.log; cat $O/ransac.json; echo; cat $O/ransac_big.json; tail -2 $O/ransac.err
