#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_ransac_gpu.py -m gpu -q -x > $O/l_pytest.log 2>&1; echo "rc=$?" >> $O/l_pytest.log
tail -4 $O/l_pytest.log
timeout 600 python tools/bench_ransac.py > $O/l_ransac.json 2> $O/l_ransac.err; cut -c1-700 $O/l_ransac.json
timeout 300 python tools/trace_lpk.py > $O/l_trace_lpk.txt 2>&1; cat $O/l_trace_lpk.txt
