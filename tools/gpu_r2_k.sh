#!/bin/bash
# upload-lane sweep of the float entry points + the single-GPU entry points after the pinned download change
mkdir -p gpurun_out; O=gpurun_out
nproc > $O/k_host.txt; free -g >> $O/k_host.txt
timeout 900 python -m pytest tests/test_bruteforce_gpu.py tests/test_ransac_gpu.py tests/test_cascade_gpu.py tests/test_kmedians_gpu.py tests/test_quantize_gpu.py -m gpu -q -x > $O/k_pytest.log 2>&1; echo "rc=$?" >> $O/k_pytest.log
tail -5 $O/k_pytest.log
timeout 900 python tools/bench_e2e_lanes.py 1 2 4 6 8 > $O/k_lanes.jsonl 2> $O/k_lanes.err
cat $O/k_lanes.jsonl; tail -3 $O/k_lanes.err
timeout 600 python tools/bench_ransac.py > $O/k_ransac.json 2> $O/k_ransac.err; cut -c1-900 $O/k_ransac.json
