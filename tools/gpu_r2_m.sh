#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_bruteforce_gpu.py tests/test_ransac_gpu.py tests/test_kmedians_gpu.py tests/test_quantize_gpu.py -m gpu -q -x > $O/m_pytest.log 2>&1; echo "rc=$?" >> $O/m_pytest.log
tail -4 $O/m_pytest.log
timeout 300 python tools/trace_lpk.py > $O/m_trace_lpk.txt 2>&1; cat $O/m_trace_lpk.txt
timeout 600 python tools/bench_e2e_lanes.py 6 > $O/m_lanes.jsonl 2> $O/m_lanes.err; cat $O/m_lanes.jsonl
timeout 600 python tools/bench_ransac.py > $O/m_ransac.json 2> $O/m_ransac.err; cut -c1-400 $O/m_ransac.json
