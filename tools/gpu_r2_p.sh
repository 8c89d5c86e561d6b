#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q > $O/p_pytest_mgpu.log 2>&1; echo "rc=$?" >> $O/p_pytest_mgpu.log
tail -6 $O/p_pytest_mgpu.log
for g in 1 2; do SPECTAVI_GPUS=$g timeout 600 python tools/bench_lp_mgpu.py 100000 >> $O/p_lp_mgpu.jsonl 2>> $O/p_lp_mgpu.err; done
cat $O/p_lp_mgpu.jsonl; tail -3 $O/p_lp_mgpu.err
