#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_bruteforce_gpu.py -m gpu -q > $O/pytest_mu.log 2>&1; echo "rc=$?" >> $O/pytest_mu.log
tail -15 $O/pytest_mu.log
