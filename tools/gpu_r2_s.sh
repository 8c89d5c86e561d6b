#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then behind the whole GPU suite
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/s_noalloc_alone.log 2>&1; echo "rc=$?" >> $O/s_noalloc_alone.log
tail -15 $O/s_noalloc_alone.log
timeout 1800 python -m pytest tests -x -q -m gpu > $O/s_pytest_all.log 2>&1; echo "rc=$?" >> $O/s_pytest_all.log
tail -6 $O/s_pytest_all.log
