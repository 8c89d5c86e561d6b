#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_l1k2_gpu.py tests/test_castle_c2.py tests/test_ratio_gpu.py -m gpu -q -x > $O/n_pytest.log 2>&1; echo "rc=$?" >> $O/n_pytest.log; tail -3 $O/n_pytest.log
for w in 0 1; do echo "SPECTAVI_K1_WARM=$w"; SPECTAVI_K1_WARM=$w timeout 300 python tools/castle_k1.py 2>&1; done > $O/n_castle_k1_warm.txt; cat $O/n_castle_k1_warm.txt
