#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_bruteforce_gpu.py tests/test_kmedians_gpu.py -m gpu -q > $O/pytest_lp.log 2>&1; echo "rc=$?" >> $O/pytest_lp.log
L="python tools/bench_lp.py 20000 5"
timeout 600 $L > $O/lp_20k.json 2> $O/lp.err && timeout 900 ncu --set full --clock-control none --import-source on -k regex:"lp_topk_tma" -c 1 -f -o $O/k1f_tma_full python tools/bench_lp.py 20000 1 > $O/ncu_lp_full.log 2>&1
tail -4 $O/pytest_lp.log; python -c "
import json; l=json.load(open('$O/lp_20k.json')); print({k:l[k] for k in ('p=1','p=2')}); print(l['roofline'])"
