#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_l1k2_gpu.py tests/test_castle_c2.py tests/test_ratio_gpu.py -m gpu -q > $O/pytest_l1.log 2>&1; echo "rc=$?" >> $O/pytest_l1.log
timeout 300 python tools/trace_e2e.py > $O/trace_e2e.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > $O/bench_n1_noextras.json 2> $O/bench_n1.err
tail -3 $O/pytest_l1.log; grep -v "piece queued" $O/trace_e2e.log; python -c "
import json; l=json.loads(open('$O/bench_n1_noextras.json').read().strip().split(chr(10))[-1]); print('value',l['value'],'ms',l['ms_per_step'],'e2e ms',l['e2e']['ms_per_step'],'ratio',l['e2e']['value']/l['value'], l['e2e']['equals_device_resident_result'], l['parity_bit_exact'])"
