#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 900 python -m pytest tests/test_cascade_gpu.py tests/test_castle_c2.py -m gpu -q > $O/pytest_c.log 2>&1; echo "rc=$?" >> $O/pytest_c.log
timeout 600 python tools/bench_cascade.py 1000000 128 5 1000 > $O/cascade_c5.json 2> $O/cascade.err
SPECTAVI_K2=legacy timeout 600 python tools/bench_cascade.py 1000000 128 5 0 > $O/cascade_c5_legacy.json 2>> $O/cascade.err
tail -4 $O/pytest_c.log; python -c "
import json
for f in ('cascade_c5','cascade_c5_legacy'):
    c=json.load(open('$O/%s.json'%f)); print(f, c['cascade_ms'], {k:(round(v['ms']*1e3,1), round(v.get('frac_of_hbm_peak',0),3)) for k,v in c['stages'].items() if isinstance(v,dict)}, c.get('parity_bit_exact'))"
tail -2 $O/cascade.err
