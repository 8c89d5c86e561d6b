#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_cascade_gpu.py tests/test_castle_c2.py -m gpu -q -x > $O/q_pytest.log 2>&1; echo "rc=$?" >> $O/q_pytest.log
tail -4 $O/q_pytest.log
timeout 600 python tools/bench_cascade.py 1000000 128 5 0 > $O/q_cascade.json 2> $O/q_cascade.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/q_cascade.json'))
print(d['cascade_ms']); print({k:v.get('ms') for k,v in d['stages'].items() if isinstance(v,dict)})
PY
timeout 300 python tools/bench_cascade.py castle > $O/q_castle.json 2>> $O/q_cascade.err; cut -c1-600 $O/q_castle.json
