#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_f32_to_u8_host.py tests/test_cascade_gpu.py tests/test_castle_c2.py -m gpu -q -x > $O/t_pytest.log 2>&1; echo "rc=$?" >> $O/t_pytest.log
tail -4 $O/t_pytest.log
timeout 900 python tools/bench_e2e_cvt.py > $O/t_cvt.jsonl 2> $O/t_cvt.err; cat $O/t_cvt.jsonl; tail -3 $O/t_cvt.err
