#!/bin/bash
mkdir -p gpurun_out; O=gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q -k "single_process_c_abi" > $O/pytest_mgpu_abi_n8.log 2>&1; echo "rc=$?" >> $O/pytest_mgpu_abi_n8.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29658 bench.py --gpus 8 --steps 5 --warmup 3 --cpu-seconds 4 > $O/bench_n8.json 2> $O/bench_n8.err; echo "rc=$?" >> $O/bench_n8.err
tail -3 $O/pytest_mgpu_abi_n8.log; python -c "
import json; l=json.loads(open('$O/bench_n8.json').read().strip().split(chr(10))[-1]); print(l['value'], l['e2e']['value'], l['e2e']['value']/l['value'], l['e2e']['equals_device_resident_result'], l['parity_bit_exact'], l['result_hash'], l['extras']['torchrun_layout_e2e']['value'], l['extras']['c5_cascade_sharded'])"; tail -2 $O/bench_n8.err
