#!/usr/bin/env python
"""C5 at N GPUs: the cascade-of-hashes matcher with the train set sharded over the ranks.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/bench_cascade_mgpu.py [rows=1000000] [dim=128] [reps=5]
Every rank holds all queries and 1/N of the train rows, hashes its shard with the SAME dictionary, probes with all
queries, and the per-rank top-2 keys are all-gathered over NCCL and merged (K5).  Timed with CUDA events per rank,
max over ranks; rank 0 also runs the whole train set alone and checks that the sharded result is bit-identical."""
import json, os, sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from spectavi_b200 import device as dev
from spectavi_b200.sharding import shard_range

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 128
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
device = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=device)
m = int(np.floor(np.log2(rows / 6.)))          # feature.py:364-367 on the WHOLE train set, as the reference would
n, g = 2, 2
gen = torch.Generator(device=device).manual_seed(3)     # the same bytes on every rank
x = torch.randint(0, 256, (rows, dim), dtype=torch.uint8, device=device, generator=gen)
y = torch.randint(0, 256, (rows, dim), dtype=torch.uint8, device=device, generator=gen)
half = rows // 2
src = torch.randint(0, rows, (half,), device=device, generator=gen)
noise = torch.randint(-3, 4, (half, dim), device=device, generator=gen)
y[:half] = (x[src].to(torch.int32) + noise).clamp(0, 255).to(torch.uint8)
dic = torch.from_numpy(dev.cascade_dictionary(dim, m, n)).to(device)
terms = dev.dict_terms_of(dic)
x0, xn = shard_range(rows, world, rank)
xs = x[x0:x0 + xn].contiguous()
keys = dev.new_keys(rows, device)
gathered = torch.empty((world, 2, rows), dtype=torch.int64, device=device)


def step():
    dev.keys_reset(keys)
    ws = dev.cascade_fold(xs, y, keys, dic, g, x0, dict_terms=terms)
    dist.all_gather_into_tensor(gathered, keys)
    dev.keys_merge(keys, gathered)
    return dev.keys_finalize(keys, 2, torch.float32), ws


for _ in range(3):
    step()
dist.barrier(); torch.cuda.synchronize()
ev = [(torch.cuda.Event(True), torch.cuda.Event(True)) for _ in range(reps)]
for a, b in ev:
    a.record(); (ci, cd), _ = step(); b.record()
dist.barrier(); torch.cuda.synchronize()
ms = torch.tensor([min(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=device)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    k1 = dev.new_keys(rows, device)
    ws = dev.cascade_fold(x, y, k1, dic, g, 0, dict_terms=terms)
    si, sd = dev.keys_finalize(k1, 2, torch.float32)
    torch.cuda.synchronize()
    print(json.dumps({"workload": "C5 cascade %d x %d x %d, m=%d n=%d g=%d, train sharded %d-way" % (rows, rows, dim, m, n, g, world),
                      "n_gpus": world, "cascade_ms(max over ranks, best of %d)" % reps: float(ms[0]),
                      "queries_per_s": rows / float(ms[0]) * 1e3, "effective_Gpairs_per_s": rows * rows / float(ms[0]) / 1e6,
                      "sharded_equals_single_gpu": bool((ci == si).all() and (cd == sd).all())}))
dist.barrier()
dist.destroy_process_group()
