"""nn_bruteforce through the C ABI with the query rows sharded over SPECTAVI_GPUS (set by the caller of this script).
    SPECTAVI_GPUS=2 python tools/bench_lp_mgpu.py [rows=100000]"""
import hashlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spectavi_b200 import feature
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
rng = np.random.RandomState(0)
x = rng.randn(rows, 132).astype("float32"); y = rng.randn(rows, 132).astype("float32")
out = {"workload": "nn_bruteforce %d x %d x 132 float32, k=2, pageable numpy through the C ABI" % (rows, rows),
       "SPECTAVI_GPUS": os.environ.get("SPECTAVI_GPUS", "1")}
for p in (1., 2.):
    feature.nn_bruteforce(x, y, p=p, mu=0., k=2)
    best = 1e30
    for _ in range(3):
        t0 = time.perf_counter(); i, d = feature.nn_bruteforce(x, y, p=p, mu=0., k=2); best = min(best, time.perf_counter() - t0)
    out["p=%g" % p] = {"e2e_ms": best * 1e3, "Gpairs_per_s": rows * rows / best / 1e9,
                       "result_hash": hashlib.blake2b(i.tobytes() + d.tobytes(), digest_size=8).hexdigest()}
print(json.dumps(out))
