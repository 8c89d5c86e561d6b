"""K1f (nn_bruteforce, float32) timing through the host-buffer C ABI: rows x rows x 132, p in {1, 2, .5}.
    python tools/bench_lp.py [rows] [reps]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(rows=20000, reps=5):
    from spectavi_b200 import feature
    rng = np.random.RandomState(0)
    x = rng.randn(rows, 132).astype("float32")
    y = rng.randn(rows, 132).astype("float32")
    out = {"workload": "nn_bruteforce %d x %d x 132 float32, k=2" % (rows, rows)}
    for p in (1., 2., .5):
        for _ in range(2):
            feature.nn_bruteforce(x, y, p=p, mu=0., k=2)
        ts = []
        for _ in range(reps):
            t0 = time.perf_counter()
            feature.nn_bruteforce(x, y, p=p, mu=0., k=2)
            ts.append(time.perf_counter() - t0)
        ms = float(np.median(ts) * 1e3)
        out["p=%g" % p] = {"e2e_ms": ms, "Gpairs_per_s": rows * rows / ms / 1e6}
    return out


if __name__ == "__main__":
    print(json.dumps(run(*[int(v) for v in sys.argv[1:]])))
