"""SURVEY.md 8f rank 3, measured: KMedians::find_nearest_neighbours at the castle pair's
size (C2-like: ~10 k x 10 k x 132, c = 5, the wrapper's cluster-count rule) through the
host-buffer C ABI, with the reference's own find_nearest_neighbours (oracle/_ref, 1 thread:
the reference does not parallelise it) timed on the same clusters, and parity on every row.
    python tools/bench_kmedians.py [rows] [dim] [reps]
Prints one JSON line."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(rows=10000, dim=132, reps=10, c=5, k=2):
    from oracle import pyoracle as po
    from spectavi_b200 import feature
    rng = np.random.RandomState(5)
    x = rng.randint(0, 256, (rows, dim)).astype("float32")
    y = np.clip(x[rng.randint(0, rows, rows)] + rng.randn(rows, dim) * 9, 0, 255).astype("float32")
    nm = int(np.round(np.sqrt(rows / c) * c))
    out = {"workload": "KMedians::find_nearest_neighbours %d x %d x %d, %d clusters a side, c=%d k=%d" % (rows, rows, dim, nm, c, k)}
    if po.have_ref():
        t0 = time.perf_counter()
        ri, rd, xl, yl, xm, ym, sec = po.ref_kmedians_nn(x, y, nm, nm, c, k, seed=3, want_seconds=True)
        out["cpu_reference"] = {"kind": "reference", "cores": 1, "find_nearest_neighbours_ms": sec * 1e3,
                                "whole_nn_kmedians_ms": (time.perf_counter() - t0) * 1e3}
    else:   # no reference build on this box: any clustering is a valid input, parity then against the oracle
        xm, ym = x[rng.choice(rows, nm, replace=False)], y[rng.choice(rows, nm, replace=False)]
        xl = po.bruteforce(xm, x, 1, 1., 0.)[0][:, 0].astype("int32")
        yl = po.bruteforce(ym, y, 1, 1., 0.)[0][:, 0].astype("int32")
        ri, rd, _ = po.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, c, k)
    for _ in range(3):
        gi, gd = feature.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, c=c, k=k)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        gi, gd = feature.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, c=c, k=k)
        ts.append(time.perf_counter() - t0)
    out["gpu_e2e_ms"] = float(np.median(ts) * 1e3)
    out["parity_rows"] = int(rows)
    out["parity_bit_exact"] = bool((gi == ri).all() and (gd.view(np.uint32) == rd.view(np.uint32)).all())
    if "cpu_reference" in out:
        out["speedup_vs_reference"] = out["cpu_reference"]["find_nearest_neighbours_ms"] / out["gpu_e2e_ms"]
    return out


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    print(json.dumps(run(*a)))
