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
    # the FP32 issue ceilings K1f's inner loop is built from, measured now (pipe_peak.cu)
    import ctypes as ct
    from spectavi_b200._libspectavi import clib
    o = (ct.c_double * 64)(); names = (ct.c_char_p * 32)(); lanes = (ct.c_double * 32)()
    clib.spv_pipe_peak.restype = ct.c_int
    clib.spv_pipe_peak.argtypes = [ct.c_int, ct.c_int, ct.c_void_p, ct.c_void_p, ct.c_void_p]
    n = clib.spv_pipe_peak(0, 4000, o, names, lanes)
    out["pipe_peak"] = {names[v].decode(): {"ops_per_clk_sm": o[2 * v], "gops": o[2 * v + 1]} for v in range(n)
                        if names[v].decode().startswith(("FADD", "f32", "FFMA"))}
    return out


if __name__ == "__main__":
    print(json.dumps(run(*[int(v) for v in sys.argv[1:]])))
