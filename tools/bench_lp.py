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
    # device-resident: K1f alone, CUDA events on the launching stream, inputs already in HBM
    import torch
    from spectavi_b200 import device as dev
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    for p in (1., 2.):
        for _ in range(3):
            dev.nn_bruteforce(dx, dy, p=p, k=2)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            dev.nn_bruteforce(dx, dy, p=p, k=2)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        out["p=%g" % p]["device_ms"] = ms
        out["p=%g" % p]["T_pair_columns_per_s"] = rows * rows * 132 / ms / 1e9
    # the FP32 issue ceilings K1f's inner loop is built from, measured now (pipe_peak.cu)
    import ctypes as ct
    from spectavi_b200._libspectavi import clib
    o = (ct.c_double * 64)(); names = (ct.c_char_p * 32)(); lanes = (ct.c_double * 32)()
    clib.spv_pipe_peak.restype = ct.c_int
    clib.spv_pipe_peak.argtypes = [ct.c_int, ct.c_int, ct.c_void_p, ct.c_void_p, ct.c_void_p]
    n = clib.spv_pipe_peak(0, 4000, o, names, lanes)
    out["pipe_peak"] = {names[v].decode(): {"ops_per_clk_sm": o[2 * v], "gops": o[2 * v + 1]} for v in range(n)
                        if names[v].decode().startswith(("FADD", "f32", "FFMA"))}
    pk = out["pipe_peak"]
    out["roofline"] = {"bound": "fp32 pipe", "unit": "T pair-columns/s",
                       "p=1": {"achieved": out["p=1"]["T_pair_columns_per_s"], "peak": pk["f32 sub + |.|add (2 FADD)"]["gops"] / 1e3},
                       "p=2": {"achieved": out["p=2"]["T_pair_columns_per_s"], "peak": pk["f32 sub + mul + add (FADD FMUL FADD)"]["gops"] / 1e3}}
    for q in ("p=1", "p=2"):
        out["roofline"][q]["frac"] = out["roofline"][q]["achieved"] / out["roofline"][q]["peak"]
    return out


if __name__ == "__main__":
    print(json.dumps(run(*[int(v) for v in sys.argv[1:]])))
