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
        # the register-staged kernel (lp_topk.cu) on the same data, for the record
        os.environ["SPECTAVI_K1F"] = "v1"
        try:
            dev.nn_bruteforce(dx, dy, p=p, k=2)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(reps):
                dev.nn_bruteforce(dx, dy, p=p, k=2)
            e1.record()
            torch.cuda.synchronize()
            out["p=%g" % p]["device_ms_register_staged_kernel"] = e0.elapsed_time(e1) / reps
        finally:
            os.environ.pop("SPECTAVI_K1F", None)
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
    # the ceiling of each pair-column formulation, scalar and packed (f32x2), measured now; the roofline is the better one
    forms = {"p=1": ("f32 sub + |.|add (2 FADD)", "f32x2 sub + |.|add (2 FADD2 per 2)"),
             "p=2": ("f32 sub + mul + add (FADD FMUL FADD)", "f32x2 sub + mul + add (FADD2 2xFMUL FADD2 per 2)")}
    out["roofline"] = {"bound": "fp32 pipe", "unit": "T pair-columns/s", "kernel": "lp_topk_tma_kernel"}
    for q, (sc, pkd) in forms.items():
        ps, pp = pk[sc]["gops"] / 1e3, pk.get(pkd, {"gops": 0})["gops"] / 1e3
        out["roofline"][q] = {"achieved": out[q]["T_pair_columns_per_s"], "peak_scalar_form": ps, "peak_packed_form": pp,
                              "peak": max(ps, pp), "frac": out[q]["T_pair_columns_per_s"] / max(ps, pp),
                              "frac_of_scalar_form": out[q]["T_pair_columns_per_s"] / ps}
    return out


if __name__ == "__main__":
    print(json.dumps(run(*[int(v) for v in sys.argv[1:]])))
