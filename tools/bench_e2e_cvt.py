"""End-to-end time of nn_cascading_hash (pageable float32 numpy in) against where the float32 -> uint8 conversion
happens (SPECTAVI_CVT=device: K0 after the upload; default: in the staging threads) and SPECTAVI_CVT_LANES.
    python tools/bench_e2e_cvt.py        # one subprocess per setting (the library reads the variables once)"""
import hashlib, json, os, subprocess, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one():
    from spectavi_b200 import feature
    rng = np.random.RandomState(3)
    out = {"cvt": os.environ.get("SPECTAVI_CVT", "host"), "lanes": os.environ.get("SPECTAVI_CVT_LANES", "default"),
           "host_cores": os.cpu_count()}
    for name, rows, dim, kw in (("c5_1M", 1_000_000, 128, dict(m=17, n=2, g=2)), ("castle_size", 27753, 144, {})):
        x = rng.randint(0, 256, (rows, dim)).astype("float32") - 128
        y = rng.randint(0, 256, (rows, dim)).astype("float32") - 128
        feature.nn_cascading_hash(x, y, k=2, **kw)
        best = 1e30
        for _ in range(4):
            t0 = time.perf_counter(); i, d = feature.nn_cascading_hash(x, y, k=2, **kw); best = min(best, time.perf_counter() - t0)
        out[name + "_e2e_ms"] = round(best * 1e3, 3)
        out[name + "_hash"] = hashlib.blake2b(i.tobytes() + d.tobytes(), digest_size=8).hexdigest()
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "one":
        one()
    else:
        for cvt, lanes in (("device", None), ("host", "4"), ("host", "8"), ("host", "12"), ("host", "16")):
            env = dict(os.environ)
            if cvt == "device": env["SPECTAVI_CVT"] = "device"
            if lanes: env["SPECTAVI_CVT_LANES"] = lanes
            subprocess.run([sys.executable, os.path.abspath(__file__), "one"], env=env, check=False)
