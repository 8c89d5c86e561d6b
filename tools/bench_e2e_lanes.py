"""End-to-end time of the float entry points (pageable numpy in, numpy out) against SPECTAVI_UPLOAD_LANES.

    python tools/bench_e2e_lanes.py            # sweeps 1, 2, 4, 6, 8 lanes, one subprocess each (the library reads the
                                               # variable once), prints one JSON line per setting
"""
import json, os, subprocess, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def best_of(fn, reps):
    fn()
    best = 1e30
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); best = min(best, time.perf_counter() - t0)
    return best * 1e3


def one():
    from spectavi_b200 import feature
    rng = np.random.RandomState(3)
    out = {"lanes": os.environ.get("SPECTAVI_UPLOAD_LANES", "default"), "host_cores": os.cpu_count()}
    rows, dim = 1_000_000, 128
    x = rng.randint(0, 256, (rows, dim)).astype("float32") - 128
    y = rng.randint(0, 256, (rows, dim)).astype("float32") - 128
    t0 = time.perf_counter(); z = x.copy(); dt = time.perf_counter() - t0
    out["numpy_copy_512MB_GBps(1 thread, read + first-touch write)"] = x.nbytes / dt / 1e9
    del z
    out["cascade_1M_e2e_ms"] = best_of(lambda: feature.nn_cascading_hash(x, y, k=2, m=17, n=2, g=2), 3)
    out["cascade_h2d_GBps"] = (x.nbytes + y.nbytes) / out["cascade_1M_e2e_ms"] / 1e6
    del x, y
    a = rng.rand(20000, 132).astype("float32"); b = rng.rand(20000, 132).astype("float32")
    out["nn_bruteforce_20k_p1_e2e_ms"] = best_of(lambda: feature.nn_bruteforce(a, b, p=1., mu=0., k=2), 5)
    out["nn_bruteforce_20k_p2_e2e_ms"] = best_of(lambda: feature.nn_bruteforce(a, b, p=2., mu=0., k=2), 5)
    a = rng.rand(100000, 132).astype("float32"); b = rng.rand(100000, 132).astype("float32")
    out["nn_bruteforce_100k_p1_e2e_ms"] = best_of(lambda: feature.nn_bruteforce(a, b, p=1., mu=0., k=2), 2)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "one":
        one()
    else:
        for lanes in (sys.argv[1:] or ["1", "2", "4", "6", "8"]):
            env = dict(os.environ, SPECTAVI_UPLOAD_LANES=lanes)
            subprocess.run([sys.executable, os.path.abspath(__file__), "one"], env=env, check=False)
