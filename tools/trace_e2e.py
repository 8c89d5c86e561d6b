"""Host-side timeline of the drop-in call (SPECTAVI_TRACE=1): C3 and the castle pair, third call of each."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spectavi_b200 import feature
rng = np.random.RandomState(1)
for name, (xr, yr, dim) in (("c3", (200000, 200000, 128)), ("castle-size", (27753, 27715, 144))):
    x = rng.randint(0, 256, (xr, dim)).astype("uint8"); y = rng.randint(0, 256, (yr, dim)).astype("uint8")
    os.environ.pop("SPECTAVI_TRACE", None)
    for _ in range(2):
        feature.nn_bruteforcel1k2(x, y)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter(); feature.nn_bruteforcel1k2(x, y); ts.append((time.perf_counter() - t0) * 1e3)
    print(name, "e2e ms:", " ".join("%.3f" % t for t in ts), flush=True)
    os.environ["SPECTAVI_TRACE"] = "1"
    t0 = time.perf_counter(); feature.nn_bruteforcel1k2(x, y); print(name, "traced call %.3f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
