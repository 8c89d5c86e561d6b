"""Host-side timeline of nn_cascading_hash (SPECTAVI_TRACE=1) at C5 and at the castle pair's size."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spectavi_b200 import feature
rng = np.random.RandomState(1)
for name, rows, dim, kw in (("c5", 1_000_000, 128, dict(m=17, n=2, g=2)), ("castle-size", 27753, 144, {})):
    x = rng.randint(0, 256, (rows, dim)).astype("float32") - 128
    y = rng.randint(0, 256, (rows, dim)).astype("float32") - 128
    os.environ.pop("SPECTAVI_TRACE", None)
    for _ in range(2):
        feature.nn_cascading_hash(x, y, k=2, **kw)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter(); feature.nn_cascading_hash(x, y, k=2, **kw); ts.append((time.perf_counter() - t0) * 1e3)
    print(name, "e2e ms:", " ".join("%.3f" % t for t in ts), flush=True)
    os.environ["SPECTAVI_TRACE"] = "1"
    t0 = time.perf_counter(); feature.nn_cascading_hash(x, y, k=2, **kw); print(name, "traced call %.3f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
