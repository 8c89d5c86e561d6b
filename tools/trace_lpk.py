"""Host-side timeline of nn_bruteforce (SPECTAVI_TRACE=1), 20 000 x 20 000 x 132 float32."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spectavi_b200 import feature
rng = np.random.RandomState(1)
x = rng.randn(20000, 132).astype("float32"); y = rng.randn(20000, 132).astype("float32")
for _ in range(3):
    feature.nn_bruteforce(x, y, p=1., mu=0., k=2)
ts = []
for _ in range(5):
    t0 = time.perf_counter(); feature.nn_bruteforce(x, y, p=1., mu=0., k=2); ts.append((time.perf_counter() - t0) * 1e3)
print("e2e ms:", " ".join("%.3f" % t for t in ts), flush=True)
os.environ["SPECTAVI_TRACE"] = "1"
for _ in range(2):
    t0 = time.perf_counter(); feature.nn_bruteforce(x, y, p=1., mu=0., k=2); print("traced call %.3f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
