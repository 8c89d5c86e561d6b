"""e2e of nn_bruteforcel1k2 at the castle pair's size, best and median of 30 calls (run once per library variant)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spectavi_b200 import feature
rng = np.random.RandomState(1)
for (xr, yr, dim) in ((27753, 27715, 144), (60000, 60000, 128)):
    x = rng.randint(0, 256, (xr, dim)).astype("uint8"); y = rng.randint(0, 256, (yr, dim)).astype("uint8")
    for _ in range(3):
        feature.nn_bruteforcel1k2(x, y)
    ts = []
    for _ in range(30):
        t0 = time.perf_counter(); feature.nn_bruteforcel1k2(x, y); ts.append((time.perf_counter() - t0) * 1e3)
    print(os.path.basename(os.environ.get("SPECTAVI_B200_LIB", "default")), (xr, yr, dim), "best %.3f median %.3f ms" % (min(ts), float(np.median(ts))), flush=True)
