"""Developer timing loop for K1 (device-resident, CUDA events)."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from spectavi_b200 import device as dev

def run(xr, yr, dim, reps=3):
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randint(0, 256, (xr, dim), dtype=torch.uint8, device="cuda", generator=g)
    y = torch.randint(0, 256, (yr, dim), dtype=torch.uint8, device="cuda", generator=g)
    keys = dev.new_keys(yr, "cuda")
    dev.l1k2_fold(x, y, keys)
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        dev.keys_reset(keys)
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); dev.l1k2_fold(x, y, keys); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    gp = xr * yr / best / 1e6
    print("x %8d y %8d dim %4d  %9.3f ms  %8.1f Gpairs/s  %7.1f Gsad4/s (%.1f%% of 18530)" %
          (xr, yr, dim, best, gp, gp * dim / 4, gp * dim / 4 / 185.30), flush=True)

if __name__ == "__main__":
    for s in sys.argv[1:]:
        xr, yr, dim = map(int, s.split("x"))
        run(xr, yr, dim)
