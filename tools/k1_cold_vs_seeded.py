"""Two K1 launches on a castle-sized uniform problem for ncu: #1 cold (keys reset), #2 over the seeded keys."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from spectavi_b200 import device as dev
rng = np.random.RandomState(0)
x = torch.from_numpy(rng.randint(0, 256, (27753, 144)).astype("uint8")).cuda()
y = torch.from_numpy(rng.randint(0, 256, (27715, 144)).astype("uint8")).cuda()
keys = dev.new_keys(y.shape[0], "cuda")
dev.keys_reset(keys)
dev.l1k2_fold(x, y, keys, 0)
dev.l1k2_fold(x, y, keys, 0)
torch.cuda.synchronize()
print("done")
