import sys, os, numpy as np, torch
sys.path.insert(0, os.getcwd())
from spectavi_b200 import device as dev, feature
g = np.load("tests/golden/castle_c2.npz")
s1 = np.hstack([g["k1"], g["d1"].astype("float32")]); s2 = np.hstack([g["k2"], g["d2"].astype("float32")])
xu = (feature.normalize_to_ubyte_and_multiple_16_dim(s1) + 128).astype("uint8"); yu = (feature.normalize_to_ubyte_and_multiple_16_dim(s2) + 128).astype("uint8")
rng = np.random.RandomState(0)
for name, (x, y) in (("castle", (xu, yu)), ("uniform same shape", (rng.randint(0,256,xu.shape).astype("uint8"), rng.randint(0,256,yu.shape).astype("uint8")))):
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    keys = dev.new_keys(y.shape[0], "cuda")
    ms = []
    for _ in range(6):
        dev.keys_reset(keys); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); dev.l1k2_fold(dx, dy, keys, 0); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    ideal = x.shape[0] * y.shape[0] * (x.shape[1] / 4) / 18.2e12 * 1e3
    print(name, x.shape, y.shape, "K1 one launch ms", min(ms), "ideal at pipe peak", ideal)
    t = []
    import time
    for _ in range(6):
        t0 = time.perf_counter(); feature.nn_bruteforcel1k2(x, y); t.append((time.perf_counter() - t0) * 1e3)
    print("   e2e ms", min(t))
# where does a castle-sized launch lose time?  K1 alone against the train-row count (uniform bytes, 27 715 x 144 queries)
print("sweep: xrows, K1 ms, ideal ms, ms above ideal")
yq = torch.from_numpy(rng.randint(0, 256, yu.shape).astype("uint8")).cuda()
for xr in (1734, 3469, 6938, 13876, 27753, 55506, 111012):
    dx = torch.from_numpy(rng.randint(0, 256, (xr, xu.shape[1])).astype("uint8")).cuda()
    keys = dev.new_keys(yq.shape[0], "cuda")
    ms = []
    for _ in range(6):
        dev.keys_reset(keys); e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); dev.l1k2_fold(dx, yq, keys, 0); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    ideal = xr * yq.shape[0] * (xu.shape[1] / 4) / 18.2e12 * 1e3
    print("  %7d  %.4f  %.4f  %.4f" % (xr, min(ms), ideal, min(ms) - ideal))
# the same launch a second time WITHOUT resetting the keys: thresholds seeded, no candidate passes -> the cost of the slow path
for name, x in (("castle", xu), ("uniform", rng.randint(0, 256, xu.shape).astype("uint8"))):
    dx = torch.from_numpy(x).cuda(); dy = torch.from_numpy(yu if name == "castle" else yq.cpu().numpy()).cuda()
    keys = dev.new_keys(dy.shape[0], "cuda"); dev.l1k2_fold(dx, dy, keys, 0)
    ms = []
    for _ in range(6):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); dev.l1k2_fold(dx, dy, keys, 0); e1.record(); torch.cuda.synchronize(); ms.append(e0.elapsed_time(e1))
    print(name, "second pass over the same rows (nothing to insert) ms", min(ms))
