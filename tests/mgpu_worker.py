"""Worker of tests/test_multigpu.py.  Two modes:
  torchrun ... mgpu_worker.py ranks     one process per GPU, NCCL all-gather + K5 merge
  SPECTAVI_GPUS=N python mgpu_worker.py abi   single process, N GPUs inside the C ABI
Both compare with the single-GPU result (bit for bit) and with the oracle on a sample."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def data(xrows, yrows, dim, alphabet):
    rng = np.random.RandomState(99)
    x = rng.randint(0, alphabet, (xrows, dim)).astype("uint8")
    y = rng.randint(0, alphabet, (yrows, dim)).astype("uint8")
    y[: yrows // 2] = x[rng.randint(0, xrows, yrows // 2)]
    return x, y


def run_ranks():
    import torch
    import torch.distributed as dist
    from spectavi_b200 import device as dev
    from spectavi_b200.sharding import shard_range
    from oracle import pyoracle as po
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    for (xr, yr, dim, alpha) in ((20001, 3000, 128, 256), (5003, 700, 144, 4), (1, 64, 16, 256)):
        x, y = data(xr, yr, dim, alpha)
        x0, xn = shard_range(xr, world, rank)
        dx = torch.from_numpy(np.ascontiguousarray(x[x0:x0 + xn])).cuda()
        dy = torch.from_numpy(y).cuda()
        idx, d = dev.nn_bruteforcel1k2_sharded(dx, dy, x0)
        # single-GPU result on this rank
        keys = dev.new_keys(yr, "cuda")
        dev.l1k2_fold(torch.from_numpy(x).cuda(), dy, keys, 0)
        sidx, sd = dev.keys_finalize(keys)
        assert torch.equal(idx, sidx) and torch.equal(d, sd), "sharded != single GPU (%d,%d,%d)" % (xr, yr, dim)
        oi, od = po.l1k2(x, y[:200], 4)
        assert (idx[:200].cpu().numpy().view(np.uint64) == oi).all() and (d[:200].cpu().numpy() == od).all()
        # cascade with the same dictionary on every rank
        m = 8 if xr > 1000 else 4
        dic = torch.from_numpy(dev.cascade_dictionary(dim, m, 2, seed=5))
        ci, cd = dev.nn_cascading_hash_sharded(dx, dy, dic, 2, x0)
        ck = dev.new_keys(yr, "cuda")
        ws = dev.cascade_fold(torch.from_numpy(x).cuda(), dy, ck, dic, 2, 0)
        si, sdist = dev.keys_finalize(ck, 2, torch.float32)
        torch.cuda.synchronize()
        del ws
        assert torch.equal(ci, si) and torch.equal(cd, sdist), "sharded cascade != single GPU"
    dist.barrier()
    if rank == 0:
        print("MGPU_RANKS_OK world=%d" % world)
    dist.destroy_process_group()


def run_abi():
    from spectavi_b200 import feature
    from spectavi_b200 import device as dev
    from oracle import pyoracle as po
    for (xr, yr, dim, alpha) in ((20001, 3000, 128, 256), (5003, 700, 144, 4), (3, 64, 16, 256)):
        x, y = data(xr, yr, dim, alpha)
        gi, gd = feature.nn_bruteforcel1k2(x, y)
        oi, od = po.l1k2(x, y, 4)
        assert (gi == oi).all() and (gd == od).all(), "SPECTAVI_GPUS l1k2 differs from the oracle"
        xf, yf = x.astype("float32") - 128, y.astype("float32") - 128
        m = 8 if xr > 1000 else 4
        ci, cd = feature.nn_cascading_hash(xf, yf, k=2, m=m, n=2, g=2)
        oi, od = po.cascade(xf, yf, dev.cascade_dictionary(dim, m, 2), 2, 4)
        assert (ci == oi).all() and (cd == od).all(), "SPECTAVI_GPUS cascade differs from the oracle"
    # nn_bruteforce shards the QUERY rows once there are 8192 per GPU (host_ctx.cu HostCall::lpk)
    rng = np.random.RandomState(5)
    xf = rng.randn(1500, 36).astype("float32")
    yf = rng.randn(8192 * 2 + 77, 36).astype("float32")
    for pw, k, mu in ((1., 2, 0.), (2., 3, 0.), (1., 2, .05)):
        gi, gd = feature.nn_bruteforce(xf, yf, p=pw, mu=mu, k=k)
        oi, od = po.bruteforce(xf, yf, k, pw, mu)
        assert (gi.reshape(oi.shape) == oi).all() and (gd.reshape(od.shape).view(np.uint32) == od.view(np.uint32)).all(), \
            "SPECTAVI_GPUS nn_bruteforce differs from the oracle"
    print("MGPU_ABI_OK gpus=%s" % os.environ.get("SPECTAVI_GPUS"))


if __name__ == "__main__":
    run_ranks() if sys.argv[1] == "ranks" else run_abi()
