"""K1 parity on the GPU: libspectavi.so's nn_bruteforcel1k2 (through the
reference's ctypes call shape) against the CPU oracle, bit-exact on idx AND
dist, including the lowest-index tie-break and the sentinels."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def feature():
    from spectavi_b200 import feature
    return feature


def _rand(rng, rows, dim, alphabet):
    return rng.randint(0, alphabet, (rows, dim)).astype("uint8")


@pytest.mark.parametrize("shape", [
    (200, 200, 144),      # reference test/test_feature.py:102-121
    (4096, 4096, 128),    # BASELINE.json configs[0]
    (1, 5, 16), (2, 3, 16), (3, 1, 32),     # sentinels / yrows==1 squeeze
    (63, 255, 128), (65, 257, 128), (1000, 777, 128),   # ragged tiles
    (5000, 300, 256), (700, 500, 384), (300, 200, 512),  # wider rows
    (300, 200, 1024), (1500, 130, 528), (70, 65, 4096),  # dim > 512: the wide-row kernel (reference: any dim % 16 == 0)
])
@pytest.mark.parametrize("alphabet", [256, 4])
def test_l1k2_bit_exact(po, feature, shape, alphabet):
    xr, yr, dim = shape
    rng = np.random.RandomState(xr * 7 + yr + alphabet)
    x, y = _rand(rng, xr, dim, alphabet), _rand(rng, yr, dim, alphabet)
    gi, gd = feature.nn_bruteforcel1k2(x, y)
    oi, od = po.l1k2(x, y, 4)
    oi, od = oi.squeeze(), od.squeeze()
    assert gi.dtype == np.uint64 and gd.dtype == np.int32
    assert gi.shape == oi.shape
    assert (gd == od).all()
    assert (gi == oi).all()


def test_l1k2_reference_test_distribution(po, feature):
    # test/test_feature.py:112-120: uniform [0,256) u8, 200 x 200 x 144, distances
    # equal numpy's all-pairs L1 top-2
    np.random.seed(0xdeadbeef)
    x = np.random.uniform(0, 256, (200, 144)).astype("uint8")
    y = np.random.uniform(0, 256, (200, 144)).astype("uint8")
    gi, gd = feature.nn_bruteforcel1k2(x, y)
    ni, nd = po.numpy_l1k2(x, y)
    assert (gd == nd).all() and (gi == ni).all()


def test_l1k2_empty_train(feature):
    x = np.zeros((0, 16), "uint8")
    y = np.zeros((4, 16), "uint8")
    gi, gd = feature.nn_bruteforcel1k2(x, y)
    assert (gi == np.uint64(0xFFFFFFFFFFFFFFFF)).all() and (gd == 2**31 - 1).all()


def test_l1k2_golden(feature):
    """The committed fixtures were produced by the reference's own headers
    (tests/golden/make_golden.py -> oracle/_ref)."""
    from conftest import golden
    for name in ("l1k2_uniform_200x200x144.npz", "l1k2_ties_257x129x128.npz", "l1k2_ties_257x129x16.npz"):
        g = golden(name)
        gi, gd = feature.nn_bruteforcel1k2(g["x"], g["y"])
        assert (gi == g["idx"]).all() and (gd == g["dist"]).all(), name
    g = golden("l1k2_sentinels.npz")
    for xr in (0, 1, 2):
        gi, gd = feature.nn_bruteforcel1k2(np.ascontiguousarray(g["x"][:xr]), g["y"])
        assert (gi == g["idx%d" % xr]).all() and (gd == g["dist%d" % xr]).all(), xr
    g = golden("l1k2_sift_surogre.npz")
    s = int(g["split"])
    gi, gd = feature.nn_bruteforcel1k2(np.ascontiguousarray(g["desc"][:s]), np.ascontiguousarray(g["desc"][s:]))
    assert (gi == g["idx"]).all() and (gd == g["dist"]).all()


def test_l1k2_large_sampled(po, feature):
    # 200k x 20k: every query row checked against the oracle on a sample
    rng = np.random.RandomState(5)
    x = _rand(rng, 200_000, 128, 256)
    y = _rand(rng, 20_000, 128, 256)
    gi, gd = feature.nn_bruteforcel1k2(x, y)
    sel = rng.choice(20_000, 256, replace=False)
    oi, od = po.l1k2(x, np.ascontiguousarray(y[sel]), 8)
    assert (gi[sel] == oi).all() and (gd[sel] == od).all()


@pytest.mark.parametrize("shape,alphabet", [((30_000, 20_000, 32), 256), ((9_000, 41_000, 16), 4), ((70_000, 8_193, 64), 256)])
def test_l1k2_streamed_upload_full_check(po, feature, shape, alphabet):
    """Shapes past the first piece of the host path's streamed upload (8 192 rows a side, host_ctx.cu): several
    query pieces against the first train piece, then further train pieces folded into the same keys with seeded
    thresholds.  Every row checked, incl. a tie-heavy alphabet (seeded keys may carry larger indices than later
    candidates)."""
    xr, yr, dim = shape
    rng = np.random.RandomState(xr + yr)
    x, y = _rand(rng, xr, dim, alphabet), _rand(rng, yr, dim, alphabet)
    gi, gd = feature.nn_bruteforcel1k2(x, y)
    oi, od = po.l1k2(x, y, 8)
    assert (gd == od).all() and (gi == oi).all()


def test_l1k2_fold_in_chunks_any_order_and_twice(po):
    """Device API: folding the train set chunk by chunk in DESCENDING index order (every stored key has a larger
    index than the next candidates), and folding every chunk twice, gives the one-launch result: no duplicate
    of the best key in rank 1 (fold_key skips a key that is already there)."""
    import torch
    from spectavi_b200 import device as dev
    rng = np.random.RandomState(31)
    for alphabet, dim in ((256, 128), (3, 16)):
        x, y = _rand(rng, 5000, dim, alphabet), _rand(rng, 700, dim, alphabet)
        dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
        keys = dev.new_keys(700, dx.device)
        bounds = [0, 64, 1000, 1001, 3000, 5000]
        for lo, hi in reversed(list(zip(bounds[:-1], bounds[1:]))):
            for _ in range(2):
                dev.l1k2_fold(dx[lo:hi].contiguous(), dy, keys, lo)
        gi, gd = dev.keys_finalize(keys)
        oi, od = po.l1k2(x, y, 4)
        assert (gi.cpu().numpy().view(np.uint64) == oi).all() and (gd.cpu().numpy() == od).all()


def test_l1k2_concurrent_callers(po, feature):
    """The reference's entry points are re-entrant (no globals); here every concurrent caller works on its own
    context (streams + buffers) taken from a pool (SURVEY 8b 'Threading').  Four threads, different inputs, each result exact."""
    import threading
    rng = np.random.RandomState(77)
    jobs = [(_rand(rng, 2000 + 300 * t, 128, 256), _rand(rng, 700 + 100 * t, 128, 256)) for t in range(4)]
    want = [po.l1k2(x, y, 2) for x, y in jobs]
    got = [None] * 4

    def work(t):
        for _ in range(5):
            got[t] = feature.nn_bruteforcel1k2(*jobs[t])

    th = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    for t in range(4):
        assert (got[t][0] == want[t][0]).all() and (got[t][1] == want[t][1]).all()
