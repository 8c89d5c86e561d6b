"""K1f parity on the GPU: nn_bruteforce / nn_bruteforcei through the reference's
ctypes call shape against (a) golden vectors produced by the reference's own
BruteForceNn.h, (b) the C oracle, (c) numpy -- indices exact, distances
BIT-EXACT for p in {1, 2, .5} (same operations in the same order) and within
4 float ulps for the general pow() branch."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def feature():
    from spectavi_b200 import feature
    return feature


def test_bruteforce_golden(feature):
    from conftest import golden
    g = golden("bruteforce_300x250x132.npz")
    x, y = g["x"], g["y"]
    for tag, p, k in (("p2", 2., 2), ("p1", 1., 2), ("p05", .5, 2), ("p2k5", 2., 5), ("p2mu", 2., 2)):
        # p2mu: mu = .05, the reference's approximate prune (BruteForceNn.h:80-86), emulated in its visiting order
        gi, gd = feature.nn_bruteforce(x, y, p=p, mu=.05 if tag == "p2mu" else 0., k=k)
        assert (gi == g["idx_" + tag]).all(), tag
        assert (gd.view(np.uint32) == g["dist_" + tag].view(np.uint32)).all(), tag
    gi, gd = feature.nn_bruteforce(x, y, p=1.5, mu=0., k=2)
    assert (gi == g["idx_p15"]).all()
    assert np.allclose(gd, g["dist_p15"], rtol=4 * 2.0**-23, atol=0)
    gi, gd = feature.nn_bruteforce(x, y, p=2., mu=0., k=2, use_int=True)
    assert gd.dtype == np.int32
    assert (gi == g["idx_int_p2"]).all() and (gd == g["dist_int_p2"]).all()


def test_bruteforce_reference_test_shape(po, feature):
    # test/test_feature.py:67-81: 1000 x 1000 x 132 randn, p=2, mu=0, k=2 -- indices equal numpy's
    np.random.seed(0xdeadbeef)
    x = np.random.randn(1000, 132).astype("float32")
    y = np.random.randn(1000, 132).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=2., mu=0., k=2)
    ni, _ = po.numpy_topk(x, y, 2, 2)
    assert (gi == ni).all()
    oi, od = po.bruteforce(x, y, 2, 2., 0.)
    assert (gi == oi).all() and (gd.view(np.uint32) == od.view(np.uint32)).all()


@pytest.mark.parametrize("shape", [(1000, 300, 128), (37, 19, 5), (129, 257, 17), (2500, 64, 33), (5, 7, 132)])
@pytest.mark.parametrize("p", [1., 2., .5])
@pytest.mark.parametrize("k", [1, 2, 7])
def test_bruteforce_bit_exact_vs_oracle(po, feature, shape, p, k):
    xr, yr, dim = shape
    rng = np.random.RandomState(xr + yr + dim + k)
    x = rng.randn(xr, dim).astype("float32")
    y = rng.randn(yr, dim).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=p, mu=0., k=k)
    oi, od = po.bruteforce(x, y, k, p, 0.)
    assert (gi.reshape(oi.shape) == oi).all()
    assert (gd.reshape(od.shape).view(np.uint32) == od.view(np.uint32)).all()


def test_bruteforce_staged_upload_path(po, feature):
    # more than 1 MB of input: the rows go up through the upload lanes (host_ctx.cu upload()), cut at 256 KB grains that
    # fall inside rows and across the x / y boundary
    rng = np.random.RandomState(77)
    x = rng.randn(3001, 132).astype("float32")
    y = rng.randn(2503, 132).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=1., mu=0., k=2)
    oi, od = po.bruteforce(x, y, 2, 1., 0.)
    assert (gi == oi).all() and (gd.view(np.uint32) == od.view(np.uint32)).all()


def test_bruteforce_fewer_rows_than_k(po, feature):
    # BruteForceNn.h:111-118 writes back to front: 3 neighbours of k=5 land in columns 2..4
    rng = np.random.RandomState(3)
    x = rng.randn(3, 20).astype("float32")
    y = rng.randn(6, 20).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=1., mu=0., k=5)
    oi, od = po.bruteforce(x, y, 5, 1., 0.)
    assert (gi == oi).all() and (gd == od).all()
    assert (gd[:, :2] == 0).all() and (gi[:, :2] == 0).all()


@pytest.mark.parametrize("p", [1., 2., .5, 1.5])
def test_bruteforcei_vs_oracle(po, feature, p):
    rng = np.random.RandomState(11)
    x = rng.randn(700, 40).astype("float32")
    y = rng.randn(90, 40).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=p, mu=0., k=3, use_int=True)
    xi, yi = np.round(100 * x).astype("int32"), np.round(100 * y).astype("int32")
    oi, od = po.bruteforce(xi, yi, 3, p, 0., use_int=True)
    if p == 1.5:   # pow(): a 1-ulp difference before truncation can move an integer by 1
        assert np.abs(gd.astype(np.int64) - od).max() <= 40
    else:
        assert (gi == oi).all() and (gd == od).all()


def test_bruteforce_ties_lexicographic(po, feature):
    # small-alphabet integers force equal distances: selection is lexicographic (dist, idx)
    rng = np.random.RandomState(5)
    x = rng.randint(0, 3, (400, 8)).astype("float32")
    y = rng.randint(0, 3, (50, 8)).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=1., mu=0., k=4)
    ni, nd = po.numpy_topk(x, y, 4, 1)
    assert (gi == ni).all() and (gd == nd.astype("float32")).all()


@pytest.mark.parametrize("k", [1, 2, 16, 17, 40])
@pytest.mark.parametrize("use_int", [False, True])
def test_bruteforce_tie_stress_many_tiles(po, feature, k, use_int):
    """Equal distances everywhere (3-symbol rows), train rows over many tiles and splits, k on both sides of the
    first-tile bound's limit (16): the result is the lexicographic (dist, idx) top-k whatever order candidates
    reach the row's list in."""
    rng = np.random.RandomState(k + 100 * use_int)
    x = rng.randint(0, 3, (3000, 12)).astype("float32")
    y = rng.randint(0, 3, (300, 12)).astype("float32")
    if use_int:
        x, y = x / 100., y / 100.          # the wrapper rounds 100 * v to int32 (feature.py:286-287)
    gi, gd = feature.nn_bruteforce(x, y, p=1., mu=0., k=k, use_int=use_int)
    gi, gd = gi.reshape(300, k), gd.reshape(300, k)
    xs, ys = (np.round(100 * x), np.round(100 * y)) if use_int else (x, y)
    ni, nd = po.numpy_topk(xs, ys, k, 1)
    assert (gi == ni).all() and (gd == nd.astype(gd.dtype)).all()


@pytest.mark.parametrize("p,k,use_int", [(1., 2, False), (2., 5, False), (.5, 1, False), (2., 3, True)])
def test_bruteforce_device_resident(po, p, k, use_int):
    """Part 2 of the header: the same kernel on device pointers (torch owns the memory and the stream)."""
    import torch
    from spectavi_b200 import device as dev
    rng = np.random.RandomState(int(p * 10) + k)
    x = rng.randn(1500, 36).astype("float32")
    y = rng.randn(700, 36).astype("float32")
    if use_int:
        x, y = np.round(100 * x).astype("int32"), np.round(100 * y).astype("int32")
    gi, gd = dev.nn_bruteforce(torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda(), p=p, k=k)
    oi, od = po.bruteforce(x, y, k, p, 0., use_int=use_int)
    assert (gi.cpu().numpy().view(np.uint64) == oi).all()
    assert (gd.cpu().numpy().view(np.uint32) == od.view(np.uint32)).all()
    # fewer train rows than k: right-aligned, zero-filled in front
    gi, gd = dev.nn_bruteforce(torch.from_numpy(x[:2]).cuda(), torch.from_numpy(y).cuda(), p=p, k=4)
    oi, od = po.bruteforce(x[:2], y, 4, p, 0., use_int=use_int)
    assert (gi.cpu().numpy().view(np.uint64) == oi).all() and (gd.cpu().numpy().view(np.uint32) == od.view(np.uint32)).all()


@pytest.mark.parametrize("xr,k,use_int", [(1000, 200, False), (300, 129, False), (150, 200, False), (700, 300, True)])
def test_bruteforce_k_above_128(po, feature, xr, k, use_int):
    """The reference takes any k (BruteForceNn.h:93-118); above 128 the search runs in passes of 128, each
    admitting only keys beyond the last one delivered.  Incl. fewer train rows than k (right-aligned, zero-filled)
    and tie-heavy integer rows."""
    rng = np.random.RandomState(k + xr)
    if use_int:
        x = rng.randint(0, 4, (xr, 12)).astype("float32") / 100.
        y = rng.randint(0, 4, (60, 12)).astype("float32") / 100.
    else:
        x = rng.randn(xr, 20).astype("float32")
        y = rng.randn(60, 20).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=1., mu=0., k=k, use_int=use_int)
    if use_int:
        ni, nd = po.numpy_topk(np.round(100 * x), np.round(100 * y), k, 1)
        assert (gi == ni).all() and (gd == nd.astype(gd.dtype)).all()
    else:
        oi, od = po.bruteforce(x, y, k, 1., 0.)
        assert (gi == oi).all() and (gd.view(np.uint32) == od.view(np.uint32)).all()


@pytest.mark.parametrize("shape,k", [((3000, 700, 132), 2), ((777, 129, 5), 7), ((300, 300, 33), 128), ((5000, 64, 16), 1)])
@pytest.mark.parametrize("p", [1., 2., .5, 1.5])
def test_bruteforce_tma_kernel_equals_register_staged_kernel(po, feature, shape, k, p):
    """float32 rows take the TMA-fed kernel (lp_topk_tma.cu: column-major copies, producer warp, packed FADD2);
    SPECTAVI_K1F=v1 forces the register-staged one.  Same operations in the same order: identical bits, ragged
    tiles / panels included (TMA zero fill adds +0.0)."""
    import os
    xr, yr, dim = shape
    rng = np.random.RandomState(xr + k)
    x = rng.randn(xr, dim).astype("float32")
    y = rng.randn(yr, dim).astype("float32")
    os.environ.pop("SPECTAVI_K1F", None)
    gi, gd = feature.nn_bruteforce(x, y, p=p, mu=0., k=k)
    os.environ["SPECTAVI_K1F"] = "v1"
    try:
        vi, vd = feature.nn_bruteforce(x, y, p=p, mu=0., k=k)
    finally:
        os.environ.pop("SPECTAVI_K1F", None)
    assert (gi == vi).all() and (gd.view(np.uint32) == vd.view(np.uint32)).all()
    if p != 1.5:
        oi, od = po.bruteforce(x, y, k, p, 0.)
        assert (gi.reshape(oi.shape) == oi).all() and (gd.reshape(od.shape).view(np.uint32) == od.view(np.uint32)).all()


@pytest.mark.parametrize("p,mu,k,use_int", [(2., .05, 2, False), (1., .3, 3, False), (.5, .02, 1, False), (2., 1.5, 2, True),
                                             (1., 2.0, 4, True), (2., -.1, 2, False)])
def test_bruteforce_mu_equals_reference_semantics(po, feature, p, mu, k, use_int):
    """mu != 0 makes the reference's search approximate: a train row is dropped as soon as
    distp + mu * (columns left) > worst, which depends on the visiting order.  The library walks the rows in the
    reference's order (one thread per query row), so idx and dist equal the oracle's restatement (itself checked
    against the reference build and its golden vector) bit for bit -- and differ from the exact top-k."""
    rng = np.random.RandomState(int(p * 10) + k)
    x = rng.randn(700, 36).astype("float32")
    y = rng.randn(90, 36).astype("float32")
    gi, gd = feature.nn_bruteforce(x, y, p=p, mu=mu, k=k, use_int=use_int)
    xs, ys = (np.round(100 * x).astype("int32"), np.round(100 * y).astype("int32")) if use_int else (x, y)
    oi, od = po.bruteforce(xs, ys, k, p, mu, use_int=use_int)
    assert (gi.reshape(oi.shape) == oi).all()
    assert (gd.reshape(od.shape).view(np.uint32) == od.view(np.uint32)).all()
