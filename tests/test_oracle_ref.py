"""The C restatement against the reference's own headers compiled through the
Eigen shim (oracle/_ref).  Runs wherever oracle/_ref was built (the build
container; it also ships prebuilt to the GPU box).  CPU only."""
import numpy as np
import pytest

from oracle import pyoracle

pytestmark = pytest.mark.skipif(not pyoracle.have_ref(),
                                reason="oracle/_ref not built here")


@pytest.mark.parametrize("shape", [(64, 50, 16), (300, 200, 128),
                                   (513, 77, 144), (1000, 3, 256)])
@pytest.mark.parametrize("alphabet", [256, 3])
def test_l1k2_matches_reference(po, shape, alphabet):
    xr, yr, dim = shape
    rng = np.random.RandomState(xr + alphabet)
    x = rng.randint(0, alphabet, (xr, dim)).astype("uint8")
    y = rng.randint(0, alphabet, (yr, dim)).astype("uint8")
    ri, rd = po.ref_l1k2(x, y, 2)
    for prune in (True, False):
        oi, od = po.l1k2(x, y, 2, prune)
        assert (oi == ri).all() and (od == rd).all()


@pytest.mark.parametrize("p", [1., 2., .5, 3.])
@pytest.mark.parametrize("k", [1, 2, 4])
def test_bruteforce_f32_matches_reference(po, p, k):
    rng = np.random.RandomState(int(p * 10) + k)
    x = rng.randn(200, 132).astype("float32")
    y = rng.randn(100, 132).astype("float32")
    ri, rd = po.ref_bruteforce(x, y, k, p, 0.)
    oi, od = po.bruteforce(x, y, k, p, 0.)
    assert (oi == ri).all() and (od == rd).all()


def test_bruteforce_mu_matches_reference(po):
    rng = np.random.RandomState(11)
    x = rng.randn(300, 64).astype("float32")
    y = rng.randn(100, 64).astype("float32")
    for mu in (.01, .1):
        ri, rd = po.ref_bruteforce(x, y, 2, 2., mu)
        oi, od = po.bruteforce(x, y, 2, 2., mu)
        assert (oi == ri).all() and (od == rd).all()


def test_cascade_matches_reference(po):
    rng = np.random.RandomState(5)
    x = rng.randint(-128, 128, (500, 128)).astype("float32")
    y = rng.randint(-128, 128, (200, 128)).astype("float32")
    ri, rd, dic, codes, cnt = po.ref_cascade(x, y, 6, 3, 2, seed=99)
    assert (po.hashcodes(x, dic) == codes).all()
    oi, od, ocnt = po.cascade(x, y, dic, 2, want_counts=True)
    assert (ocnt == cnt).all() and (od == rd).all()
    # unordered_set scan order vs ascending idx: picks may differ only among
    # candidates at exactly the reported distance
    bx = (x + 128).astype("uint8").astype("int64")
    by = (y + 128).astype("uint8").astype("int64")
    for r, c in np.argwhere(oi != ri):
        assert np.abs(bx[int(ri[r, c])] - by[r]).sum() == int(od[r, c])
