"""SURVEY.md 8f rank 3 on the CPU: the oracle's restatement of the filter overload
(BruteForceNn.h:37-120) and of KMedians::find_nearest_neighbours (KMedians.h:260-295)
against golden vectors made by the reference's own headers (tests/golden/make_kmedians.py)
and, where oracle/_ref is built, against the reference live."""
import numpy as np
import pytest

from conftest import golden
from oracle import pyoracle

FILTER_CASES = (("k3p1", 3, 1., "fx", "fy"), ("k1p2", 1, 2., "fx", "fy"), ("k3p05", 3, .5, "fx2", "fy"),
                ("allx", 2, 1., "none", "fy"), ("ally", 2, 1., "fx", "none"))


def filter_case(g, tag):
    fx, fy = g["f_fx"], g["f_fy"]
    sel = {"fx": fx, "fy": fy, "fx2": fx[:2], "none": fx[:0]}
    for t, k, p, a, b in FILTER_CASES:
        if t == tag:
            return k, p, sel[a], sel[b]
    raise KeyError(tag)


@pytest.mark.parametrize("tag", [c[0] for c in FILTER_CASES])
def test_oracle_filter_overload_golden(po, tag):
    g = golden("kmedians_nn.npz")
    x, y = g["km_x"].astype("float32"), g["km_y"]
    k, p, fx, fy = filter_case(g, tag)
    i0 = np.full((y.shape[0], k), 77, "uint64")
    d0 = np.full((y.shape[0], k), -3, "float32")
    oi, od = po.bruteforce_filtered(x, y, fx, fy, k, p, 0., i0, d0)
    assert (oi == g["f_idx_" + tag]).all()
    assert (od.view(np.uint32) == g["f_dist_" + tag].view(np.uint32)).all()
    if tag == "k3p05":      # two candidates, k = 3: column 0 keeps the caller's value
        rows = fy.astype(int)
        assert (oi[rows, 0] == 77).all() and (od[rows, 0] == -3).all()
    untouched = np.setdiff1d(np.arange(y.shape[0]), fy.astype(int)) if len(fy) else []
    assert (oi[untouched] == 77).all()


def test_oracle_kmedians_nn_golden(po):
    g = golden("kmedians_nn.npz")
    x, y = g["km_x"].astype("float32"), g["km_y"]
    oi, od, near = po.kmedians_find_nearest_neighbours(x, y, g["km_medians_x"], g["km_labels_x"], g["km_medians_y"],
                                                       g["km_labels_y"], int(g["km_c"]), int(g["km_k"]))
    assert (oi == g["km_idx"]).all()
    assert (od.view(np.uint32) == g["km_dist"].view(np.uint32)).all()
    # the approximation itself is the reference's (test/test_feature.py:84-87 calls it "very disappointing"):
    # on these rows about a third of the true nearest neighbours survive the cluster filter
    bi, _ = po.bruteforce(x, y, 2, 1., 0.)
    assert 0.2 < (bi[:, 0] == oi[:, 0]).mean() < 0.5


def test_oracle_kmedians_empty_cluster_means_all_rows(po):
    """BruteForceNn.h:45-46: an empty filter is `no filter`.  A query cluster without members makes the
    reference search ALL query rows against that cluster's candidates; later clusters overwrite their rows."""
    rng = np.random.RandomState(3)
    x = rng.rand(90, 8).astype("float32")
    y = rng.rand(60, 8).astype("float32")
    xm = rng.rand(4, 8).astype("float32")
    ym = rng.rand(3, 8).astype("float32")
    xl = rng.randint(0, 4, 90)
    yl = rng.randint(0, 2, 60) * 2          # cluster 1 of y is empty
    oi, od, near = po.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, 2, 2)
    cand1 = np.flatnonzero(np.isin(xl, near[1].astype(int)))
    rows0 = np.flatnonzero(yl == 0)         # written by group 0, then overwritten by the all-rows group 1
    wi, wd = po.bruteforce_filtered(x, y, cand1, rows0, 2, 1., 0.)
    assert (oi[rows0] == wi[rows0]).all()


needs_ref = pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref not built here")


@needs_ref
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_oracle_kmedians_nn_matches_reference_live(po, seed):
    rng = np.random.RandomState(seed)
    x = rng.rand(400 + 37 * seed, 20).astype("float32")
    y = rng.rand(300, 20).astype("float32")
    ri, rd, xl, yl, xm, ym = po.ref_kmedians_nn(x, y, 9 + seed, 7, 3, 2, seed=seed)
    assert xl.min() >= 0 and yl.min() >= 0
    oi, od, _ = po.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, 3, 2)
    assert (oi == ri).all() and (od == rd).all()


@needs_ref
@pytest.mark.parametrize("k", [1, 2, 5])
def test_oracle_filter_overload_matches_reference_live(po, k):
    rng = np.random.RandomState(k)
    x = rng.rand(700, 24).astype("float32")
    y = rng.rand(500, 24).astype("float32")
    fx = rng.choice(700, 200, replace=False)
    fy = rng.choice(500, 120, replace=False)
    for a, b in ((fx, fy), ([], fy), (fx, []), (fx[:1], fy)):
        i0 = rng.randint(0, 99, (500, k)).astype("uint64")
        d0 = rng.rand(500, k).astype("float32")
        oi, od = po.bruteforce_filtered(x, y, a, b, k, 1., 0., i0.copy(), d0.copy())
        ri, rd = po.ref_bruteforce_filtered(x, y, a, b, k, 1., 0., i0.copy(), d0.copy())
        assert (oi == ri).all() and (od == rd).all()
