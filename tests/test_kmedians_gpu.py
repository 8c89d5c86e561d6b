"""SURVEY.md 8f rank 3 on the GPU: the filter overload of the generic matcher and
KMedians::find_nearest_neighbours through the C ABI, bit-exact against the golden
vectors of the reference's own headers and against the oracle."""
import numpy as np
import pytest

from conftest import golden
from test_kmedians_oracle import FILTER_CASES, filter_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def feature():
    from spectavi_b200 import feature
    return feature


def _eq(a, b):
    return bool((a.view(np.uint32) == b.view(np.uint32)).all())


@pytest.mark.parametrize("tag", [c[0] for c in FILTER_CASES])
def test_filter_overload_golden(feature, tag):
    g = golden("kmedians_nn.npz")
    x, y = g["km_x"].astype("float32"), g["km_y"]
    k, p, fx, fy = filter_case(g, tag)
    i0 = np.full((y.shape[0], k), 77, "uint64")
    d0 = np.full((y.shape[0], k), -3, "float32")
    gi, gd = feature.nn_bruteforce_filtered(x, y, fx, fy, p=p, mu=0., k=k, nn_idx=i0, nn_dist=d0)
    assert gi is i0 and gd is d0            # in-out, like the reference's Eigen::Ref outputs
    assert (gi == g["f_idx_" + tag]).all()
    assert _eq(gd, g["f_dist_" + tag])


def test_kmedians_nn_golden(feature):
    g = golden("kmedians_nn.npz")
    x, y = g["km_x"].astype("float32"), g["km_y"]
    gi, gd = feature.kmedians_find_nearest_neighbours(x, y, g["km_medians_x"], g["km_labels_x"], g["km_medians_y"],
                                                      g["km_labels_y"], c=int(g["km_c"]), k=int(g["km_k"]))
    assert gi.dtype == np.uint64 and gd.dtype == np.float32 and gi.shape == (900, 2)
    assert (gi == g["km_idx"]).all()
    assert _eq(gd, g["km_dist"])


@pytest.mark.parametrize("shape", [(3000, 2500, 132, 5, 2), (700, 900, 20, 3, 4), (5000, 300, 64, 2, 1), (260, 260, 7, 1, 3)])
def test_kmedians_nn_equals_oracle(po, feature, shape):
    xr, yr, dim, c, k = shape
    rng = np.random.RandomState(xr + dim)
    x = rng.rand(xr, dim).astype("float32")
    y = rng.rand(yr, dim).astype("float32")
    nmx = int(np.round(np.sqrt(xr / c) * c))      # feature.py:331-332
    nmy = int(np.round(np.sqrt(yr / c) * c))
    # any clustering is a valid input: nearest of nm sampled rows, deliberately unbalanced
    xm = x[rng.choice(xr, nmx, replace=False)]
    ym = y[rng.choice(yr, nmy, replace=False)]
    xl = po.bruteforce(xm, x, 1, 1., 0.)[0][:, 0].astype("int32")
    yl = po.bruteforce(ym, y, 1, 1., 0.)[0][:, 0].astype("int32")
    oi, od, _ = po.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, c, k)
    gi, gd = feature.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, c=c, k=k)
    gi, gd = gi.reshape(oi.shape), gd.reshape(od.shape)      # NdArray.asarray squeezes k = 1 like the reference's
    assert (gi == oi).all() and _eq(gd, od)


def test_kmedians_empty_clusters_follow_the_reference_quirk(po, feature):
    rng = np.random.RandomState(3)
    x = rng.rand(900, 16).astype("float32")
    y = rng.rand(600, 16).astype("float32")
    xm = rng.rand(6, 16).astype("float32")
    ym = rng.rand(5, 16).astype("float32")
    xl = rng.randint(0, 3, 900) * 2                 # train clusters 1, 3, 5 are empty
    yl = np.where(rng.rand(600) < .5, 0, 3)         # query clusters 1, 2, 4 are empty
    for c in (1, 2):
        oi, od, _ = po.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, c, 2)
        gi, gd = feature.kmedians_find_nearest_neighbours(x, y, xm, xl, ym, yl, c=c, k=2)
        assert (gi == oi).all() and _eq(gd, od)


@pytest.mark.parametrize("k", [1, 2, 7])
@pytest.mark.parametrize("p", [1., 2., .5])
def test_filter_overload_equals_oracle(po, feature, k, p):
    rng = np.random.RandomState(k + int(10 * p))
    x = rng.randn(2000, 40).astype("float32")
    y = rng.randn(1500, 40).astype("float32")
    fx = rng.choice(2000, 777, replace=False)
    fy = rng.choice(1500, 401, replace=False)
    for a, b in ((fx, fy), (np.r_[fx, fx[:50]], np.r_[fy, fy]), (fx[:k - 1], fy), ([], fy[:3]), (fx[:300], [])):
        i0 = rng.randint(0, 99, (1500, k)).astype("uint64")
        d0 = rng.rand(1500, k).astype("float32")
        ua, ub = np.unique(np.asarray(a, "uint64")), np.unique(np.asarray(b, "uint64"))   # set semantics
        if len(a) and not len(ua):
            continue
        oi, od = po.bruteforce_filtered(x, y, ua, ub, k, p, 0., i0.copy(), d0.copy())
        gi, gd = feature.nn_bruteforce_filtered(x, y, a, b, p=p, mu=0., k=k, nn_idx=i0.copy(), nn_dist=d0.copy())
        assert (gi == oi).all() and _eq(gd, od)


def test_filter_overload_rejects_bad_ids(feature):
    x = np.zeros((10, 4), "float32")
    with pytest.raises(ValueError):
        feature.nn_bruteforce_filtered(x, x, [10], [0])
    with pytest.raises(ValueError):
        feature.kmedians_find_nearest_neighbours(x, x, x[:2], np.full(10, 2), x[:2], np.zeros(10), c=1)


def test_filter_overload_tie_stress(po, feature):
    """Equal distances everywhere: with the filters in ascending order the oracle's visiting order is ascending row
    id, so its result is the lexicographic (dist, idx) top-k -- what the GPU returns for any list order."""
    rng = np.random.RandomState(8)
    x = rng.randint(0, 3, (2500, 10)).astype("float32")
    y = rng.randint(0, 3, (400, 10)).astype("float32")
    fx = np.sort(rng.choice(2500, 1900, replace=False))
    fy = np.sort(rng.choice(400, 333, replace=False))
    for k in (1, 3, 16, 20):
        oi, od = po.bruteforce_filtered(x, y, fx, fy, k, 1., 0.)
        gi, gd = feature.nn_bruteforce_filtered(x, y, fx[::-1].copy(), fy[::-1].copy(), p=1., mu=0., k=k)
        assert (gi == oi).all() and _eq(gd, od)
