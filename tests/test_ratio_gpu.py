"""K6 parity on the GPU: ratio test + match compaction
against the numpy lines of the reference's example
(example/ex01_essential_estimation.py:102-105) applied to the oracle's output,
and against the C oracle's ratio mask.  Bit-exact: the pass mask, the match
count, the order and every (train idx, query row) pair."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NONE = np.uint64(0xFFFFFFFFFFFFFFFF)


@pytest.fixture(scope="module")
def feature():
    from spectavi_b200 import feature
    return feature


def _numpy_matches(idx, dist, min_ratio):
    # the reference's own lines, verbatim semantics
    with np.errstate(divide="ignore", invalid="ignore"):
        ratio = dist[:, 1] / dist[:, 0].astype("float64")
    pass_idx = ratio >= min_ratio
    idx0, _ = idx.T
    return idx0[pass_idx], np.arange(idx.shape[0], dtype=np.uint64)[pass_idx], pass_idx


@pytest.mark.parametrize("shape,alphabet", [
    ((3000, 2500, 128), 256), ((4096, 4096, 128), 256),
    ((700, 1025, 16), 4),        # tie stress: many d1 == d2 and d1 == 0 (inf / nan branches)
    ((1, 300, 16), 256),         # second neighbour missing: sentinel INT_MAX takes part in the division
    ((5, 1, 32), 256), ((300, 257, 144), 256),
])
@pytest.mark.parametrize("min_ratio", [1.75, 1.0, 1.2, 0.5])
def test_match_bruteforce_equals_numpy_ratio_test(po, feature, shape, alphabet, min_ratio):
    xr, yr, dim = shape
    rng = np.random.RandomState(xr + 3 * yr + alphabet)
    x = rng.randint(0, alphabet, (xr, dim)).astype("uint8")
    y = rng.randint(0, alphabet, (yr, dim)).astype("uint8")
    # plant near-duplicates so that some ratios are large / infinite
    nplant = min(yr, xr) // 3
    src = rng.randint(0, xr, nplant)
    y[:nplant] = x[src]
    y[:nplant // 2, 0] ^= 1
    oi, od = po.l1k2(x, y, 4)
    ex, ey, mask = _numpy_matches(oi.reshape(-1, 2), od.reshape(-1, 2), min_ratio)
    assert (mask == po.ratio_mask(od.reshape(-1, 2), min_ratio)).all()
    gx, gy = feature.match_bruteforcel1k2(x, y, min_ratio)
    assert gx.dtype == np.uint64 and gy.dtype == np.uint64
    assert gx.shape == ex.shape
    assert (gy == ey).all() and (gx == ex).all()


def test_match_equals_library_output_then_numpy(feature):
    rng = np.random.RandomState(5)
    x = rng.randint(0, 256, (20000, 128)).astype("uint8")
    y = rng.randint(0, 256, (30011, 128)).astype("uint8")
    y[::3] = np.clip(x[rng.randint(0, 20000, len(y[::3]))].astype(int) + rng.randint(-2, 3, (len(y[::3]), 128)), 0, 255)
    gi, gd = feature.nn_bruteforcel1k2(x, y)
    ex, ey, mask = _numpy_matches(gi, gd, 1.75)
    assert (feature.ratio_test(gd, 1.75) == mask).all()
    gx, gy = feature.match_bruteforcel1k2(x, y, 1.75)
    assert mask.sum() > 5000
    assert (gx == ex).all() and (gy == ey).all()


def test_match_no_rows(feature):
    x = np.zeros((0, 16), "uint8")
    y = np.zeros((7, 16), "uint8")
    gx, gy = feature.match_bruteforcel1k2(x, y, 1.75)      # INT_MAX / INT_MAX = 1 < 1.75
    assert gx.size == 0 and gy.size == 0
    gx, gy = feature.match_bruteforcel1k2(x, y, 1.0)       # ... but passes at 1.0, with the "none" index
    assert (gx == NONE).all() and (gy == np.arange(7)).all()
    gx, gy = feature.match_bruteforcel1k2(np.zeros((3, 16), "uint8"), np.zeros((0, 16), "uint8"))
    assert gx.size == 0 and gy.size == 0


@pytest.mark.parametrize("min_ratio", [1.75, 1.1])
def test_match_cascade_equals_numpy_ratio_test(feature, min_ratio):
    rng = np.random.RandomState(11)
    x = (rng.randint(0, 256, (6000, 128)) - 128).astype("float32")
    y = (rng.randint(0, 256, (4000, 128)) - 128).astype("float32")
    y[:2000] = np.clip(x[rng.randint(0, 6000, 2000)] + rng.randint(-2, 3, (2000, 128)), -128, 127)
    ci, cd = feature.nn_cascading_hash(x, y, k=2, m=8, n=4, g=2)
    assert cd.dtype == np.float32
    ex, ey, mask = _numpy_matches(ci, cd, min_ratio)         # float32 / float64, sentinel 2147483648.0f
    gx, gy = feature.match_cascading_hash(x, y, min_ratio, m=8, n=4, g=2)
    assert mask.sum() > 1000
    assert (gx == ex).all() and (gy == ey).all()


def test_device_ratio_mask_and_matches(po):
    import torch
    from spectavi_b200 import device as dev
    rng = np.random.RandomState(3)
    yr = 100_003
    d1 = rng.randint(0, 40, yr).astype(np.int64)
    d2 = d1 + rng.randint(0, 60, yr)
    idx = rng.randint(0, 2**31, (2, yr)).astype(np.int64)
    keys_np = np.stack([(d1 << 32) | idx[0], (d2 << 32) | idx[1]])
    keys_np[:, ::977] = -1                                    # "none" rows
    keys = torch.from_numpy(keys_np).cuda()
    for f32 in (False, True):
        for r in (1.75, 1.0, 3.0):
            fi, fd = dev.keys_finalize(keys, 2, torch.float32 if f32 else torch.int32)
            ex, ey, mask = _numpy_matches(fi.cpu().numpy().view(np.uint64), fd.cpu().numpy(), r)
            m, gm = dev.ratio_matches(keys, r, dist_f32=f32, want_mask=True)
            assert (gm.cpu().numpy().astype(bool) == mask).all()
            assert (gm.cpu().numpy() == po.ratio_mask(fd.cpu().numpy(), r)).all()
            m = m.cpu().numpy().view(np.uint64)
            assert m.shape[0] == mask.sum()
            assert (m[:, 0] == ex).all() and (m[:, 1] == ey).all()
