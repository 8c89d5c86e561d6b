"""K7 parity on the GPU: the descriptor quantiser against
the reference's numpy function (spectavi/feature.py:384-407, mirrored verbatim
in spectavi_b200.feature.normalize_to_ubyte_and_multiple_16_dim).  Bit-exact:
every output element compared with ==, NaN positions included."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def feature():
    from spectavi_b200 import feature
    return feature


def _same(a, b):
    return a.shape == b.shape and bool(((a == b) | (np.isnan(a) & np.isnan(b))).all())


@pytest.mark.parametrize("rows,dim", [(1000, 132), (300, 128), (200, 144), (10007, 132), (1, 20), (2, 16), (70001, 40)])
@pytest.mark.parametrize("kind", ["randn", "sift", "uniform"])
def test_normalize_equals_numpy(feature, rows, dim, kind):
    rng = np.random.RandomState(rows + dim)
    if kind == "randn":            # test/test_feature.py:138-141
        x = rng.randn(rows, dim).astype("float32")
    elif kind == "sift":           # integer-valued descriptors + float keypoint columns (sur-ogre.sift rows)
        x = rng.randint(0, 256, (rows, dim)).astype("float32")
        x[:, :4] = (rng.rand(rows, 4) * 1000).astype("float32")
    else:
        x = (rng.rand(rows, dim) * 255).astype("float32")
    with np.errstate(divide="ignore", invalid="ignore"):
        want = feature.normalize_to_ubyte_and_multiple_16_dim(x)
    got = feature.normalize_to_ubyte_and_multiple_16_dim_gpu(x)
    assert got.dtype == np.float32 and want.dtype == np.float32
    assert _same(got, want)
    assert got.shape[1] % 16 == 0


def test_normalize_constant_column_is_nan_like_numpy(feature):
    rng = np.random.RandomState(1)
    x = rng.randn(50, 20).astype("float32")
    x[:, 3] = 7.0
    with np.errstate(divide="ignore", invalid="ignore"):
        want = feature.normalize_to_ubyte_and_multiple_16_dim(x)
    got = feature.normalize_to_ubyte_and_multiple_16_dim_gpu(x)
    assert np.isnan(want[:, 3]).all() and _same(got, want)


def test_normalize_real_sift_rows(feature):
    z = np.load(os.path.join(GOLD, "l1k2_sift_surogre.npz"))
    x = z[[k for k in z.files if z[k].ndim == 2 and z[k].shape[1] >= 128][0]].astype("float32")
    want = feature.normalize_to_ubyte_and_multiple_16_dim(x)
    got = feature.normalize_to_ubyte_and_multiple_16_dim_gpu(x)
    assert _same(got, want)


def test_device_quantiser_feeds_matcher(po, feature):
    """ex01:92-99: normalise both sets, +128 -> uint8, brute-force match; the device path
    (K7 -> K1, nothing leaves HBM in between) equals numpy quantiser -> oracle matcher."""
    import torch
    from spectavi_b200 import device as dev
    rng = np.random.RandomState(9)
    x = rng.randint(0, 256, (3000, 132)).astype("float32")
    y = np.clip(x[rng.randint(0, 3000, 2000)] + rng.randint(-20, 21, (2000, 132)), 0, 255).astype("float32")
    _x = feature.normalize_to_ubyte_and_multiple_16_dim(x)
    _y = feature.normalize_to_ubyte_and_multiple_16_dim(y)
    oi, od = po.l1k2((_x + 128).astype("uint8"), (_y + 128).astype("uint8"), 4)
    fx, ux = dev.normalize_rows(torch.from_numpy(x).cuda())
    fy, uy = dev.normalize_rows(torch.from_numpy(y).cuda())
    assert (fx.cpu().numpy() == _x).all() and (ux.cpu().numpy() == (_x + 128).astype("uint8")).all()
    keys = dev.new_keys(2000, "cuda")
    dev.l1k2_fold(ux, uy, keys)
    gi, gd = dev.keys_finalize(keys)
    assert (gi.cpu().numpy().view(np.uint64) == oi).all() and (gd.cpu().numpy() == od).all()
