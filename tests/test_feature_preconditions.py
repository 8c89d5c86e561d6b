"""The C entry points abort() on a violated precondition (the reference's failures are C++ exceptions escaping
extern "C", src/BruteForceNnL1K2.h:74-81); the Python wrappers therefore raise ValueError BEFORE the ctypes call,
so a bad argument never kills the interpreter.  No GPU needed: nothing reaches the library."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def feature():
    from spectavi_b200 import feature
    return feature


def test_l1k2_dim_not_multiple_of_16(feature):
    a = np.zeros((4, 20), "uint8")
    with pytest.raises(ValueError):
        feature.nn_bruteforcel1k2(a, a)
    with pytest.raises(ValueError):
        feature.match_bruteforcel1k2(a, a)


@pytest.mark.parametrize("kw", [dict(m=40), dict(m=0), dict(m=8, k=1), dict(m=14, g=13, n=2), dict(m=8, n=0),
                                dict(m=8, g=-1), dict(m=14, n=(1 << 12) + 1, g=12)])
def test_cascade_parameters(feature, kw):
    a = np.zeros((400, 32), "float32")
    with pytest.raises(ValueError):
        feature.nn_cascading_hash(a, a, **kw)


def test_cascade_dim(feature):
    a = np.zeros((400, 24), "float32")
    with pytest.raises(ValueError):
        feature.nn_cascading_hash(a, a, m=8)
    with pytest.raises(ValueError):
        feature.match_cascading_hash(a, a, m=8)


def test_bruteforce_k(feature):
    a = np.zeros((10, 8), "float32")
    with pytest.raises(ValueError):
        feature.nn_bruteforce(a, a, k=0)
    with pytest.raises(ValueError):
        feature.nn_bruteforce_filtered(a, a, [0], [0], k=129)
    with pytest.raises(ValueError):
        feature.kmedians_find_nearest_neighbours(a, a, a[:2], np.zeros(10, "int32"), a[:2], np.zeros(10, "int32"), c=1, k=200)
