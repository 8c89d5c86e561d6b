"""CPU-side checks of the drop-in boundary: the library loads without a GPU,
exports every symbol the header declares, and NdArray has the reference's
layout (ctypes_ndarray/src/NdArray.h:20-29).  No compute calls here."""
import ctypes as ct

from abi_symbols import declared_symbols


def test_header_symbols_exported():
    from spectavi_b200._libspectavi import clib
    names = declared_symbols()
    for must in ("nn_bruteforcel1k2", "nn_bruteforce", "nn_bruteforcei", "nn_cascading_hash", "spv_l1k2_device"):
        assert must in names
    missing = [n for n in names if not hasattr(clib, n)]
    assert not missing, missing


def test_ndarray_layout():
    from spectavi_b200.cndarray import NdArray
    assert ct.sizeof(NdArray) == 64
    offs = [getattr(NdArray, f).offset for f, _ in NdArray._fields_]
    assert offs == [0, 8, 16, 24, 32, 40, 48, 56]


def test_ndarray_protocol_squeeze():
    import numpy as np
    from spectavi_b200.cndarray import NdArray
    a = NdArray(dtype="int32")
    a.m_ndim = 2
    a.np_shape[0], a.np_shape[1] = 1, 2
    addr = a._alloc_cb(8)
    assert addr == a.arr.ctypes.data
    out = a.asarray()
    assert out.shape == (2,) and out.dtype == np.int32
