"""Python-3 half of the NdArray output protocol.

Same class name, field layout and ``asarray()`` semantics as the reference's
``cndarray.NdArray`` (ctypes_ndarray/cndarray/ndarray.py:47-122), so code that
does ``NdArray(dtype='uint64')`` / ``.asarray()`` keeps working: the C side
sets ``m_ndim``/``m_shape`` and calls ``m_alloc(nbytes)``, Python answers with
the address of a fresh zero-filled numpy buffer that it owns, and
``asarray()`` is ``view(dtype).reshape(shape).squeeze()`` (so yrows == 1
collapses to 1-D exactly as the reference does, ndarray.py:113-122).

Differences forced by Python 3 (SURVEY.md 8b last row): ``m_dtype`` is stored
as bytes, and buffer addresses come from ``ndarray.ctypes.data`` instead of
numpy's private ``_ctype_ndarray``.
"""
import ctypes as ct

import numpy as np

allocator_type = ct.CFUNCTYPE(ct.c_void_p, ct.c_size_t)
set_dtype_type = ct.CFUNCTYPE(None, ct.c_char_p)


class NdArray(ct.Structure):
    """numpy-owned output array handed to C as ``NdArray*`` (64-byte POD,
    include/spectavi_b200.h; reference ctypes_ndarray/src/NdArray.h:20-29)."""

    _fields_ = [('m_data', ct.c_void_p),
                ('m_ndim', ct.c_int),
                ('m_shape', ct.POINTER(ct.c_uint64)),
                ('m_sizeofdtype', ct.c_int),
                ('m_alloc', allocator_type),
                ('m_realloc', allocator_type),
                ('m_set_dtype', set_dtype_type),
                ('m_dtype', ct.c_char_p)]

    def __init__(self, shape=None, dtype='float64', maxdims=256):
        super().__init__()
        self.arr = None
        self._dtype = np.dtype(dtype)

        def set_dtype(newdtype):
            self._dtype = np.dtype(newdtype.decode())
            self.m_dtype = newdtype
            self.m_sizeofdtype = self._dtype.itemsize

        def allocator(nbytes):
            self.arr = np.zeros((nbytes,), dtype='u1')
            return self.arr.ctypes.data

        def reallocator(nbytes):
            self.arr = np.resize(self.arr, (nbytes,))
            return self.arr.ctypes.data

        # keep the thunks alive for the lifetime of the struct
        self._alloc_cb = allocator_type(allocator)
        self._realloc_cb = allocator_type(reallocator)
        self._set_dtype_cb = set_dtype_type(set_dtype)
        self.m_sizeofdtype = self._dtype.itemsize
        if shape is None:
            self.np_shape = np.zeros(maxdims, dtype='uint64')
            ndim = 0
        else:
            shape = np.asarray(shape)
            self.np_shape = shape.astype('uint64')
            ndim = len(shape)
            nbytes = int(np.prod(shape)) * self.m_sizeofdtype
            self.arr = np.zeros((nbytes,), dtype='u1')
            self.m_data = self.arr.ctypes.data
        self.m_ndim = ndim
        self.m_shape = self.np_shape.ctypes.data_as(ct.POINTER(ct.c_uint64))
        self.m_alloc = self._alloc_cb
        self.m_realloc = self._realloc_cb
        self.m_set_dtype = self._set_dtype_cb
        self.m_dtype = self._dtype.name.encode()

    def asarray(self):
        shape = tuple(int(s) for s in self.np_shape[:self.m_ndim])
        return self.arr.view(self._dtype).reshape(shape).squeeze()
