"""Host-side mirror of the reference's NN wrappers (spectavi/feature.py:202-304,
340-407): same names, argument meaning, dtypes and return shapes; the ctypes
prototypes are the reference's own (feature.py:208-243,346-358), bound to the
B200 ``libspectavi.so``.  Everything below the ctypes call runs on the GPU;
this module adds no CPU compute beyond what the reference's wrappers do in
numpy themselves (``use_int`` rounding, the auto-``m`` rule, the quantiser).
"""
import ctypes as ct

import numpy as np
from numpy.ctypeslib import ndpointer

from ._libspectavi import clib
from .cndarray import NdArray

_nn_bruteforce = clib.nn_bruteforce
_nn_bruteforce.restype = None
_nn_bruteforce.argtypes = [ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                           ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                           ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                           ct.c_float, ct.c_float,
                           ct.POINTER(NdArray), ct.POINTER(NdArray)]

_nn_bruteforcei = clib.nn_bruteforcei
_nn_bruteforcei.restype = None
_nn_bruteforcei.argtypes = [ndpointer(ct.c_int, flags="C_CONTIGUOUS"),
                            ndpointer(ct.c_int, flags="C_CONTIGUOUS"),
                            ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                            ct.c_float, ct.c_float,
                            ct.POINTER(NdArray), ct.POINTER(NdArray)]

_nn_bruteforcel1k2 = clib.nn_bruteforcel1k2
_nn_bruteforcel1k2.restype = None
_nn_bruteforcel1k2.argtypes = [ndpointer(ct.c_ubyte, flags="C_CONTIGUOUS"),
                               ndpointer(ct.c_ubyte, flags="C_CONTIGUOUS"),
                               ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                               ct.POINTER(NdArray), ct.POINTER(NdArray)]

_nn_cascading_hash = clib.nn_cascading_hash
_nn_cascading_hash.restype = None
_nn_cascading_hash.argtypes = [ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                               ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                               ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                               ct.c_int, ct.c_int, ct.c_int,
                               ct.POINTER(NdArray), ct.POINTER(NdArray)]


_match_l1k2 = clib.spv_match_bruteforcel1k2
_match_l1k2.restype = None
_match_l1k2.argtypes = [ndpointer(ct.c_ubyte, flags="C_CONTIGUOUS"),
                        ndpointer(ct.c_ubyte, flags="C_CONTIGUOUS"),
                        ct.c_int, ct.c_int, ct.c_int, ct.c_double,
                        ct.POINTER(NdArray)]

_match_cashash = clib.spv_match_cascading_hash
_match_cashash.restype = None
_match_cashash.argtypes = [ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                           ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                           ct.c_int, ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                           ct.c_int, ct.c_double, ct.POINTER(NdArray)]


_normalize = clib.spv_normalize_to_ubyte
_normalize.restype = None
_normalize.argtypes = [ndpointer(ct.c_float, flags="C_CONTIGUOUS"), ct.c_int, ct.c_int, ct.POINTER(NdArray)]


_u64list = ndpointer(ct.c_uint64, flags="C_CONTIGUOUS")
_nn_filtered = clib.spv_nn_bruteforce_filtered
_nn_filtered.restype = None
_nn_filtered.argtypes = [ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                         ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                         ct.c_int, ct.c_int, ct.c_int, ct.c_int, ct.c_float, ct.c_float,
                         _u64list, ct.c_int, _u64list, ct.c_int,
                         ndpointer(ct.c_uint64, flags="C_CONTIGUOUS"),
                         ndpointer(ct.c_float, flags="C_CONTIGUOUS")]

_kmedians_fnn = clib.spv_kmedians_find_nearest_neighbours
_kmedians_fnn.restype = None
_kmedians_fnn.argtypes = [ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                          ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                          ct.c_int, ct.c_int, ct.c_int,
                          ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                          ndpointer(ct.c_int32, flags="C_CONTIGUOUS"), ct.c_int,
                          ndpointer(ct.c_float, flags="C_CONTIGUOUS"),
                          ndpointer(ct.c_int32, flags="C_CONTIGUOUS"), ct.c_int,
                          ct.c_int, ct.c_int, ct.POINTER(NdArray), ct.POINTER(NdArray)]


LP_MAX_K_FILTERED = 128     # csrc/spv_common.h LP_MAX_K: the filter overload runs single-pass

clib.spv_cascade_ws_bytes.restype = ct.c_size_t
clib.spv_cascade_ws_bytes.argtypes = [ct.c_int] * 6


def _check_l1_dim(dim):
    # BruteForceNnL1K2.h:78-81 throws (a C++ exception escaping extern "C", i.e. abort); the C entry point here
    # aborts too, so the precondition is raised in Python first
    if dim <= 0 or dim % 16 != 0:
        raise ValueError("data dimension (%d) must be a positive multiple of 16" % dim)


def _check_cascade(xrows, yrows, dim, k, m, n, g):
    _check_l1_dim(dim)
    if k < 2:
        raise ValueError("k = %d: the cascade matcher always produces 2 neighbours" % k)
    if not 1 <= m <= 31:
        raise ValueError("m = %d outside [1, 31] (codes are int32, CascadingHashNn.h:102-111)" % m)
    if n < 1 or g < 0:
        raise ValueError("need n >= 1 and g >= 0")
    if clib.spv_cascade_ws_bytes(xrows, yrows, dim, m, n, g) == 0:
        raise ValueError("unsupported cascade parameters dim=%d m=%d n=%d g=%d (min(g, m) <= 12, n * 2^g <= 2^24, "
                         "n * rows < 2^31, and the hash tables of one pass must fit shared memory)" % (dim, m, n, g))


def nn_bruteforce(x, y, p=.5, mu=0., k=2, use_int=False):
    """k nearest rows of ``x`` for every row of ``y`` under sum |d|^p
    (reference feature.py:246-289).  Returns ``(idx uint64, dist)`` of shape
    ``(yrows, k)``, ascending; ``dist`` is float32, or int32 with ``use_int``
    (inputs rounded to ``round(100 * v)`` as the reference does, :286-287)."""
    xrows, xdim = x.shape
    yrows, ydim = y.shape
    assert ydim == xdim
    dim = xdim
    if k < 1:
        raise ValueError("k must be >= 1")
    nn_idx = NdArray(dtype='uint64')
    if not use_int:
        nn_dist = NdArray(dtype='float32')
        _nn_bruteforce(x, y, xrows, yrows, dim, k, p, mu, nn_idx, nn_dist)
    else:
        nn_dist = NdArray(dtype='int32')
        xi = np.round(100 * x).astype('int32')
        yi = np.round(100 * y).astype('int32')
        _nn_bruteforcei(xi, yi, xrows, yrows, dim, k, p, mu, nn_idx, nn_dist)
    return nn_idx.asarray(), nn_dist.asarray()


def nn_bruteforce_filtered(x, y, filter_x, filter_y, p=1., mu=0., k=2, nn_idx=None, nn_dist=None):
    """``BruteForceNn::find_neighbours(out_idx, out_dist, filter_x, filter_y, k)``
    (reference src/BruteForceNn.h:37-120): k nearest rows of ``x`` among
    ``filter_x`` for every row of ``y`` in ``filter_y``; an empty filter means all
    rows.  ``nn_idx`` / ``nn_dist`` (yrows x k) are updated in place on the rows of
    ``filter_y`` (zeros are allocated when omitted) and returned."""
    xrows, dim = x.shape
    yrows, ydim = y.shape
    assert ydim == dim
    if nn_idx is None:
        nn_idx = np.zeros((yrows, k), 'uint64')
    if nn_dist is None:
        nn_dist = np.zeros((yrows, k), 'float32')
    assert nn_idx.shape == (yrows, k) and nn_dist.shape == (yrows, k)
    if not 1 <= k <= LP_MAX_K_FILTERED:
        raise ValueError("k = %d outside [1, %d]" % (k, LP_MAX_K_FILTERED))
    fx = np.ascontiguousarray(np.asarray(filter_x, 'uint64').ravel())
    fy = np.ascontiguousarray(np.asarray(filter_y, 'uint64').ravel())
    if (fx.size and fx.max() >= xrows) or (fy.size and fy.max() >= yrows):
        raise ValueError("filter holds a row id outside the matrix")   # the C entry point aborts on this
    _nn_filtered(x, y, xrows, yrows, dim, k, p, mu, fx, fx.size, fy, fy.size, nn_idx, nn_dist)
    return nn_idx, nn_dist


def kmedians_find_nearest_neighbours(x, y, medians_x, labels_x, medians_y, labels_y, c=5, k=2):
    """``KMedians::find_nearest_neighbours`` (reference src/KMedians.h:260-295),
    the matching half of ``nn_kmedians`` (feature.py:328-337), given both
    clusterings: for every cluster of ``y`` the ``c`` nearest medians of ``x``
    (L1), then the k nearest rows (L1) of each of its rows among the members of
    those clusters.  ``labels_*``: cluster id of every row; ``medians_*``:
    (clusters, dim) float32.  Returns ``(idx uint64, dist float32)``, (yrows, k)."""
    xrows, dim = x.shape
    yrows, ydim = y.shape
    assert ydim == dim and medians_x.shape[1] == dim and medians_y.shape[1] == dim
    lx = np.ascontiguousarray(labels_x, 'int32')
    ly = np.ascontiguousarray(labels_y, 'int32')
    assert lx.shape == (xrows,) and ly.shape == (yrows,)
    if lx.min() < 0 or lx.max() >= medians_x.shape[0] or ly.min() < 0 or ly.max() >= medians_y.shape[0]:
        raise ValueError("label outside [0, number of medians)")           # the C entry point aborts on this
    if not 1 <= c <= min(medians_x.shape[0], LP_MAX_K_FILTERED):
        raise ValueError("c must be in [1, min(number of train medians, %d)]" % LP_MAX_K_FILTERED)
    if not 1 <= k <= LP_MAX_K_FILTERED:
        raise ValueError("k = %d outside [1, %d]" % (k, LP_MAX_K_FILTERED))
    nn_idx = NdArray(dtype='uint64')
    nn_dist = NdArray(dtype='float32')
    _kmedians_fnn(x, y, xrows, yrows, dim, np.ascontiguousarray(medians_x, 'float32'), lx, medians_x.shape[0],
                  np.ascontiguousarray(medians_y, 'float32'), ly, medians_y.shape[0], c, k, nn_idx, nn_dist)
    return nn_idx.asarray(), nn_dist.asarray()


def nn_bruteforcel1k2(x, y, nthreads=1):
    """Exact L1 2-nearest-neighbour of every ``y`` row among ``x`` rows; uint8
    inputs, ``dim % 16 == 0`` (reference feature.py:292-304).  ``nthreads`` is
    kept for call compatibility and ignored (the work runs on the GPU)."""
    xrows, xdim = x.shape
    yrows, ydim = y.shape
    assert ydim == xdim
    dim = xdim
    _check_l1_dim(dim)
    nn_idx = NdArray(dtype='uint64')
    nn_dist = NdArray(dtype='int32')
    _nn_bruteforcel1k2(x, y, xrows, yrows, dim, nthreads, nn_idx, nn_dist)
    return nn_idx.asarray(), nn_dist.asarray()


def nn_cascading_hash(x, y, k=2, m=None, n=2, g=2):
    """Cascade-of-hashes matcher (reference feature.py:360-376): ``n`` tables of
    ``m``-bit random-hyperplane codes, multi-probe on the ``g`` least confident
    bits, exact L1 k=2 re-rank of the candidates.  ``m=None`` applies the
    reference's ~6-rows-per-bucket rule and falls back to brute force (int32
    distances, as the reference) when that gives fewer than 4 bits."""
    xrows, xdim = x.shape
    yrows, ydim = y.shape
    assert ydim == xdim
    if m is None:
        mrows = max([xrows, yrows])
        m = int(np.floor(np.log2(mrows / 6.)))
        if m < 4:
            return nn_bruteforcel1k2((x + 128).astype('uint8'),
                                     (y + 128).astype('uint8'), nthreads=8)
    dim = xdim
    _check_cascade(xrows, yrows, dim, k, m, n, g)
    cashash_idx = NdArray(dtype='uint64')
    cashash_dist = NdArray(dtype='float32')
    _nn_cascading_hash(x, y, xrows, yrows, dim, k, m, n, g, cashash_idx,
                       cashash_dist)
    return cashash_idx.asarray(), cashash_dist.asarray()


def normalize_to_ubyte_and_multiple_16_dim(x, dtype='float32'):
    """Column de-mean, scale to [-128, 127], round half-to-even, zero-pad the
    columns to a multiple of 16 (reference feature.py:384-407; host numpy in
    the reference too -- a constant column yields NaN there and here)."""
    x0 = x - np.mean(x, axis=0, keepdims=True)
    hi = np.max(x0, axis=0, keepdims=True)
    lo = np.min(x0, axis=0, keepdims=True)
    norm = np.max(np.stack([hi, -lo]), axis=0)
    x0 = np.round(x0 / norm * 128)
    x0[x0 > 127] = 127
    x0[x0 < -128] = -128
    xrows, dim = x0.shape
    new_dim = int(np.ceil(dim / 16.) * 16)
    xx = np.zeros([xrows, new_dim])
    xx[:, :dim] = x0
    return xx.astype(dtype)


def normalize_to_ubyte_and_multiple_16_dim_gpu(x):
    """``normalize_to_ubyte_and_multiple_16_dim`` for a float32 matrix, computed on
    the GPU (K7) and bit-identical to the numpy function above."""
    x = np.ascontiguousarray(x, dtype='float32')
    rows, dim = x.shape
    out = NdArray(dtype='float32')
    _normalize(x, rows, dim, out)
    return out.arr.view('float32').reshape(rows, -1)


def ratio_test(nn_dist, min_ratio=1.75):
    """The example's ratio test (example/ex01_essential_estimation.py:102-103):
    ``second / first >= min_ratio`` in float64; x/0 -> inf passes, 0/0 -> nan
    fails."""
    with np.errstate(divide='ignore', invalid='ignore'):
        ratio = nn_dist[:, 1] / nn_dist[:, 0].astype('float64')
    return ratio >= min_ratio


def _split_matches(out):
    # not NdArray.asarray(): its squeeze() (cndarray/ndarray.py:113-122) would drop the pair axis of a single match
    m = out.arr.view('uint64').reshape(-1, 2)
    return m[:, 0].copy(), m[:, 1].copy()


def match_bruteforcel1k2(x, y, min_ratio=1.75):
    """Brute-force L1 k=2 match + ratio test + compaction on the GPU: what
    example/ex01_essential_estimation.py:95-105 computes with
    ``nn_bruteforcel1k2`` followed by the numpy ratio test, returned as
    ``(idx_x, idx_y)`` -- ``idx0[pass_idx]`` and ``arange(yrows)[pass_idx]``."""
    xrows, xdim = x.shape
    yrows, ydim = y.shape
    assert ydim == xdim
    _check_l1_dim(xdim)
    out = NdArray(dtype='uint64')
    _match_l1k2(x, y, xrows, yrows, xdim, float(min_ratio), out)
    return _split_matches(out)


def match_cascading_hash(x, y, min_ratio=1.75, m=None, n=2, g=2):
    """Cascade-of-hashes match + ratio test + compaction on the GPU (the
    'cascading-hash' branch of example/ex01_essential_estimation.py:100-105);
    same ``m`` rule and brute-force fallback as ``nn_cascading_hash``."""
    xrows, xdim = x.shape
    yrows, ydim = y.shape
    assert ydim == xdim
    if m is None:
        mrows = max([xrows, yrows])
        m = int(np.floor(np.log2(mrows / 6.)))
        if m < 4:
            return match_bruteforcel1k2((x + 128).astype('uint8'),
                                        (y + 128).astype('uint8'), min_ratio)
    _check_cascade(xrows, yrows, xdim, 2, m, n, g)
    out = NdArray(dtype='uint64')
    _match_cashash(x, y, xrows, yrows, xdim, m, n, g, float(min_ratio), out)
    return _split_matches(out)
