"""ctypes front-ends for the CPU checkers (TEST INFRASTRUCTURE ONLY).

Three independent checkers for the descriptor-matching path:

* ``liboracle.so``  -- C restatement (oracle/spv_oracle.c) of
  src/BruteForceNnL1K2.h, src/BruteForceNn.h, src/CascadingHashNn.h.
* ``_ref/libspectavi_ref.so`` -- the reference's own headers compiled from
  /root/reference through oracle/eigen_shim (oracle/ref_driver.cpp).  Present
  only where it was built (this container, or shipped prebuilt to the GPU box).
* numpy -- stable lexicographic (dist, idx) sort, the independent cross-check
  (restates test/test_feature.py:10-26 with a *stable* sort).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs import this module.  The product never does.
"""
import ctypes as ct
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(_HERE, "liboracle.so")
_REF_SO = os.path.join(_HERE, "_ref", "libspectavi_ref.so")

NONE_IDX = np.uint64(0xFFFFFFFFFFFFFFFF)
INT_MAX = np.int32(2**31 - 1)


def build(quiet=True):
    """Compile liboracle.so (and _ref when /root/reference is present)."""
    out = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True,
                         text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if not quiet:
        print(out.stdout)


_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
_u64p = np.ctypeslib.ndpointer(np.uint64, flags="C_CONTIGUOUS")

_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_ORACLE_SO):
            build()
        L = ct.CDLL(_ORACLE_SO)
        L.oracle_l1k2.restype = ct.c_int
        L.oracle_l1k2.argtypes = [_u8p, _u8p, ct.c_int, ct.c_int, ct.c_int,
                                  ct.c_int, ct.c_int, _u64p, _i32p]
        L.oracle_l1k2_filtered.restype = ct.c_int
        L.oracle_l1k2_filtered.argtypes = [_u8p, _u8p, ct.c_int, ct.c_int,
                                           ct.c_int, _i64p, _i32p, ct.c_int,
                                           _u64p, _i32p]
        L.oracle_bruteforce.restype = ct.c_int
        L.oracle_bruteforce.argtypes = [ct.c_void_p, ct.c_void_p, ct.c_int,
                                        ct.c_int, ct.c_int, ct.c_int,
                                        ct.c_float, ct.c_float, ct.c_int,
                                        _u64p, ct.c_void_p]
        L.oracle_bruteforce_filtered.restype = ct.c_int
        L.oracle_bruteforce_filtered.argtypes = [_f32p, _f32p, ct.c_int,
                                                 ct.c_int, ct.c_int, ct.c_int,
                                                 ct.c_float, ct.c_float,
                                                 _u64p, ct.c_int, _u64p,
                                                 ct.c_int, _u64p, _f32p]
        L.oracle_hashcodes.restype = ct.c_int
        L.oracle_hashcodes.argtypes = [_f32p, ct.c_int, ct.c_int, _f32p,
                                       ct.c_int, ct.c_int, _i32p]
        L.oracle_cascade.restype = ct.c_int
        L.oracle_cascade.argtypes = [_f32p, _f32p, ct.c_int, ct.c_int,
                                     ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                                     _f32p, ct.c_int, _u64p, _f32p,
                                     ct.c_void_p]
        L.oracle_probe_codes.restype = ct.c_int
        L.oracle_probe_codes.argtypes = [_f32p, ct.c_int, _f32p, ct.c_int,
                                         ct.c_int, _i32p]
        L.oracle_ratio_mask_i32.restype = None
        L.oracle_ratio_mask_i32.argtypes = [_i32p, ct.c_int, ct.c_double, _u8p]
        L.oracle_ratio_mask_f32.restype = None
        L.oracle_ratio_mask_f32.argtypes = [_f32p, ct.c_int, ct.c_double, _u8p]
        _lib = L
    return _lib


def have_ref():
    return os.path.exists(_REF_SO)


def ref():
    """The reference's own code (oracle/_ref).  Raises if it was never built."""
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libspectavi_ref.so not built "
                               "(needs /root/reference at build time)")
        R = ct.CDLL(_REF_SO)
        R.ref_l1k2.restype = ct.c_int
        R.ref_l1k2.argtypes = [_u8p, _u8p, ct.c_int, ct.c_int, ct.c_int,
                               ct.c_int, _u64p, _i32p]
        R.ref_bruteforce_f32.restype = ct.c_int
        R.ref_bruteforce_f32.argtypes = [_f32p, _f32p, ct.c_int, ct.c_int,
                                         ct.c_int, ct.c_int, ct.c_float,
                                         ct.c_float, _u64p, _f32p]
        R.ref_bruteforce_i32.restype = ct.c_int
        R.ref_bruteforce_i32.argtypes = [_i32p, _i32p, ct.c_int, ct.c_int,
                                         ct.c_int, ct.c_int, ct.c_float,
                                         ct.c_float, _u64p, _i32p]
        R.ref_bruteforce_filtered_f32.restype = ct.c_int
        R.ref_bruteforce_filtered_f32.argtypes = [_f32p, _f32p, ct.c_int,
                                                  ct.c_int, ct.c_int, ct.c_int,
                                                  ct.c_float, ct.c_float,
                                                  _u64p, ct.c_int, _u64p,
                                                  ct.c_int, _u64p, _f32p]
        R.ref_kmedians_nn.restype = ct.c_int
        R.ref_kmedians_nn.argtypes = [_f32p, _f32p, ct.c_int, ct.c_int,
                                      ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                                      ct.c_int, ct.c_uint, _u64p, _f32p,
                                      _i32p, _i32p, _f32p, _f32p,
                                      ct.POINTER(ct.c_double)]
        R.ref_cascade.restype = ct.c_int
        R.ref_cascade.argtypes = [_f32p, _f32p, ct.c_int, ct.c_int, ct.c_int,
                                  ct.c_int, ct.c_int, ct.c_int, ct.c_int,
                                  ct.c_uint, _u64p, _f32p, ct.c_void_p,
                                  ct.c_void_p, ct.c_void_p]
        R.ref_cascade_candidates.restype = ct.c_int
        R.ref_cascade_candidates.argtypes = [_f32p, _f32p, ct.c_int, ct.c_int,
                                             ct.c_int, ct.c_int, ct.c_int,
                                             ct.c_int, ct.c_uint, ct.c_int,
                                             _i32p, ct.c_int]
        _ref = R
    return _ref


# --------------------------------------------------------------------------- C
def l1k2(x, y, nthreads=1, prune=True):
    """oracle_l1k2: restates BruteForceNnL1K2.h:84-145 (u8, k=2)."""
    x = np.ascontiguousarray(x, np.uint8)
    y = np.ascontiguousarray(y, np.uint8)
    assert x.shape[1] == y.shape[1]
    idx = np.zeros((y.shape[0], 2), np.uint64)
    dist = np.zeros((y.shape[0], 2), np.int32)
    rc = lib().oracle_l1k2(x, y, x.shape[0], y.shape[0], x.shape[1], nthreads,
                           int(prune), idx, dist)
    if rc != 0:
        raise ValueError("oracle_l1k2: dim %% 16 != 0 (reference throws)")
    return idx, dist


def l1k2_filtered(x, y, offsets, cands, nthreads=1):
    x = np.ascontiguousarray(x, np.uint8)
    y = np.ascontiguousarray(y, np.uint8)
    idx = np.zeros((y.shape[0], 2), np.uint64)
    dist = np.zeros((y.shape[0], 2), np.int32)
    rc = lib().oracle_l1k2_filtered(
        x, y, x.shape[0], y.shape[0], x.shape[1],
        np.ascontiguousarray(offsets, np.int64),
        np.ascontiguousarray(cands, np.int32), nthreads, idx, dist)
    assert rc == 0
    return idx, dist


def bruteforce(x, y, k=2, p=.5, mu=0., use_int=False):
    """oracle_bruteforce: restates BruteForceNn.h:37-120.  Output buffers start
    as zeros like the numpy-allocated ones of the reference wrapper."""
    if use_int:
        x = np.ascontiguousarray(x, np.int32)
        y = np.ascontiguousarray(y, np.int32)
        dist = np.zeros((y.shape[0], k), np.int32)
    else:
        x = np.ascontiguousarray(x, np.float32)
        y = np.ascontiguousarray(y, np.float32)
        dist = np.zeros((y.shape[0], k), np.float32)
    idx = np.zeros((y.shape[0], k), np.uint64)
    rc = lib().oracle_bruteforce(x.ctypes.data, y.ctypes.data, x.shape[0],
                                 y.shape[0], x.shape[1], k, p, mu,
                                 int(use_int), idx, dist.ctypes.data)
    assert rc == 0
    return idx, dist


def _flist(f):
    return np.ascontiguousarray(np.asarray(f, np.uint64).ravel())


def bruteforce_filtered(x, y, filter_x, filter_y, k=2, p=1., mu=0., idx=None,
                        dist=None):
    """oracle_bruteforce_filtered: BruteForceNn.h:37-120 with explicit filter
    lists (empty = all rows).  idx / dist are in-out (zeros when omitted)."""
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    idx = np.zeros((y.shape[0], k), np.uint64) if idx is None else idx
    dist = np.zeros((y.shape[0], k), np.float32) if dist is None else dist
    fx, fy = _flist(filter_x), _flist(filter_y)
    rc = lib().oracle_bruteforce_filtered(x, y, x.shape[0], y.shape[0],
                                          x.shape[1], k, p, mu, fx, fx.size,
                                          fy, fy.size, idx, dist)
    assert rc == 0
    return idx, dist


def kmedians_find_nearest_neighbours(x, y, medians_x, labels_x, medians_y,
                                     labels_y, c=5, k=2, idx=None, dist=None):
    """KMedians::find_nearest_neighbours (src/KMedians.h:260-295) given the two
    clusterings: x = `other` (train side), y = `this` (query side), as
    nn_kmedians passes them (Spectavi.cpp:318).  For every query cluster: the c
    nearest train medians under L1 (:268-273), then the filtered brute-force
    search of the cluster's rows against the union of those train clusters
    (:275-293), groups in ascending order.  An empty filter means ALL rows to
    BruteForceNn (BruteForceNn.h:45-46) and that quirk is kept."""
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    labels_x = np.asarray(labels_x)
    labels_y = np.asarray(labels_y)
    nmx, nmy = medians_x.shape[0], medians_y.shape[0]
    idx = np.zeros((y.shape[0], k), np.uint64) if idx is None else idx
    dist = np.zeros((y.shape[0], k), np.float32) if dist is None else dist
    near, _ = bruteforce(medians_x, medians_y, k=c, p=1., mu=0.)
    xl = [np.flatnonzero(labels_x == g) for g in range(nmx)]
    for g in range(nmy):
        fy = np.flatnonzero(labels_y == g)
        fx = np.concatenate([xl[int(j)] for j in near[g]]) if c else np.zeros(0)
        bruteforce_filtered(x, y, fx, fy, k, 1., 0., idx, dist)
    return idx, dist, near


def hashcodes(data, dictionary):
    """oracle_hashcodes: CascadingHashNn.h:102-135. dictionary [n][dim][m]."""
    data = np.ascontiguousarray(data, np.float32)
    d = np.ascontiguousarray(dictionary, np.float32)
    n, dim, m = d.shape
    codes = np.zeros((data.shape[0], n), np.int32)
    rc = lib().oracle_hashcodes(data, data.shape[0], dim, d, m, n, codes)
    assert rc == 0
    return codes


def probe_codes(yrow, dict_t, g):
    yrow = np.ascontiguousarray(yrow, np.float32)
    d = np.ascontiguousarray(dict_t, np.float32)
    dim, m = d.shape
    out = np.zeros(1 << g, np.int32)
    rc = lib().oracle_probe_codes(yrow, dim, d, m, g, out)
    assert rc == 0
    return out


def cascade(x, y, dictionary, g, nthreads=1, want_counts=False):
    """oracle_cascade: CascadingHashNn.h:187-245 with an injected dictionary."""
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    d = np.ascontiguousarray(dictionary, np.float32)
    n, dim, m = d.shape
    assert dim == x.shape[1] == y.shape[1]
    idx = np.zeros((y.shape[0], 2), np.uint64)
    dist = np.zeros((y.shape[0], 2), np.float32)
    cnt = np.zeros(y.shape[0], np.int32)
    rc = lib().oracle_cascade(x, y, x.shape[0], y.shape[0], dim, m, n, g, d,
                              nthreads, idx, dist, cnt.ctypes.data)
    if rc != 0:
        raise ValueError("oracle_cascade: bad parameters")
    return (idx, dist, cnt) if want_counts else (idx, dist)


def ratio_mask(dist, min_ratio=1.75):
    """ex01_essential_estimation.py:102-103 evaluated by the C oracle."""
    dist = np.ascontiguousarray(dist)
    mask = np.zeros(dist.shape[0], np.uint8)
    if dist.dtype == np.int32:
        lib().oracle_ratio_mask_i32(dist, dist.shape[0], min_ratio, mask)
    else:
        lib().oracle_ratio_mask_f32(dist.astype(np.float32), dist.shape[0],
                                    min_ratio, mask)
    return mask.astype(bool)


# ------------------------------------------------------------------- reference
def ref_l1k2(x, y, nthreads=1):
    x = np.ascontiguousarray(x, np.uint8)
    y = np.ascontiguousarray(y, np.uint8)
    idx = np.zeros((y.shape[0], 2), np.uint64)
    dist = np.zeros((y.shape[0], 2), np.int32)
    rc = ref().ref_l1k2(x, y, x.shape[0], y.shape[0], x.shape[1], nthreads,
                        idx, dist)
    if rc != 0:
        raise ValueError("reference threw (dim mismatch / dim % 16)")
    return idx, dist


def ref_bruteforce(x, y, k=2, p=.5, mu=0., use_int=False):
    idx = np.zeros((y.shape[0], k), np.uint64)
    if use_int:
        x = np.ascontiguousarray(x, np.int32)
        y = np.ascontiguousarray(y, np.int32)
        dist = np.zeros((y.shape[0], k), np.int32)
        ref().ref_bruteforce_i32(x, y, x.shape[0], y.shape[0], x.shape[1], k,
                                 p, mu, idx, dist)
    else:
        x = np.ascontiguousarray(x, np.float32)
        y = np.ascontiguousarray(y, np.float32)
        dist = np.zeros((y.shape[0], k), np.float32)
        ref().ref_bruteforce_f32(x, y, x.shape[0], y.shape[0], x.shape[1], k,
                                 p, mu, idx, dist)
    return idx, dist


def ref_bruteforce_filtered(x, y, filter_x, filter_y, k=2, p=1., mu=0.,
                            idx=None, dist=None):
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    idx = np.zeros((y.shape[0], k), np.uint64) if idx is None else idx
    dist = np.zeros((y.shape[0], k), np.float32) if dist is None else dist
    fx, fy = _flist(filter_x), _flist(filter_y)
    ref().ref_bruteforce_filtered_f32(x, y, x.shape[0], y.shape[0], x.shape[1],
                                      k, p, mu, fx, fx.size, fy, fy.size, idx,
                                      dist)
    return idx, dist


def ref_kmedians_nn(x, y, nmx, nmy, c=5, k=2, seed=7, want_seconds=False):
    """The reference's nn_kmedians (Spectavi.cpp:305-319), shuffle seed pinned,
    plus the clustering it used: (idx, dist, labels_x, labels_y, medians_x,
    medians_y[, seconds spent in find_nearest_neighbours alone])."""
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    idx = np.zeros((y.shape[0], k), np.uint64)
    dist = np.zeros((y.shape[0], k), np.float32)
    xl = np.full(x.shape[0], -1, np.int32)
    yl = np.full(y.shape[0], -1, np.int32)
    xm = np.zeros((nmx, x.shape[1]), np.float32)
    ym = np.zeros((nmy, x.shape[1]), np.float32)
    sec = ct.c_double(0.)
    rc = ref().ref_kmedians_nn(x, y, x.shape[0], y.shape[0], x.shape[1], nmx,
                               nmy, c, k, seed, idx, dist, xl, yl, xm, ym,
                               ct.byref(sec))
    assert rc == 0
    if want_seconds:
        return idx, dist, xl, yl, xm, ym, sec.value
    return idx, dist, xl, yl, xm, ym


def ref_cascade(x, y, m, n, g, seed=1234, k=2):
    """Reference CascadingHashNn with its random_device pinned to `seed`.
    Returns idx, dist, dictionary [n][dim][m], x codes [xrows][n], counts."""
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    dim = x.shape[1]
    idx = np.zeros((y.shape[0], k), np.uint64)
    dist = np.zeros((y.shape[0], k), np.float32)
    d = np.zeros((n, dim, m), np.float32)
    codes = np.zeros((x.shape[0], n), np.int32)
    cnt = np.zeros(y.shape[0], np.int32)
    rc = ref().ref_cascade(x, y, x.shape[0], y.shape[0], dim, k, m, n, g, seed,
                           idx, dist, d.ctypes.data, codes.ctypes.data,
                           cnt.ctypes.data)
    if rc != 0:
        raise ValueError("reference threw")
    return idx, dist, d, codes, cnt


# ----------------------------------------------------------------------- numpy
def numpy_topk(x, y, k=2, p=1, batch=256):
    """Independent oracle: all-pairs distances, then the k lexicographically
    smallest (dist, idx) per query via a stable argsort.  Integer inputs are
    summed in int64 (exact); float inputs in float64 (not bit-equal to the
    reference's float32 column-order sum -- use for index checks on tie-free
    data, as test/test_feature.py:67-81 does)."""
    x = np.asarray(x)
    y = np.asarray(y)
    integer = np.issubdtype(x.dtype, np.integer)
    xa = x.astype(np.int64 if integer else np.float64)
    ya = y.astype(np.int64 if integer else np.float64)
    idx = np.full((y.shape[0], k), NONE_IDX, np.uint64)
    dist = np.zeros((y.shape[0], k), xa.dtype)
    for s in range(0, y.shape[0], batch):
        diff = np.abs(xa[None, :, :] - ya[s:s + batch, None, :])
        if p == 1:
            d = diff.sum(-1)
        elif p == 2:
            d = (diff * diff).sum(-1)
        else:
            d = (diff ** p).sum(-1)
        order = np.argsort(d, axis=1, kind="stable")[:, :k]
        kk = order.shape[1]
        idx[s:s + batch, :kk] = order
        dist[s:s + batch, :kk] = np.take_along_axis(d, order, axis=1)
    return idx, dist


def numpy_l1k2(x, y, batch=256):
    """numpy_topk specialised to the BruteForceNnL1K2 output convention:
    missing neighbours are (INT_MAX, size_t(-1)), BruteForceNnL1K2.h:99-103."""
    idx, dist = numpy_topk(np.asarray(x, np.uint8), np.asarray(y, np.uint8),
                           2, 1, batch)
    dist = dist.astype(np.int64)
    n = np.asarray(x).shape[0]
    if n < 2:
        dist[:, n:] = int(INT_MAX)
    return idx, dist.astype(np.int32)
