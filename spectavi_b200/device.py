"""Device-resident API over torch tensors (Part 2 of include/spectavi_b200.h).

torch is plumbing here: it owns the device memory, the CUDA stream and -- for
the one-process-per-GPU layout -- the NCCL process group.  Every computation
is a kernel of libspectavi.so launched on torch's current stream; nothing in
this module computes with torch ops.
"""
import ctypes as ct

import torch

from ._libspectavi import clib

_vp = ct.c_void_p
clib.spv_keys_reset.restype = ct.c_int
clib.spv_keys_reset.argtypes = [_vp, ct.c_int, _vp]
clib.spv_l1k2_device.restype = ct.c_int
clib.spv_l1k2_device.argtypes = [_vp, _vp, ct.c_int, ct.c_int, ct.c_int, ct.c_uint32, _vp, _vp]
clib.spv_keys_merge.restype = ct.c_int
clib.spv_keys_merge.argtypes = [_vp, _vp, ct.c_int, ct.c_int, _vp]
clib.spv_keys_finalize_i32.restype = ct.c_int
clib.spv_keys_finalize_i32.argtypes = [_vp, ct.c_int, ct.c_int, _vp, _vp, _vp]
clib.spv_keys_finalize_f32.restype = ct.c_int
clib.spv_keys_finalize_f32.argtypes = [_vp, ct.c_int, ct.c_int, _vp, _vp, _vp]
clib.spv_launch_count.restype = ct.c_uint64
clib.spv_launch_count.argtypes = []
clib.spv_last_error.restype = ct.c_char_p
clib.spv_last_error.argtypes = []


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed: %s" % (what, clib.spv_last_error().decode()))


def _stream():
    return ct.c_void_p(torch.cuda.current_stream().cuda_stream)


def launch_count():
    return int(clib.spv_launch_count())


def new_keys(yrows, device):
    """uint64 keys[2][yrows] (held as int64 storage), reset to 'none'."""
    keys = torch.empty((2, yrows), dtype=torch.int64, device=device)
    keys_reset(keys)
    return keys


def keys_reset(keys):
    _check(clib.spv_keys_reset(keys.data_ptr(), keys.shape[1], _stream()), "spv_keys_reset")


def l1k2_fold(x, y, keys, xoffset=0):
    """K1: fold the L1 top-2 of query rows ``y`` against train rows ``x`` (both
    uint8 CUDA tensors, C-contiguous, dim % 16 == 0) into ``keys``; reported
    indices are ``xoffset + row``."""
    assert x.is_cuda and y.is_cuda and x.dtype == torch.uint8 and y.dtype == torch.uint8
    assert x.is_contiguous() and y.is_contiguous() and x.shape[1] == y.shape[1]
    assert keys.shape == (2, y.shape[0]) and keys.is_contiguous()
    _check(clib.spv_l1k2_device(x.data_ptr(), y.data_ptr(), x.shape[0], y.shape[0], x.shape[1],
                                int(xoffset), keys.data_ptr(), _stream()), "spv_l1k2_device")


def keys_merge(keys, gathered):
    """K5: keys <- top-2(keys U every key buffer in ``gathered`` [nsrc, 2, yrows])."""
    assert gathered.is_contiguous() and gathered.shape[1:] == keys.shape
    _check(clib.spv_keys_merge(keys.data_ptr(), gathered.data_ptr(), gathered.shape[0], keys.shape[1],
                               _stream()), "spv_keys_merge")


def keys_finalize(keys, kcols=2, dist_dtype=torch.int32):
    """keys -> (idx uint64 [yrows, kcols] (as int64 bits), dist [yrows, kcols])."""
    yrows = keys.shape[1]
    idx = torch.zeros((yrows, kcols), dtype=torch.int64, device=keys.device)
    dist = torch.zeros((yrows, kcols), dtype=dist_dtype, device=keys.device)
    fn = clib.spv_keys_finalize_i32 if dist_dtype == torch.int32 else clib.spv_keys_finalize_f32
    _check(fn(keys.data_ptr(), yrows, kcols, idx.data_ptr(), dist.data_ptr(), _stream()), "spv_keys_finalize")
    return idx, dist


clib.spv_ratio_ws_bytes.restype = ct.c_size_t
clib.spv_ratio_ws_bytes.argtypes = [ct.c_int]
clib.spv_ratio_matches_device.restype = ct.c_int
clib.spv_ratio_matches_device.argtypes = [_vp, ct.c_int, ct.c_double, ct.c_int, _vp, _vp, _vp, _vp, ct.c_size_t, _vp]


def ratio_matches(keys, min_ratio=1.75, dist_f32=False, want_mask=False):
    """K6: ratio test (second / float64(first) >= min_ratio, example/ex01:102-103) and
    order-preserving compaction on a key buffer.  Returns ``matches`` int64 [count, 2]
    (train idx, query row) -- one device->host read of the count -- and, with
    ``want_mask``, the uint8 pass mask [yrows]."""
    yrows = keys.shape[1]
    nbytes = int(clib.spv_ratio_ws_bytes(yrows))
    ws = torch.empty(nbytes + 8, dtype=torch.uint8, device=keys.device)
    count = torch.zeros(1, dtype=torch.int32, device=keys.device)
    matches = torch.empty((yrows, 2), dtype=torch.int64, device=keys.device)
    mask = torch.empty(yrows, dtype=torch.uint8, device=keys.device) if want_mask else None
    _check(clib.spv_ratio_matches_device(keys.data_ptr(), yrows, float(min_ratio), int(bool(dist_f32)),
                                         mask.data_ptr() if want_mask else None, matches.data_ptr(), count.data_ptr(),
                                         ws.data_ptr(), nbytes, _stream()), "spv_ratio_matches_device")
    out = matches[:int(count.item())]
    return (out, mask) if want_mask else out


clib.spv_normalize_ws_bytes.restype = ct.c_size_t
clib.spv_normalize_ws_bytes.argtypes = [ct.c_int]
clib.spv_normalize_new_dim.restype = ct.c_int
clib.spv_normalize_new_dim.argtypes = [ct.c_int]
clib.spv_normalize_device.restype = ct.c_int
clib.spv_normalize_device.argtypes = [_vp, ct.c_int, ct.c_int, _vp, _vp, _vp, ct.c_size_t, _vp]


def normalize_rows(x, want_f32=True, want_u8=True):
    """K7: normalize_to_ubyte_and_multiple_16_dim (feature.py:384-407) on a float32 CUDA
    matrix.  Returns (float32 [rows, ceil16(dim)] or None, uint8 (= value + 128) or None)."""
    assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.dim() == 2
    rows, dim = x.shape
    nd = int(clib.spv_normalize_new_dim(dim))
    nbytes = int(clib.spv_normalize_ws_bytes(dim))
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    of = torch.empty((rows, nd), dtype=torch.float32, device=x.device) if want_f32 else None
    ou = torch.empty((rows, nd), dtype=torch.uint8, device=x.device) if want_u8 else None
    _check(clib.spv_normalize_device(x.data_ptr(), rows, dim, of.data_ptr() if want_f32 else None,
                                     ou.data_ptr() if want_u8 else None, ws.data_ptr(), nbytes, _stream()),
           "spv_normalize_device")
    return of, ou


clib.spv_nn_bruteforce_ws_bytes.restype = ct.c_size_t
clib.spv_nn_bruteforce_ws_bytes.argtypes = [ct.c_int, ct.c_int, ct.c_int]
clib.spv_nn_bruteforce_ws_bytes_dim.restype = ct.c_size_t
clib.spv_nn_bruteforce_ws_bytes_dim.argtypes = [ct.c_int, ct.c_int, ct.c_int, ct.c_int]
clib.spv_nn_bruteforce_device.restype = ct.c_int
clib.spv_nn_bruteforce_device.argtypes = [_vp, _vp, ct.c_int, ct.c_int, ct.c_int, ct.c_int, ct.c_int, ct.c_double, _vp, _vp,
                                          _vp, ct.c_size_t, _vp]


def nn_bruteforce(x, y, p=.5, k=2):
    """K1f on CUDA tensors: float32 or int32 rows, sum |x - y|^p top-k (BruteForceNn.h:37-120).
    Returns (idx int64 [yrows, k] holding the uint64 bits, dist float32 / int32 [yrows, k])."""
    assert x.is_cuda and y.is_cuda and x.dtype == y.dtype and x.dtype in (torch.float32, torch.int32)
    assert x.is_contiguous() and y.is_contiguous() and x.dim() == 2 and y.dim() == 2 and x.shape[1] == y.shape[1]
    xrows, dim = x.shape
    yrows = y.shape[0]
    nbytes = int(clib.spv_nn_bruteforce_ws_bytes_dim(xrows, yrows, dim, k))
    if nbytes == 0:
        raise ValueError("unsupported k")
    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    idx = torch.empty((yrows, k), dtype=torch.int64, device=x.device)
    dist = torch.empty((yrows, k), dtype=x.dtype, device=x.device)
    _check(clib.spv_nn_bruteforce_device(x.data_ptr(), y.data_ptr(), int(x.dtype == torch.int32), xrows, yrows, dim, k,
                                         float(p), idx.data_ptr(), dist.data_ptr(), ws.data_ptr(), nbytes, _stream()),
           "spv_nn_bruteforce_device")
    return idx, dist


def nn_bruteforcel1k2_sharded(x_shard, y, xoffset, group=None):
    """Brute-force L1 k=2 with the train set sharded over the ranks of ``group``
    (SURVEY.md 8e): every rank folds its shard, the key buffers are
    all-gathered over NCCL and merged by K5 on every rank.  Returns
    ``(idx, dist)`` device tensors, identical on all ranks and bit-identical to
    the single-GPU result."""
    import torch.distributed as dist
    keys = new_keys(y.shape[0], y.device)
    l1k2_fold(x_shard, y, keys, xoffset)
    if group is not None or (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        world = dist.get_world_size(group)
        gathered = torch.empty((world,) + tuple(keys.shape), dtype=keys.dtype, device=keys.device)
        dist.all_gather_into_tensor(gathered, keys, group=group)
        keys_merge(keys, gathered)
    return keys_finalize(keys)


# ------------------------------------------------------------------ cascade of hashes
import numpy as _np

clib.spv_f32_to_u8_device.restype = ct.c_int
clib.spv_f32_to_u8_device.argtypes = [_vp, _vp, ct.c_size_t, _vp]
clib.spv_cascade_ws_bytes.restype = ct.c_size_t
clib.spv_cascade_ws_bytes.argtypes = [ct.c_int] * 6
clib.spv_cascade_device.restype = ct.c_int
clib.spv_cascade_device.argtypes = [_vp, _vp] + [ct.c_int] * 6 + [_vp, ct.c_int, ct.c_uint32, _vp, _vp, ct.c_size_t, _vp]
clib.spv_hash_codes_device.restype = ct.c_int
clib.spv_hash_codes_device.argtypes = [_vp] + [ct.c_int] * 5 + [_vp, ct.c_int, _vp, _vp, _vp, ct.c_size_t, _vp]
clib.spv_cascade_dictionary.restype = None
clib.spv_cascade_dictionary.argtypes = [ct.c_uint, ct.c_int, ct.c_int, ct.c_int, ct.c_int, _vp]
clib.spv_cascade_seed.restype = ct.c_uint
clib.spv_cascade_seed.argtypes = []


def cascade_seed():
    return int(clib.spv_cascade_seed())


def cascade_dictionary(dim, m, n, seed=None, quantize=True):
    """The hyperplanes nn_cascading_hash uses for these parameters: float32 [n, dim, m] (numpy)."""
    out = _np.zeros((n, dim, m), _np.float32)
    clib.spv_cascade_dictionary(cascade_seed() if seed is None else int(seed), dim, m, n, int(bool(quantize)),
                                out.ctypes.data)
    return out


def f32_to_u8(t):
    """K0: float32 rows holding integers in [-128, 127] -> uint8 (value + 128)."""
    assert t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
    out = torch.empty(t.shape, dtype=torch.uint8, device=t.device)
    _check(clib.spv_f32_to_u8_device(t.data_ptr(), out.data_ptr(), t.numel(), _stream()), "spv_f32_to_u8_device")
    return out


def _ws(xrows, yrows, dim, m, n, g, device):
    nbytes = int(clib.spv_cascade_ws_bytes(xrows, yrows, dim, m, n, g))
    if nbytes == 0:
        raise ValueError("cascade: unsupported parameters dim=%d m=%d n=%d g=%d" % (dim, m, n, g))
    return torch.empty(nbytes, dtype=torch.uint8, device=device), nbytes


def dict_terms_of(dictionary):
    """How K2 may multiply by this dictionary exactly: 0 when every coefficient is q / 32 with q an
    int8 in [-127, 127] (the library's own quantised dictionary: tcgen05 kind::i8 on the raw uint8
    rows), 1 when every coefficient is exact in bf16, else 3 (exact 3-term bf16 split)."""
    d = dictionary if torch.is_tensor(dictionary) else torch.from_numpy(_np.ascontiguousarray(dictionary))
    d = d.float()
    q = d * 32.0
    if bool(((q == q.round()) & (q.abs() <= 127)).all()):
        return 0
    return 1 if bool((d.bfloat16().float() == d).all()) else 3


def hash_codes(rows_u8, dictionary, g=0, dict_terms=None):
    """K2 alone: (codes int32 [n, rows], flips int64 [n, rows] or None)."""
    n, dim, m = dictionary.shape
    assert rows_u8.is_cuda and rows_u8.dtype == torch.uint8 and rows_u8.shape[1] == dim and rows_u8.is_contiguous()
    rows = rows_u8.shape[0]
    d = dictionary.to(rows_u8.device, torch.float32).contiguous()
    ws, nbytes = _ws(rows, rows, dim, m, n, g, rows_u8.device)
    codes = torch.zeros((n, rows), dtype=torch.int32, device=rows_u8.device)
    flips = torch.zeros((n, rows), dtype=torch.int64, device=rows_u8.device)
    terms = dict_terms_of(d) if dict_terms is None else dict_terms
    _check(clib.spv_hash_codes_device(rows_u8.data_ptr(), rows, dim, m, n, g, d.data_ptr(), terms, codes.data_ptr(),
                                      flips.data_ptr(), ws.data_ptr(), nbytes, _stream()), "spv_hash_codes_device")
    return codes, (flips if g > 0 else None)


def cascade_fold(x_u8, y_u8, keys, dictionary, g, xoffset=0, dict_terms=None):
    """K2-K4: fold the cascade-of-hashes top-2 of query rows y_u8 among train rows x_u8 into keys."""
    n, dim, m = dictionary.shape
    assert x_u8.dtype == torch.uint8 and y_u8.dtype == torch.uint8 and x_u8.is_contiguous() and y_u8.is_contiguous()
    assert x_u8.shape[1] == dim and y_u8.shape[1] == dim and keys.shape == (2, y_u8.shape[0])
    d = dictionary.to(y_u8.device, torch.float32).contiguous()
    ws, nbytes = _ws(x_u8.shape[0], y_u8.shape[0], dim, m, n, g, y_u8.device)
    terms = dict_terms_of(d) if dict_terms is None else dict_terms
    _check(clib.spv_cascade_device(x_u8.data_ptr(), y_u8.data_ptr(), x_u8.shape[0], y_u8.shape[0], dim, m, n, g,
                                   d.data_ptr(), terms, int(xoffset), keys.data_ptr(), ws.data_ptr(), nbytes, _stream()),
           "spv_cascade_device")
    return ws   # keep alive until the stream has consumed it


def nn_cascading_hash_sharded(x_shard_u8, y_u8, dictionary, g, xoffset, group=None):
    """The cascade matcher with the train set sharded over the ranks of ``group``: every rank
    hashes its shard with the SAME dictionary, probes with all queries, and the per-rank top-2
    keys are all-gathered over NCCL and merged.  Returns (idx, dist float32)."""
    import torch.distributed as dist
    keys = new_keys(y_u8.shape[0], y_u8.device)
    ws = cascade_fold(x_shard_u8, y_u8, keys, dictionary, g, xoffset)
    if group is not None or (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        world = dist.get_world_size(group)
        gathered = torch.empty((world,) + tuple(keys.shape), dtype=keys.dtype, device=keys.device)
        dist.all_gather_into_tensor(gathered, keys, group=group)
        keys_merge(keys, gathered)
    out = keys_finalize(keys, 2, torch.float32)
    del ws
    return out
