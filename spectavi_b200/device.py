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
