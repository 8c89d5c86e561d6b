"""spv_f32_to_u8_host: the host-side twin of K0 (f32_to_u8_kernel; CascadingHashNn.h:236-239), applied by
nn_cascading_hash's staging threads.  CPU: against the device rule restated in numpy -- cvt.rzi.s32.f32 (truncate,
saturate, NaN -> 0), + 128 wrapping, low byte -- on ordinary, fractional, huge, infinite and NaN inputs, at counts
that exercise the 32-wide AVX2 body and its scalar tail.  GPU: against K0 itself on the same values."""
import ctypes as ct

import numpy as np
import pytest


def _lib():
    from spectavi_b200._libspectavi import clib
    clib.spv_f32_to_u8_host.restype = None
    clib.spv_f32_to_u8_host.argtypes = [ct.c_void_p, ct.c_void_p, ct.c_size_t]
    return clib


def _values(n, seed):
    rng = np.random.RandomState(seed)
    v = rng.randint(-128, 128, n).astype("float32")                      # what the Python wrapper passes
    frac = rng.rand(n) < .25
    v[frac] = (rng.rand(int(frac.sum())) * 700 - 350).astype("float32")  # fractional, beyond [-128, 127]
    special = np.array([np.nan, np.inf, -np.inf, 2147483648.0, 2147483520.0, -2147483648.0, -2147483904.0, 3e9, -3e9, 1e30,
                        -1e30, -0.0, 0.99, -0.99, 127.5, -128.5, 255.0, 256.0, -129.0, 1e-40], "float32")
    pos = rng.choice(n, min(n, special.size), replace=False)
    v[pos] = special[:pos.size]
    return v


def _device_rule(v):
    with np.errstate(invalid="ignore"):
        t = np.trunc(v.astype("float64"))
    t = np.where(np.isnan(t), 0.0, t)
    i = np.clip(t, -2147483648.0, 2147483647.0).astype("int64")
    return ((i + 128) & 0xFF).astype("uint8")


@pytest.mark.parametrize("n", [1, 31, 32, 33, 64, 1000, 4096 + 17, 1 << 16])
def test_host_conversion_equals_device_rule(n):
    lib = _lib()
    v = _values(n, n)
    out = np.full(n + 8, 0xAB, "uint8")
    lib.spv_f32_to_u8_host(v.ctypes.data, out.ctypes.data, n)
    assert (out[:n] == _device_rule(v)).all()
    assert (out[n:] == 0xAB).all()          # nothing written past the end


@pytest.mark.gpu
def test_host_conversion_equals_k0():
    import torch
    from spectavi_b200 import device as dev
    lib = _lib()
    v = _values(1 << 18, 5)
    out = np.zeros(v.size, "uint8")
    lib.spv_f32_to_u8_host(v.ctypes.data, out.ctypes.data, v.size)
    k0 = dev.f32_to_u8(torch.from_numpy(v).cuda()).cpu().numpy().reshape(-1)
    assert (out == k0).all()
