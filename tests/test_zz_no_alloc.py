"""spv_ratio_matches_device_static: K6 (ratio test + match compaction) with its scratch in a file-scope device array
inside the library, for callers that cannot call cudaMalloc.  It has no counterpart in the reference; the check is
against the existing entry point spv_ratio_matches_device on the same key buffers -- pass mask, match count, order and
every (train idx, query row) pair, bit for bit (integer / index work) -- on two calls in a row with different inputs,
the second one smaller, so that whatever the first call left in the shared array is still there.

Collected last on purpose.  Needs nothing from the other tests: it binds the symbol
itself and builds its own inputs, so it passes alone in a fresh process and after the rest of the suite."""
import ctypes as ct

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

_vp = ct.c_void_p


@pytest.fixture(scope="module")
def lib():
    from spectavi_b200._libspectavi import clib
    clib.spv_ratio_static_max_rows.restype = ct.c_int
    clib.spv_ratio_static_max_rows.argtypes = []
    clib.spv_ratio_matches_device_static.restype = ct.c_int
    clib.spv_ratio_matches_device_static.argtypes = [_vp, ct.c_int, ct.c_double, ct.c_int, _vp, _vp, _vp, _vp]
    clib.spv_last_error.restype = ct.c_char_p
    return clib


def _keys(torch, dev, xrows, yrows, dim, alphabet, seed):
    """a key buffer as K1 leaves it (small alphabets give d1 == d2 and d1 == 0: the inf / nan branches)"""
    rng = np.random.RandomState(seed)
    x = torch.from_numpy(rng.randint(0, alphabet, (xrows, dim)).astype("uint8")).cuda()
    y = torch.from_numpy(rng.randint(0, alphabet, (yrows, dim)).astype("uint8")).cuda()
    keys = dev.new_keys(yrows, x.device)
    dev.l1k2_fold(x, y, keys)
    return keys


def _static(torch, lib, keys, min_ratio, dist_f32):
    yrows = keys.shape[1]
    # outputs are the caller's; allocated BEFORE the call under test
    count = torch.full((1,), -1, dtype=torch.int32, device=keys.device)
    matches = torch.full((yrows, 2), -7, dtype=torch.int64, device=keys.device)
    mask = torch.full((yrows,), 9, dtype=torch.uint8, device=keys.device)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    rc = lib.spv_ratio_matches_device_static(keys.data_ptr(), yrows, float(min_ratio), int(dist_f32), mask.data_ptr(),
                                             matches.data_ptr(), count.data_ptr(),
                                             _vp(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.spv_last_error().decode()
    assert torch.cuda.memory_allocated() == before
    n = int(count.item())
    return matches[:n].cpu().numpy(), mask.cpu().numpy(), n


@pytest.mark.parametrize("min_ratio,dist_f32", [(1.2, False), (1.05, True)])
def test_static_scratch_equals_workspace_entry_point_twice_in_a_row(lib, min_ratio, dist_f32):
    import torch
    from spectavi_b200 import device as dev
    # first call: 70 001 rows = 274 blocks of counts; second call: 1 031 rows = 5 blocks, the array still holds the
    # first call's scanned offsets behind them; third: the first input again
    calls = [_keys(torch, dev, 900, 70001, 32, 6, 1), _keys(torch, dev, 50, 1031, 16, 3, 2)]
    calls.append(calls[0])
    for keys in calls:
        want, want_mask = dev.ratio_matches(keys, min_ratio, dist_f32, want_mask=True)
        got, got_mask, n = _static(torch, lib, keys, min_ratio, dist_f32)
        want = want.cpu().numpy()
        assert n == want.shape[0] and 0 < n <= keys.shape[1]
        assert (got_mask == want_mask.cpu().numpy()).all()
        assert (got == want).all()


def test_static_scratch_empty_and_limit(lib):
    import torch
    from spectavi_b200 import device as dev
    limit = lib.spv_ratio_static_max_rows()
    assert limit == 1 << 22
    # beyond the limit: refused with a message, nothing launched (the pointers are never touched)
    assert lib.spv_ratio_matches_device_static(None, limit + 1, 1.75, 0, None, None, None, None) == -1
    msg = lib.spv_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg
    assert lib.spv_ratio_matches_device_static(None, -1, 1.75, 0, None, None, None, None) == -1
    # zero rows: count = 0
    count = torch.full((1,), 5, dtype=torch.int32, device="cuda")
    keys = dev.new_keys(1, count.device)
    assert lib.spv_ratio_matches_device_static(keys.data_ptr(), 0, 1.75, 0, None, None, count.data_ptr(), None) == 0
    assert int(count.item()) == 0
    # at the limit itself the call is accepted: all-"no neighbour" keys, INT_MAX / INT_MAX = 1.0 >= 1.0 passes everywhere
    big = dev.new_keys(limit, count.device)
    matches = torch.empty((limit, 2), dtype=torch.int64, device="cuda")
    assert lib.spv_ratio_matches_device_static(big.data_ptr(), limit, 1.0, 0, None, matches.data_ptr(), count.data_ptr(),
                                               _vp(torch.cuda.current_stream().cuda_stream)) == 0
    assert int(count.item()) == limit
    assert bool((matches[:, 1] == torch.arange(limit, device="cuda")).all()) and bool((matches[:, 0] == -1).all())
