"""Cascade-of-hashes parity on the GPU (K0, K2-K4).

* hash codes and probe-code sets: bit-exact vs the C oracle for the library's
  quantised dictionary; equal to the codes the reference's own CascadingHashNn.h
  produced (golden files) for its unquantised fp32 dictionaries;
* whole matcher: idx AND dist bit-exact vs the oracle fed the same dictionary
  (both use the lexicographic (dist, idx) rule); vs the reference golden
  results: distances exact, indices equal except where the reference's
  unordered_set scan order picked another member of an exact tie;
* through the drop-in nn_cascading_hash (dictionary from spv_cascade_dictionary);
* the reference's own acceptance test (test/test_feature.py:123-151)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
NONE_IDX = np.uint64(0xFFFFFFFFFFFFFFFF)


@pytest.fixture(scope="module")
def dev():
    import torch  # noqa: F401
    from spectavi_b200 import device
    return device


@pytest.fixture(scope="module")
def feature():
    from spectavi_b200 import feature
    return feature


def _quantised_rows(rng, rows, dim, alphabet=256):
    return (rng.randint(0, alphabet, (rows, dim)) - 128).astype("float32")


def _gpu_cascade(dev, x, y, dictionary, g):
    import torch
    xu = torch.from_numpy((x + 128).astype("uint8")).cuda()
    yu = torch.from_numpy((y + 128).astype("uint8")).cuda()
    keys = dev.new_keys(y.shape[0], "cuda")
    ws = dev.cascade_fold(xu, yu, keys, torch.from_numpy(dictionary), g)
    idx, dist = dev.keys_finalize(keys, 2, torch.float32)
    torch.cuda.synchronize()
    del ws
    return idx.cpu().numpy().view(np.uint64), dist.cpu().numpy()


def _probe_sets(codes, flips, nb):
    """all 2^nb probe codes of every (table, row) from base code + flip positions"""
    n, rows = codes.shape
    out = np.zeros((n, rows, 1 << nb), np.int64)
    for ii in range(1 << nb):
        c = codes.astype(np.int64).copy()
        for q in range(nb):
            bit = (flips >> (5 * q)) & 31
            c = (c & ~(1 << bit)) | (((ii >> q) & 1) << bit)
        out[:, :, ii] = c
    return out


@pytest.mark.parametrize("shape", [(1000, 128, 17, 2, 2), (300, 144, 8, 16, 5), (257, 16, 5, 3, 7), (129, 512, 31, 2, 3),
                                   (500, 256, 24, 5, 0), (64, 128, 9, 9, 4)])
@pytest.mark.parametrize("terms", [0, 1, 3])    # 0: tcgen05 kind::i8 on the raw uint8 rows; 1 / 3: bf16 terms
def test_hash_codes_bit_exact_quantised_dictionary(po, dev, shape, terms):
    import torch
    rows, dim, m, n, g = shape
    rng = np.random.RandomState(rows + m)
    x = _quantised_rows(rng, rows, dim)
    d = dev.cascade_dictionary(dim, m, n, seed=rows)
    assert dev.dict_terms_of(d) == 0
    codes, flips = dev.hash_codes(torch.from_numpy((x + 128).astype("uint8")).cuda(), torch.from_numpy(d), g,
                                  dict_terms=terms)
    codes = codes.cpu().numpy()
    want = po.hashcodes(x, d)                      # [rows][n]
    assert (codes.T == want).all()
    if g > 0:
        nb = min(g, m)
        sets = _probe_sets(codes, flips.cpu().numpy(), nb)
        for t in range(n):
            for r in range(0, rows, max(1, rows // 40)):
                want_set = sorted(set(int(c) for c in po.probe_codes(x[r], d[t], nb)))
                assert sorted(set(int(c) for c in sets[t, r])) == want_set, (t, r)


@pytest.mark.parametrize("name,g_", [("cascade_200x200x144_m8n16g5.npz", 5), ("cascade_400x300x128_m9n2g2.npz", 2)])
def test_cascade_golden_from_reference(dev, name, g_):
    import torch
    from conftest import golden
    g = golden(name)
    x, y, d = g["x"].astype("float32"), g["y"].astype("float32"), g["dictionary"]
    assert dev.dict_terms_of(d) == 3               # the reference's dictionary is plain N(0,1) fp32
    codes, _ = dev.hash_codes(torch.from_numpy((x + 128).astype("uint8")).cuda(), torch.from_numpy(d), 0)
    assert (codes.cpu().numpy().T == g["codes"]).all()
    idx, dist = _gpu_cascade(dev, x, y, d, g_)
    assert (dist == g["dist"]).all()
    bx, by = (x + 128).astype("int64"), (y + 128).astype("int64")
    diff = np.argwhere(idx != g["idx"])
    for r, c in diff:                               # only exact ties may be resolved differently
        assert np.abs(bx[int(g["idx"][r, c])] - by[r]).sum() == int(dist[r, c])
    assert len(diff) <= 0.02 * idx.size


@pytest.mark.parametrize("shape", [
    (200, 200, 144, 8, 16, 5),       # the reference test's parameters
    (400, 300, 128, 9, 2, 2),
    (5000, 1000, 128, 9, 2, 2),      # auto-m regime: ~10 rows per bucket
    (3000, 500, 128, 4, 2, 2),       # ~190 rows per bucket: several list drains
    (2000, 64, 128, 2, 1, 0),        # 500 rows per bucket, g = 0: the direct-from-global path
    (300, 100, 16, 26, 2, 3),        # m > 24: bucket key is the low 24 bits, full code verified
    (150, 40, 512, 12, 3, 12),       # g == m: every code is probed
    (1, 7, 128, 6, 2, 2), (2, 7, 128, 6, 2, 2),   # fewer than two candidates
    (400, 90, 1024, 8, 2, 2), (300, 50, 528, 10, 3, 3),   # dim > 512 (reference: any dim % 16 == 0): bf16 projection, run-time chunk loop in K4
])
def test_cascade_bit_exact_vs_oracle(po, dev, shape):
    xr, yr, dim, m, n, g = shape
    rng = np.random.RandomState(xr + yr + m)
    x, y = _quantised_rows(rng, xr, dim), _quantised_rows(rng, yr, dim)
    if xr >= 2000:                                   # plant true neighbours so that distances matter
        y[: yr // 2] = np.clip(x[rng.randint(0, xr, yr // 2)] + rng.randint(-3, 4, (yr // 2, dim)), -128, 127)
    d = dev.cascade_dictionary(dim, m, n, seed=7)
    gi, gd = _gpu_cascade(dev, x, y, d, g)
    oi, od = po.cascade(x, y, d, g, 4)
    assert (gd == od).all()
    assert (gi == oi).all()


def test_cascade_ties_and_duplicates(po, dev):
    # low-entropy rows: many equal distances, every row in many probed buckets
    rng = np.random.RandomState(9)
    x, y = _quantised_rows(rng, 700, 32, 3), _quantised_rows(rng, 90, 32, 3)
    d = dev.cascade_dictionary(32, 5, 6, seed=3)
    gi, gd = _gpu_cascade(dev, x, y, d, 3)
    oi, od = po.cascade(x, y, d, 3, 2)
    assert (gd == od).all() and (gi == oi).all()


def test_nn_cascading_hash_drop_in(po, dev, feature):
    rng = np.random.RandomState(21)
    x, y = _quantised_rows(rng, 3000, 128), _quantised_rows(rng, 400, 128)
    y[:200] = np.clip(x[rng.randint(0, 3000, 200)] + rng.randint(-2, 3, (200, 128)), -128, 127)
    gi, gd = feature.nn_cascading_hash(x, y, k=2, m=8, n=4, g=2)
    assert gi.dtype == np.uint64 and gd.dtype == np.float32 and gi.shape == (400, 2)
    d = dev.cascade_dictionary(128, 8, 4)           # what the entry point drew (fixed seed)
    oi, od = po.cascade(x, y, d, 2, 4)
    assert (gd == od).all() and (gi == oi).all()
    # auto m (feature.py:364-367): floor(log2(3000 / 6)) = 8
    gi2, gd2 = feature.nn_cascading_hash(x, y)
    d2 = dev.cascade_dictionary(128, 8, 2)
    oi2, od2 = po.cascade(x, y, d2, 2, 4)
    assert (gd2 == od2).all() and (gi2 == oi2).all()
    # k > 2: columns 0, 1 as for k = 2, the rest stay zero (Spectavi.cpp:329-334)
    gi3, gd3 = feature.nn_cascading_hash(x, y, k=4, m=8, n=4, g=2)
    assert (gi3[:, :2] == gi).all() and (gd3[:, :2] == gd).all() and (gi3[:, 2:] == 0).all() and (gd3[:, 2:] == 0).all()


def test_nn_cascading_hash_small_falls_back_to_brute_force(po, feature):
    # feature.py:364-371: fewer than 96 rows -> m < 4 -> exact L1 k=2 with int32 distances
    rng = np.random.RandomState(2)
    x, y = _quantised_rows(rng, 50, 32), _quantised_rows(rng, 20, 32)
    gi, gd = feature.nn_cascading_hash(x, y)
    oi, od = po.l1k2((x + 128).astype("uint8"), (y + 128).astype("uint8"), 1)
    assert gd.dtype == np.int32 and (gi == oi).all() and (gd == od).all()


def test_cascade_sentinels(dev):
    # queries whose probed buckets are empty: idx = size_t(-1), dist = (float)INT_MAX
    x = np.full((4, 16), -128, "float32")
    y = np.full((3, 16), 127, "float32")
    d = dev.cascade_dictionary(16, 12, 1, seed=1)
    gi, gd = _gpu_cascade(dev, x, y, d, 0)
    assert (gi == NONE_IDX).all() and (gd == np.float32(2147483648.0)).all()


def test_reference_acceptance_test(po, feature):
    # test/test_feature.py:123-151: 200 x 200 randn -> normalised 144-D, m=8 n=16 g=5;
    # at most 160 of the 400 index entries may differ from the exact L1 top-2
    np.random.seed(0xdeadbeef)
    x = feature.normalize_to_ubyte_and_multiple_16_dim(np.random.randn(200, 132))
    y = feature.normalize_to_ubyte_and_multiple_16_dim(np.random.randn(200, 132))
    gi, _ = feature.nn_cascading_hash(x, y, m=8, n=16, g=5)
    ti, _ = po.numpy_l1k2((x + 128).astype("uint8"), (y + 128).astype("uint8"))
    assert (gi != ti).sum() <= 2 * round(.4 * 200)


@pytest.mark.parametrize("rows,dim,m,n,g", [(100_003, 128, 14, 2, 2), (40_001, 64, 10, 12, 3), (33_000, 256, 20, 3, 2),
                                            (5_000, 512, 12, 2, 2), (20_000, 32, 11, 30, 1)])
def test_hash_codes_i8_pipeline_many_tiles(po, dev, rows, dim, m, n, g):
    """The persistent TMA -> tcgen05 kind::i8 -> TMEM -> epilogue pipeline over hundreds of row tiles per
    CTA, ragged last tile, several table groups: codes == the bf16 kernel's == the oracle's (sampled)."""
    import torch
    rng = np.random.RandomState(rows + dim)
    x = _quantised_rows(rng, rows, dim)
    d = dev.cascade_dictionary(dim, m, n, seed=dim)
    xu = torch.from_numpy((x + 128).astype("uint8")).cuda()
    c0, f0 = dev.hash_codes(xu, torch.from_numpy(d), g, dict_terms=0)
    c1, f1 = dev.hash_codes(xu, torch.from_numpy(d), g, dict_terms=1)
    assert bool((c0 == c1).all()) and bool((f0 == f1).all())
    sel = rng.choice(rows, 3000, replace=False)
    want = po.hashcodes(x[sel], d)
    assert (c0.cpu().numpy().T[sel] == want).all()


@pytest.mark.parametrize("terms", [0, 1, 3])
def test_cascade_exact_workspace_size(po, dev, terms):
    """spv_cascade_ws_bytes must cover the layout of every dict_terms (the image size depends on the number of
    tables per pass): dim 256, m 25, n 5 is a combination where the 1-term layout is the larger one."""
    import torch
    rng = np.random.RandomState(4)
    xr, yr, dim, m, n, g = 600, 80, 256, 25, 5, 2
    x, y = _quantised_rows(rng, xr, dim), _quantised_rows(rng, yr, dim)
    d = dev.cascade_dictionary(dim, m, n, seed=11)
    xu = torch.from_numpy((x + 128).astype("uint8")).cuda()
    yu = torch.from_numpy((y + 128).astype("uint8")).cuda()
    keys = dev.new_keys(yr, xu.device)
    ws = dev.cascade_fold(xu, yu, keys, torch.from_numpy(d), g, 0, dict_terms=terms)   # allocates exactly ws_bytes
    gi, gd = dev.keys_finalize(keys, 2, torch.float32)
    del ws
    oi, od = po.cascade(x, y, d, g, 4)
    assert (gd.cpu().numpy() == od).all() and (gi.cpu().numpy().view(np.uint64) == oi).all()


def test_cascade_fold_shards_concurrently(po, dev):
    """Two train shards folded into ONE key buffer from two streams at once: K4's final fold is the same
    two-atomicMin protocol as K1's, so no update is lost and the result equals the unsharded one."""
    import torch
    rng = np.random.RandomState(12)
    xr, yr, dim, m, n, g = 6000, 3000, 128, 8, 2, 2
    x, y = _quantised_rows(rng, xr, dim), _quantised_rows(rng, yr, dim)
    y[:1500] = np.clip(x[rng.randint(0, xr, 1500)] + rng.randint(-3, 4, (1500, dim)), -128, 127)
    d = torch.from_numpy(dev.cascade_dictionary(dim, m, n, seed=5))
    xu = torch.from_numpy((x + 128).astype("uint8")).cuda()
    yu = torch.from_numpy((y + 128).astype("uint8")).cuda()
    keys = dev.new_keys(yr, xu.device)
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    with torch.cuda.stream(s1):
        w1 = dev.cascade_fold(xu[:3000].contiguous(), yu, keys, d, g, 0)
    with torch.cuda.stream(s2):
        w2 = dev.cascade_fold(xu[3000:].contiguous(), yu, keys, d, g, 3000)
    torch.cuda.synchronize()
    gi, gd = dev.keys_finalize(keys, 2, torch.float32)
    # per-shard candidate sets: the union of the two shard searches (buckets are per shard)
    a = po.cascade(x[:3000], y, d.numpy(), g, 4)
    b = po.cascade(x[3000:], y, d.numpy(), g, 4)
    NONE = np.uint64(0xFFFFFFFFFFFFFFFF)
    bi = np.where(b[0] == NONE, NONE, b[0] + np.uint64(3000))
    ci, cd = np.concatenate([a[0], bi], 1), np.concatenate([a[1], b[1]], 1)
    order = np.lexsort((ci, cd), axis=1)[:, :2] if False else np.argsort(cd + 0 * ci.astype("float64"), axis=1, kind="stable")
    # lexicographic (dist, idx) top-2 of the four candidates
    want_i, want_d = np.empty((yr, 2), np.uint64), np.empty((yr, 2), np.float32)
    for r in range(yr):
        c = sorted(zip(cd[r].tolist(), ci[r].tolist()))
        want_d[r] = [c[0][0], c[1][0]]
        want_i[r] = [c[0][1], c[1][1]]
    assert (gd.cpu().numpy() == want_d).all() and (gi.cpu().numpy().view(np.uint64) == want_i).all()
    del w1, w2
