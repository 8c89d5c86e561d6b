#!/usr/bin/env python
"""C5: cascade-of-hashes matcher vs brute force on one GPU.
  python tools/bench_cascade.py [rows=1000000] [dim=128] [reps=5] [cpu_rows=2000]
Queries: half are train rows + uniform noise in [-3, 3] (true neighbours exist), half are
independent uniform rows.  Reports
  * device-resident: cascade ms (CUDA events), queries/s, effective pairs/s, K2 alone on both sides with its
    achieved GB/s, the brute-force time on the same data, recall@1/@2 against brute force;
  * end to end: the call a user makes, feature.nn_cascading_hash on host float32 rows (the ABI's dtype,
    Spectavi.cpp:321-336: 4 bytes per element over PCIe, K0 on the device);
  * the CPU beside it: the reference's OWN CascadingHashNn (oracle/_ref, "kind": "reference"), construction
    (hashing + the serial bucket build, CascadingHashNn.h:187-206) separated from the per-query time by timing
    a 1-query call and a slice; the C port (oracle_cascade, all host cores) only where _ref is absent.
run_castle(): the example's default matcher on the example's data (example/ex01_essential_estimation.py:100-101,281):
the castle pair's SIFT rows through nn_cascading_hash with the auto-m rule; real rows are clustered, so buckets are skewed."""
import ctypes as ct
import json, os, sys, time
os.environ.setdefault("SPECTAVI_STAGE_TIMING", "1")     # events at the cascade's stage boundaries (read once by the library)
import numpy as np
import torch
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from spectavi_b200 import device as dev

HBM_PEAK_GBPS = 6552.0   # B200_PROFILING.md copy bandwidth; MEASURED_PEAKS.json when present


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return HBM_PEAK_GBPS, "B200_PROFILING.md fallback"


def timed(fn, reps):
    fn(); torch.cuda.synchronize()
    best = 1e30
    out = None
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); out = fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, out


def run(rows=1_000_000, dim=128, reps=5, cpu_rows=2000, e2e=True):
    """One C5 measurement on the current CUDA device; returns the result dict."""
    from spectavi_b200 import feature
    m = int(np.floor(np.log2(rows / 6.)))          # feature.py:364-367
    n, g = 2, 2
    gen = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randint(0, 256, (rows, dim), dtype=torch.uint8, device="cuda", generator=gen)
    y = torch.randint(0, 256, (rows, dim), dtype=torch.uint8, device="cuda", generator=gen)
    half = rows // 2
    src = torch.randint(0, rows, (half,), device="cuda", generator=gen)
    noise = torch.randint(-3, 4, (half, dim), device="cuda", generator=gen)
    y[:half] = (x[src].to(torch.int32) + noise).clamp(0, 255).to(torch.uint8)
    dic = torch.from_numpy(dev.cascade_dictionary(dim, m, n)).cuda()
    terms = dev.dict_terms_of(dic)   # 0: int8-exact dictionary -> tcgen05 kind::i8 projection
    peak, peak_src = hbm_peak()

    keys = dev.new_keys(rows, "cuda")
    def cascade():
        dev.keys_reset(keys)
        ws = dev.cascade_fold(x, y, keys, dic, g, 0, dict_terms=terms)
        return dev.keys_finalize(keys, 2, torch.float32), ws
    l0 = dev.launch_count()
    ms_c, ((ci, cd), _) = timed(cascade, reps)
    launches = (dev.launch_count() - l0) // (reps + 1)
    # per-stage times of one more call: CUDA events the library records at its stage boundaries
    from spectavi_b200._libspectavi import clib
    clib.spv_cascade_stage_ms.restype = ct.c_int
    clib.spv_cascade_stage_ms.argtypes = [ct.c_void_p, ct.c_int]
    st = (ct.c_double * 8)()
    best = None
    for _ in range(reps):
        cascade(); torch.cuda.synchronize()
        if clib.spv_cascade_stage_ms(st, 8) == 5:
            cur = [st[i] for i in range(5)]
            best = cur if best is None else [min(a, b) for a, b in zip(best, cur)]
    ms_img, ms_kx, ms_k3, ms_ky, ms_k4 = best if best else [float("nan")] * 5

    bkeys = dev.new_keys(rows, "cuda")
    def brute():
        dev.keys_reset(bkeys); dev.l1k2_fold(x, y, bkeys); return dev.keys_finalize(bkeys)
    ms_b, (bi, bd) = timed(brute, 1)
    r1 = float((ci[:, 0] == bi[:, 0]).float().mean()); r2 = float((ci == bi).float().mean())
    r1p = float((ci[:half, 0] == bi[:half, 0]).float().mean())
    bx, by = rows * (dim + 4 * n), rows * (dim + 12 * n)
    nb = n << m
    b3 = n * rows * 16 + nb * 4 * 6          # codes read twice, ids + cursor atomics; counters zeroed, scanned (r + w twice)
    b4 = rows * (n * (4 + 8) + (n << g) * 8 + 128) + float(rows) * (n << g) * (rows / float(1 << m)) * (4 + dim)   # codes / flips, ranges, query row; ids + candidate rows
    def stage(ms, nbytes, note=None):
        d = {"ms": ms, "algorithmic_bytes": int(nbytes), "GBps": nbytes / ms / 1e6, "frac_of_hbm_peak": nbytes / ms / 1e6 / peak}
        if note:
            d["note"] = note
        return d
    out = {"workload": "C5 cascade %d x %d x %d, m=%d n=%d g=%d" % (rows, rows, dim, m, n, g),
           "cascade_ms": ms_c, "queries_per_s": rows / ms_c * 1e3, "effective_Gpairs_per_s": rows * rows / ms_c / 1e6,
           "kernels_per_call": int(launches),
           "stages": {"hbm_peak_GBps": peak, "hbm_peak_source": peak_src,
                      "timing": "CUDA events recorded by the library at its stage boundaries (SPECTAVI_STAGE_TIMING), best of %d" % reps,
                      "dictionary_image": {"ms": ms_img},
                      "K2_train(projection + codes)": stage(ms_kx, bx),
                      "K3(bucket count + scan + scatter)": stage(ms_k3, b3),
                      "K2_query(+ least-confident bits)": stage(ms_ky, by),
                      "K4(probe + gather + re-rank)": stage(ms_k4, b4, "bytes = every candidate row gathered once (no cache reuse assumed); the 126 MB L2 serves about half of them")},
           "bruteforce_ms": ms_b, "bruteforce_Gpairs_per_s": rows * rows / ms_b / 1e6, "speedup_vs_bruteforce": ms_b / ms_c,
           "recall@1_all": r1, "recall@1_planted_half": r1p, "index_agreement_top2": r2}
    if e2e:
        # the call a user makes: host float32 rows (value - 128), 4 B per element over PCIe
        xf = (x.cpu().numpy().astype("float32") - 128)
        yf = (y.cpu().numpy().astype("float32") - 128)
        feature.nn_cascading_hash(xf, yf, k=2, m=m, n=n, g=g)
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter(); ei, ed = feature.nn_cascading_hash(xf, yf, k=2, m=m, n=n, g=g); best = min(best, time.perf_counter() - t0)
        out["e2e"] = {"api": "feature.nn_cascading_hash (C ABI nn_cascading_hash) on pageable float32 numpy rows",
                      "ms": best * 1e3, "queries_per_s": rows / best, "host_float_bytes": int(2 * rows * dim * 4),
                      "h2d_bytes": int(2 * rows * dim), "d2h_bytes": int(rows * 2 * 12),
                      "host_float_GBps": 2 * rows * dim * 4 / best / 1e9,
                      "note": "the float32 rows are converted to uint8 by the staging threads on their way into pinned memory, so 1 byte per element crosses PCIe",
                      "equals_device_resident_result": bool((ei == ci.cpu().numpy().view(np.uint64)).all() and (ed == cd.cpu().numpy()).all())}
    if cpu_rows > 0:
        from oracle import pyoracle as po
        xf = x.cpu().numpy().astype("float32") - 128
        yf = y[:cpu_rows].cpu().numpy().astype("float32") - 128
        cores = os.cpu_count()
        t0 = time.perf_counter(); oi, od = po.cascade(xf, yf, dic.cpu().numpy(), g, cores); dt = time.perf_counter() - t0
        out["parity_vs_oracle_rows"] = cpu_rows
        out["parity_bit_exact"] = bool((ci[:cpu_rows].cpu().numpy().view(np.uint64) == oi).all() and (cd[:cpu_rows].cpu().numpy() == od).all())
        if po.have_ref():
            # the reference's own CascadingHashNn: construction hashes ALL train rows and builds the tables serially
            nq = min(rows, 100_000)            # enough queries for the per-query share to stand out of the construction time
            yq = y[:nq].cpu().numpy().astype("float32") - 128
            t0 = time.perf_counter(); po.ref_cascade(xf, yq[:1], m, n, g, seed=7); t_build = time.perf_counter() - t0
            t0 = time.perf_counter(); ri, rd = po.ref_cascade(xf, yq, m, n, g, seed=7)[:2]; t_all = time.perf_counter() - t0
            per_q = max(t_all - t_build, 0.0) / nq
            out["cpu_baseline"] = {"kind": "reference", "cores": "8 (CascadingHashNn.h:230 default nthread) for the re-rank; hashing / tables serial",
                                   "construction_s(hash %d train rows + build tables)" % rows: t_build,
                                   "per_query_us": per_q * 1e6, "queries": nq, "whole_call_s(%d queries)" % nq: t_all,
                                   "whole_call_extrapolated_s(%d queries)" % rows: t_build + per_q * rows,
                                   "queries_per_s(whole call)": rows / (t_build + per_q * rows)}
        else:
            out["cpu_baseline"] = {"kind": "port", "cores": cores, "seconds": dt, "queries": cpu_rows, "queries_per_s": cpu_rows / dt,
                                   "note": "includes hashing ALL train rows + bucket build"}
    return out


def run_castle():
    """The example's default matcher on the example's data: castle SIFT rows, nn_cascading_hash with auto-m."""
    from spectavi_b200 import feature
    from oracle import pyoracle as po
    gold = np.load(os.path.join(ROOT, "tests", "golden", "castle_c2.npz"))
    s1 = np.hstack([gold["k1"], gold["d1"].astype("float32")])
    s2 = np.hstack([gold["k2"], gold["d2"].astype("float32")])
    xf = feature.normalize_to_ubyte_and_multiple_16_dim(s1)     # ex01:96-97: float32 rows in [-128, 127]
    yf = feature.normalize_to_ubyte_and_multiple_16_dim(s2)
    m = int(np.floor(np.log2(max(xf.shape[0], yf.shape[0]) / 6.)))
    feature.nn_cascading_hash(xf, yf)
    best = 1e30
    for _ in range(5):
        t0 = time.perf_counter(); ci, cd = feature.nn_cascading_hash(xf, yf); best = min(best, time.perf_counter() - t0)
    bi = gold["idx"].astype(np.uint64)      # the reference's own brute-force output on these rows
    mx, my = feature.match_cascading_hash(xf, yf, 1.75)
    bmask = feature.ratio_test(gold["dist"], 1.75)
    out = {"workload": "castle pair %d x %d x %d, nn_cascading_hash auto m=%d n=2 g=2" % (xf.shape[0], yf.shape[0], xf.shape[1], m),
           "e2e_ms": best * 1e3, "recall@1": float((ci[:, 0] == bi[:, 0]).mean()), "recall@2": float((ci == bi).mean()),
           "matches_after_ratio_test": int(mx.shape[0]), "bruteforce_matches_after_ratio_test": int(bmask.sum()),
           "matches_also_found_by_bruteforce": int(np.isin(my, np.nonzero(bmask)[0]).sum())}
    # device-resident time of the same call
    xu = torch.from_numpy((xf + 128).astype("uint8")).cuda(); yu = torch.from_numpy((yf + 128).astype("uint8")).cuda()
    dic = torch.from_numpy(dev.cascade_dictionary(xf.shape[1], m, 2)).cuda()
    keys = dev.new_keys(yf.shape[0], "cuda")
    def cascade():
        dev.keys_reset(keys)
        return dev.cascade_fold(xu, yu, keys, dic, 2, 0, dict_terms=0)
    out["device_ms"], _ = timed(cascade, 5)
    if po.have_ref():
        t0 = time.perf_counter(); ri, rd = po.ref_cascade(xf, yf, m, 2, 2, seed=7)[:2]; dt = time.perf_counter() - t0
        out["cpu_reference"] = {"kind": "reference", "ms": dt * 1e3, "recall@1": float((ri[:, 0] == bi[:, 0]).mean()),
                                "recall@2": float((ri == bi).mean()), "note": "its dictionary is N(0,1) fp32 from a seeded random_device stand-in"}
    return out


def run_sharded(device, world, rank, rows=1_000_000, dim=128, reps=5):
    """C5 at N GPUs under torchrun (call on every rank): train rows sharded, the SAME dictionary everywhere, all-gather +
    K5 merge; rank 0 also runs the whole train set alone and checks bit-identity.  Returns the dict on rank 0, else None."""
    import torch.distributed as dist
    from spectavi_b200.sharding import shard_range
    m = int(np.floor(np.log2(rows / 6.)))
    n, g = 2, 2
    gen = torch.Generator(device=device).manual_seed(3)     # the same bytes on every rank
    x = torch.randint(0, 256, (rows, dim), dtype=torch.uint8, device=device, generator=gen)
    y = torch.randint(0, 256, (rows, dim), dtype=torch.uint8, device=device, generator=gen)
    half = rows // 2
    src = torch.randint(0, rows, (half,), device=device, generator=gen)
    noise = torch.randint(-3, 4, (half, dim), device=device, generator=gen)
    y[:half] = (x[src].to(torch.int32) + noise).clamp(0, 255).to(torch.uint8)
    dic = torch.from_numpy(dev.cascade_dictionary(dim, m, n)).to(device)
    terms = dev.dict_terms_of(dic)
    x0, xn = shard_range(rows, world, rank)
    xs = x[x0:x0 + xn].contiguous()
    keys = dev.new_keys(rows, device)
    gathered = torch.empty((world, 2, rows), dtype=torch.int64, device=device)

    def step():
        dev.keys_reset(keys)
        ws = dev.cascade_fold(xs, y, keys, dic, g, x0, dict_terms=terms)
        dist.all_gather_into_tensor(gathered, keys)
        dev.keys_merge(keys, gathered)
        return dev.keys_finalize(keys, 2, torch.float32), ws

    for _ in range(3):
        step()
    dist.barrier(); torch.cuda.synchronize()
    ev = [(torch.cuda.Event(True), torch.cuda.Event(True)) for _ in range(reps)]
    for a, b in ev:
        a.record(); (ci, cd), _ = step(); b.record()
    dist.barrier(); torch.cuda.synchronize()
    ms = torch.tensor([min(a.elapsed_time(b) for a, b in ev)], dtype=torch.float64, device=device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    k1 = dev.new_keys(rows, device)
    ws = dev.cascade_fold(x, y, k1, dic, g, 0, dict_terms=terms)
    si, sd = dev.keys_finalize(k1, 2, torch.float32)
    torch.cuda.synchronize()
    return {"workload": "C5 cascade %d x %d x %d, m=%d n=%d g=%d, train sharded %d-way" % (rows, rows, dim, m, n, g, world),
            "cascade_ms(max over ranks, best of %d)" % reps: float(ms[0]), "queries_per_s": rows / float(ms[0]) * 1e3,
            "effective_Gpairs_per_s": rows * rows / float(ms[0]) / 1e6,
            "sharded_equals_single_gpu": bool((ci == si).all() and (cd == sd).all())}


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "castle":
        print(json.dumps(run_castle()))
    else:
        a = [int(v) for v in sys.argv[1:5]]
        d = [1_000_000, 128, 5, 2000]
        print(json.dumps(run(*(a + d[len(a):]))))
