#!/usr/bin/env python
"""bench.py -- G L1 descriptor-pair comparisons / s (k=2, 128-D) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                    [--workload c1|c3|c4|XxYxD]

One step = one brute-force L1 k=2 match of the whole workload (reference
src/BruteForceNnL1K2.h:84-145 behind nn_bruteforcel1k2, Spectavi.cpp:284-298):
    N == 1 : workload c3 = BASELINE.json configs[2], 200k query x 200k train x 128 u8
    N  > 1 : workload c4 = configs[3], 1M x 1M x 128, train rows sharded over the
             N ranks (one process per GPU), query rows replicated, per-rank top-2
             all-gathered over NCCL and merged (K5) every step  -> "strong" scaling
`value`   : pairs covered / s with inputs resident in HBM (CUDA events, max over ranks)
`e2e`     : the same through the drop-in call a user makes, feature.nn_bruteforcel1k2 (C ABI
            nn_bruteforcel1k2) on ordinary PAGEABLE numpy arrays, H2D of the descriptors and D2H
            of (idx, dist) inside the timed region.  At N > 1 that is the library's own multi-GPU
            layout (SPECTAVI_GPUS=N: one host thread per GPU, NCCL all-gather of the keys), run
            by rank 0 while the other ranks wait on a CPU (gloo) barrier; the torchrun layout
            from pinned buffers is reported beside it in `extras`.
`roofline`: K1's packed-SAD rate against the VABSDIFF4 issue peak measured in the
            same run by spv_pipe_peak (the path is ALU-bound), and
            against the nominal 64 per clk per SM
`cpu_baseline` / --impl reference: the reference's own BruteForceNnL1K2 (oracle/_ref,
            compiled from /root/reference through oracle/eigen_shim; else the C port)
            on all host cores over a bounded slice of query rows against the full
            train set (rows are independent, BruteForceNnL1K2.h:92-93).
Parity is checked inside the run at every N: the timed GPU result equals the CPU sample bit for
bit, and `result_hash` (blake2b of the full idx + dist arrays) is the same string at N = 2, 4, 8
and in the N = 1 line's `extras.c4_n1`.
"""
import argparse
import datetime
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

METRIC = "G L1 descriptor-pair comparisons/sec (k=2, 128-D) at 1/2/4/8 B200 vs CPU SSE"
UNIT = "Gpairs/s"
WORKLOADS = {"c1": (4096, 4096, 128), "c3": (200_000, 200_000, 128), "c4": (1_000_000, 1_000_000, 128)}
L2_FLUSH_BYTES = 256 << 20


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, help="c1|c3|c4 or XROWSxYROWSxDIM")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU baseline sample budget")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-peak", action="store_true", help="skip the pipe-peak microbenchmark")
    ap.add_argument("--no-extras", action="store_true", help="skip the side measurements (C1, C2, C4 on one GPU, C5, K1f, K8)")
    return ap.parse_args()


def workload_of(args, world):
    name = args.workload or ("c3" if world == 1 else "c4")
    if name in WORKLOADS:
        return name, WORKLOADS[name]
    xr, yr, d = (int(v) for v in name.lower().split("x"))
    return name, (xr, yr, d)


def workload_text(wname, xr, yr, dim, world):
    return "%s: %d train x %d query x %d-D u8, brute-force L1 k=2%s" % (
        wname, xr, yr, dim, "" if world == 1 else ", train sharded %d-way + NCCL all-gather + top-2 merge" % world)


def config_of(wname, xr, yr, dim, world):
    """The SAME dictionary in both arms (the driver compares them): what differs between the arms -- how a step is
    sampled and timed -- is in `arm_notes` and `cpu_baseline.sample`."""
    return {"workload": workload_text(wname, xr, yr, dim, world),
            "l2": "GPU arm: L2 flushed between timed steps by a %d MiB fill outside the event pair; reference arm: host cores, n/a" % (L2_FLUSH_BYTES >> 20),
            "timing": "GPU arm: CUDA events per step on the launching stream, summed, max over ranks; reference arm: perf_counter around the K steps",
            "sample": "GPU arm: the full workload every step; reference arm: a bounded slice of query rows x all train rows per step (rows are independent), see cpu_baseline.sample"}


def host_rows(rows, dim, seed, lo=0):
    """Synthetic uniform u8 descriptors (test/test_feature.py:112-115's distribution),
    generated in 64k-row blocks keyed by (seed, block) so that any shard of any
    GPU count sees the same bytes."""
    import numpy as np
    out = np.empty((rows, dim), np.uint8)
    blk = 65536
    r = 0
    while r < rows:
        b = (lo + r) // blk
        off = (lo + r) % blk
        n = min(blk - off, rows - r)
        rng = np.random.Generator(np.random.Philox(key=seed * 1_000_003 + b))
        out[r:r + n] = rng.integers(0, 256, (blk, dim), dtype=np.uint8)[off:off + n]
        r += n
    return out


def sift_like_rows(rows, seed):
    """SURVEY 8d "S" data: rows sampled with replacement from the 128 descriptor columns of the reference's
    data/sift-test/sur-ogre.sift (tests/golden/l1k2_sift_surogre.npz) plus +-2 uniform integer noise, clipped.
    Real SIFT rows are clustered, which is what the CPU's early-exit prune (BruteForceNnL1K2.h:118-121) feeds on."""
    import numpy as np
    desc = np.load(os.path.join(ROOT, "tests", "golden", "l1k2_sift_surogre.npz"))["desc"]
    rng = np.random.Generator(np.random.Philox(key=seed))
    out = np.empty((rows, 128), np.uint8)
    for r in range(0, rows, 65536):
        n = min(65536, rows - r)
        base = desc[rng.integers(0, desc.shape[0], n)].astype(np.int16)
        out[r:r + n] = np.clip(base + rng.integers(-2, 3, (n, 128), dtype=np.int16), 0, 255).astype(np.uint8)
    return out


def result_hash(idx, dist):
    """blake2b-64 of the full (idx uint64 [yrows, 2], dist int32 [yrows, 2]) result."""
    import numpy as np
    h = hashlib.blake2b(digest_size=8)
    h.update(np.ascontiguousarray(idx).view(np.uint8).tobytes())
    h.update(np.ascontiguousarray(dist).view(np.uint8).tobytes())
    return h.hexdigest()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------- CPU legs
def cpu_matcher():
    """The reference's own code when oracle/_ref was built, else the C port."""
    from oracle import pyoracle as po
    if po.have_ref():
        return "reference", po.ref_l1k2
    return "port", (lambda x, y, nt: po.l1k2(x, y, nt))


def cpu_sample(x, yfull, budget_s, cores):
    """Time the CPU matcher on a contiguous slice of query rows against the FULL
    train set, sized to about budget_s seconds; returns (kind, Gpairs/s, rows, seconds, result)."""
    import numpy as np
    kind, fn = cpu_matcher()
    probe = min(64 * cores, yfull.shape[0])
    t0 = time.perf_counter()
    fn(x, np.ascontiguousarray(yfull[:probe]), cores)
    dt = max(time.perf_counter() - t0, 1e-6)
    rows = int(min(yfull.shape[0], max(probe, probe * budget_s / dt)))
    rows = max(cores, rows // cores * cores)
    ys = np.ascontiguousarray(yfull[:rows])
    t0 = time.perf_counter()
    res = fn(x, ys, cores)
    dt = time.perf_counter() - t0
    return kind, x.shape[0] * rows / dt / 1e9, rows, dt, res


def run_reference(args):
    """--impl reference: rank 0 only, host cores only, same metric/config."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    ngpu = max(world, args.gpus)
    wname, (xr, yr, dim) = workload_of(args, ngpu)
    cores = os.cpu_count() or 1
    x = host_rows(xr, dim, 11)
    kind, fn = cpu_matcher()
    # size a per-step slice of query rows to ~ (budget / steps) seconds, at least 1 s
    probe = min(64 * cores, yr)
    yp = host_rows(probe, dim, 13)
    t0 = time.perf_counter(); fn(x, yp, cores); dt = max(time.perf_counter() - t0, 1e-6)
    per_step_s = max(1.0, min(6.0, 60.0 / max(1, args.steps + args.warmup)))
    rows = int(min(yr, max(probe, probe * per_step_s / dt)))
    rows = max(cores, rows // cores * cores)
    y = host_rows(rows, dim, 13)
    for _ in range(args.warmup):
        fn(x, y, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        fn(x, y, cores)
    dt = time.perf_counter() - t0
    val = xr * rows * args.steps / dt / 1e9
    sample = "%d of %d query rows x all %d train rows per step (rows independent; pairs covered incl. pruned)" % (rows, yr, xr)
    extras = {}
    try:   # the same arm on 1 thread (the reference test's default, spectavi/feature.py:292) and on SIFT-like rows
        r1 = max(1, rows // cores)
        t0 = time.perf_counter(); fn(x, y[:r1].copy(), 1); d1 = time.perf_counter() - t0
        extras["one_thread"] = {"value": xr * r1 / d1 / 1e9, "unit": UNIT, "cores": 1, "sample": "%d query rows x all train rows" % r1}
        if dim == 128:
            xs, ys = sift_like_rows(xr, 21), sift_like_rows(rows, 23)
            t0 = time.perf_counter(); fn(xs, ys, cores); ds = time.perf_counter() - t0
            extras["sift_like_rows"] = {"value": xr * rows / ds / 1e9, "unit": UNIT, "cores": cores,
                                        "sample": "%d SIFT-like query rows x %d SIFT-like train rows (SURVEY 8d 'S' data: the prune bites)" % (rows, xr)}
    except Exception as ex:
        extras["error"] = repr(ex)[:200]
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": ngpu,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong" if ngpu > 1 else "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic uniform u8 (Philox, seeded per 64k-row block)",
            "config": config_of(wname, xr, yr, dim, ngpu), "arm_notes": {"sample": sample, "timing": "perf_counter around the K steps"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "extras": extras}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------ GPU arm
def pipe_peak(dev_index):
    """Issue rates measured on this GPU now (pipe_peak.cu): VABSDIFF4.U8.ACC, FP32, FP64 variants."""
    import ctypes as ct
    from spectavi_b200._libspectavi import clib
    out = (ct.c_double * 128)()
    names = (ct.c_char_p * 64)()
    lanes = (ct.c_double * 64)()
    clib.spv_pipe_peak.restype = ct.c_int
    clib.spv_pipe_peak.argtypes = [ct.c_int, ct.c_int, ct.c_void_p, ct.c_void_p, ct.c_void_p]
    n = clib.spv_pipe_peak(dev_index, 4000, out, names, lanes)
    if n <= 0:
        return None
    return {names[v].decode(): {"ops_per_clk_sm": out[2 * v], "gops": out[2 * v + 1]} for v in range(n)}


def k1f_query_sharded(world, device):
    """nn_bruteforce through the C ABI with SPECTAVI_GPUS=world (already set for this process): the query rows are split
    over the GPUs, every GPU holds all train rows, no exchange (host_ctx.cu HostCall::lpk).  Beside it: the same problem
    device-resident on ONE GPU (CUDA events), and equality of the two results."""
    import numpy as np
    import torch
    from spectavi_b200 import device as dev
    from spectavi_b200 import feature
    rows = 16384 * world
    rng = np.random.RandomState(4)
    x = rng.randn(rows, 132).astype("float32"); y = rng.randn(rows, 132).astype("float32")
    feature.nn_bruteforce(x, y, p=1., mu=0., k=2)
    best = 1e30
    for _ in range(3):
        t0 = time.perf_counter(); gi, gd = feature.nn_bruteforce(x, y, p=1., mu=0., k=2); best = min(best, time.perf_counter() - t0)
    dx, dy = torch.from_numpy(x).to(device), torch.from_numpy(y).to(device)
    dev.nn_bruteforce(dx, dy, p=1., k=2)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); di, dd = dev.nn_bruteforce(dx, dy, p=1., k=2); e1.record(); torch.cuda.synchronize()
    one = e0.elapsed_time(e1)
    return {"workload": "nn_bruteforce %d x %d x 132 float32, p=1, k=2" % (rows, rows), "gpus": world,
            "e2e_ms(C ABI, pageable numpy, query rows sharded)": best * 1e3, "Gpairs_per_s_e2e": rows * rows / best / 1e9,
            "one_gpu_device_resident_ms": one, "speedup_vs_one_gpu_device_resident": one / (best * 1e3),
            "equals_one_gpu_result": bool((gi.reshape(rows, 2) == di.cpu().numpy().view(np.uint64).reshape(rows, 2)).all()
                                          and (gd.reshape(rows, 2) == dd.cpu().numpy().reshape(rows, 2)).all())}


def guarded(out, key, fn):
    """A side measurement must never cost the bench line."""
    try:
        t0 = time.perf_counter()
        out[key] = fn()
        if isinstance(out[key], dict):
            out[key]["wall_s"] = round(time.perf_counter() - t0, 2)
    except Exception as ex:
        out[key] = {"error": repr(ex)[:300]}


def extra_c1():
    """C1 = BASELINE.json configs[0], 4096 x 4096 x 128: the reference's own CPU-runnable case, timed on the GPU
    (through the C ABI) and on the host with 1 thread (the reference test's default) and all cores."""
    from spectavi_b200 import feature
    kind, fn = cpu_matcher()
    x, y = host_rows(4096, 128, 31), host_rows(4096, 128, 33)
    feature.nn_bruteforcel1k2(x, y)
    best = 1e9
    for _ in range(10):
        t0 = time.perf_counter(); gi, gd = feature.nn_bruteforcel1k2(x, y); best = min(best, time.perf_counter() - t0)
    cores = os.cpu_count() or 1
    t0 = time.perf_counter(); o1 = fn(x, y, 1); t1 = time.perf_counter() - t0
    t0 = time.perf_counter(); fn(x, y, cores); tn = time.perf_counter() - t0
    pairs = 4096. * 4096.
    return {"workload": "c1: 4096 x 4096 x 128 u8", "gpu_e2e_ms": best * 1e3, "gpu_e2e_Gpairs_per_s": pairs / best / 1e9,
            "cpu_%s_1_thread_ms" % kind: t1 * 1e3, "cpu_1_thread_Gpairs_per_s": pairs / t1 / 1e9,
            "cpu_%s_%d_threads_ms" % (kind, cores): tn * 1e3, "cpu_all_cores_Gpairs_per_s": pairs / tn / 1e9,
            "parity_bit_exact": bool((gi == o1[0].reshape(gi.shape)).all() and (gd == o1[1].reshape(gd.shape)).all())}


def extra_c2():
    """The castle pair's SIFT rows (tests/golden/castle_c2.npz) through the example's step 2 via the host-buffer API."""
    import numpy as np
    from spectavi_b200 import feature
    g = np.load(os.path.join(ROOT, "tests", "golden", "castle_c2.npz"))
    s1 = np.hstack([g["k1"], g["d1"].astype("float32")])
    s2 = np.hstack([g["k2"], g["d2"].astype("float32")])
    xu = (feature.normalize_to_ubyte_and_multiple_16_dim(s1) + 128).astype("uint8")
    yu = (feature.normalize_to_ubyte_and_multiple_16_dim(s2) + 128).astype("uint8")
    feature.nn_bruteforcel1k2(xu, yu)
    best = 1e9
    for _ in range(8):
        t0 = time.perf_counter(); gi, gd = feature.nn_bruteforcel1k2(xu, yu); best = min(best, time.perf_counter() - t0)
    feature.match_bruteforcel1k2(xu, yu, 1.75)
    tm = 1e9
    for _ in range(5):
        t0 = time.perf_counter(); mx, my = feature.match_bruteforcel1k2(xu, yu, 1.75); tm = min(tm, time.perf_counter() - t0)
    kind, cfn = cpu_matcher()
    t0 = time.perf_counter(); cfn(xu, yu, os.cpu_count() or 1); cpu_ms = (time.perf_counter() - t0) * 1e3
    return {"rows": [int(xu.shape[0]), int(yu.shape[0])], "dim": int(xu.shape[1]),
            "cpu_%s_ms(all host cores, real SIFT rows: the prune of BruteForceNnL1K2.h:118-121 bites)" % kind: cpu_ms,
            "nn_bruteforcel1k2_ms_e2e": best * 1e3, "Gpairs_per_s_e2e": xu.shape[0] * yu.shape[0] / best / 1e9,
            "match_ms_e2e": tm * 1e3, "matches": int(mx.shape[0]),
            "equals_reference_output": bool((gi == g["idx"].astype(np.uint64)).all() and (gd == g["dist"]).all())}


def extra_c3_sift(device, peak_gops):
    """C3 on SIFT-like ("S") rows: GPU speed is data-independent (no prune), the CPU's is not."""
    import numpy as np
    import torch
    from spectavi_b200 import device as dev
    xr = yr = 200_000
    hx, hy = sift_like_rows(xr, 21), sift_like_rows(yr, 23)
    dx, dy = torch.from_numpy(hx).to(device), torch.from_numpy(hy).to(device)
    keys = dev.new_keys(yr, device)
    dev.l1k2_fold(dx, dy, keys, 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for _ in range(3):
        dev.keys_reset(keys)
        e0.record(); dev.l1k2_fold(dx, dy, keys, 0); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    gi, gd = dev.keys_finalize(keys)
    cores = os.cpu_count() or 1
    kind, gps, rows, secs, (oi, od) = cpu_sample(hx, hy, 6.0, cores)
    oi, od = oi.reshape(-1, 2), od.reshape(-1, 2)
    k1 = min(ms)
    out = {"workload": "c3 on SIFT-like rows (sur-ogre.sift descriptors + noise), 200000 x 200000 x 128", "k1_ms": k1,
           "Gpairs_per_s": xr * yr / k1 / 1e6, "cpu_%s_Gpairs_per_s" % kind: gps, "cpu_cores": cores,
           "cpu_sample": "%d query rows, %.1f s" % (rows, secs),
           "parity_rows_checked": int(oi.shape[0]),
           "parity_bit_exact": bool((gi[:oi.shape[0]].cpu().numpy().view(np.uint64) == oi).all() and (gd[:oi.shape[0]].cpu().numpy() == od).all())}
    if peak_gops:
        out["frac_of_pipe_peak"] = xr * yr * 32. / (k1 * 1e-3) / 1e9 / peak_gops
    return out


def extra_c4_n1(device, peak_gops):
    """C4 (1M x 1M) on ONE GPU: the like-for-like base of the strong-scaling curve, and the hash every N > 1 line must repeat."""
    import torch
    from spectavi_b200 import device as dev
    xr = yr = 1_000_000
    dx, dy = torch.from_numpy(host_rows(xr, 128, 11)).to(device), torch.from_numpy(host_rows(yr, 128, 13)).to(device)
    keys = dev.new_keys(yr, device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for _ in range(2):
        dev.keys_reset(keys)
        e0.record(); dev.l1k2_fold(dx, dy, keys, 0); gi, gd = dev.keys_finalize(keys); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    out = {"workload": workload_text("c4", xr, yr, 128, 1), "ms_per_step": min(ms), "Gpairs_per_s": xr * yr / min(ms) / 1e6,
           "result_hash": result_hash(gi.cpu().numpy(), gd.cpu().numpy())}
    if peak_gops:
        out["frac_of_pipe_peak"] = xr * yr * 32. / (min(ms) * 1e-3) / 1e9 / peak_gops
    return out


def side_measurements(device, peak_gops):
    """Not part of the metric: BASELINE.json's other configs and the neighbours of the path, measured in the same run."""
    out = {}
    guarded(out, "c1", extra_c1)
    guarded(out, "c2_castle", extra_c2)
    guarded(out, "c3_sift_like", lambda: extra_c3_sift(device, peak_gops))
    guarded(out, "c4_n1", lambda: extra_c4_n1(device, peak_gops))

    def lp():
        import bench_lp
        return bench_lp.run(20000, 3)
    guarded(out, "k1f_nn_bruteforce_f32", lp)

    def cascade():
        import bench_cascade
        return bench_cascade.run(1_000_000, 128, 3, 1000)
    guarded(out, "c5_cascade", cascade)

    def cascade_castle():
        import bench_cascade
        return bench_cascade.run_castle()
    guarded(out, "c2_castle_cascade", cascade_castle)

    def kmed():
        import bench_kmedians
        return bench_kmedians.run(10000, 132, 5)
    guarded(out, "kmedians_nn", kmed)

    def ransac():
        import bench_ransac
        return bench_ransac.run()
    guarded(out, "k8_ransac_scoring", ransac)
    return out


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU path (use --impl reference)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
        # rank 0 later drives ALL GPUs through the C ABI while the others wait: that wait must not be an NCCL
        # kernel spinning on their SMs, so it goes through a CPU (gloo) group
        cpu_group = dist.new_group(backend="gloo", timeout=datetime.timedelta(minutes=60))
    from spectavi_b200 import device as dev
    from spectavi_b200 import feature

    wname, (xr, yr, dim) = workload_of(args, world)
    from spectavi_b200.sharding import shard_range
    x0, xn = shard_range(xr, world, rank)      # train rows of this rank
    hx = torch.from_numpy(host_rows(xn, dim, 11, lo=x0)).pin_memory()
    hy = torch.from_numpy(host_rows(yr, dim, 13)).pin_memory()
    dx, dy = hx.to(device), hy.to(device)
    keys = dev.new_keys(yr, device)
    gathered = torch.empty((world, 2, yr), dtype=torch.int64, device=device) if world > 1 else None
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(ek0=None, ek1=None):
        dev.keys_reset(keys)
        if ek0 is not None:
            ek0.record()
        dev.l1k2_fold(dx, dy, keys, x0)
        if ek1 is not None:
            ek1.record()
        if world > 1:
            dist.all_gather_into_tensor(gathered, keys)
            dev.keys_merge(keys, gathered)
        return dev.keys_finalize(keys)

    for _ in range(max(args.warmup, 3) if args.warmup else 0):
        flush.fill_(1)
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(args.steps)]
    l0 = dev.launch_count()
    barrier()
    for s in range(args.steps):
        flush.fill_(s & 0xFF)                   # L2 flush between timed iterations (outside the event pair)
        ev[s][0].record()
        idx, dd = step(ev[s][1], ev[s][2])
        ev[s][3].record()
    barrier()
    l1 = dev.launch_count()
    clocks = sampler.stop() if rank == 0 else None
    step_ms = sum(e[0].elapsed_time(e[3]) for e in ev)
    k1_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    t = torch.tensor([step_ms, k1_ms], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_ms, k1_ms = float(t[0]), float(t[1])
    pairs = float(xr) * float(yr)
    value = pairs * args.steps / (step_ms * 1e-3) / 1e9
    result_idx, result_dist = idx, dd

    # ---- torchrun layout end to end (N > 1): pinned shard + query rows up, sharded match, result to rank 0's host
    torchrun_e2e = None
    if world > 1 and not args.no_e2e:
        def estep():
            ex = hx.to(device, non_blocking=True)
            ey = hy.to(device, non_blocking=True)
            gi, gd = dev.nn_bruteforcel1k2_sharded(ex, ey, x0)
            if rank == 0:
                return gi.cpu(), gd.cpu()
            return None
        for _ in range(2):
            estep()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            estep()
        barrier()
        e_s = time.perf_counter() - t0
        tt = torch.tensor([e_s], dtype=torch.float64, device=device)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e_s = float(tt[0])
        torchrun_e2e = {"value": pairs * args.steps / e_s / 1e9, "unit": UNIT, "ms_per_step": e_s / args.steps * 1e3,
                        "api": "device.nn_bruteforcel1k2_sharded from pinned host buffers (one process per GPU), result to host on rank 0"}
        barrier()

    c5_sharded = None
    if world > 1 and not args.no_extras:      # C5 at N GPUs (BASELINE.json configs[4]); every rank takes part
        try:
            import bench_cascade
            c5_sharded = bench_cascade.run_sharded(device, world, rank)
        except Exception as ex:
            c5_sharded = {"error": repr(ex)[:300]}
        barrier()

    if rank != 0:
        dist.barrier(group=cpu_group)          # CPU-side wait: rank 0 is about to use this GPU through the C ABI
        dist.destroy_process_group()
        return 0

    # ---- rank 0 from here on.  e2e: the drop-in call on pageable numpy arrays
    e2e = None
    c_abi = {}
    if not args.no_e2e:
        xfull = np.empty((xr, dim), np.uint8)       # ordinary pageable arrays, as a user's are
        yfull = np.empty((yr, dim), np.uint8)
        if world == 1:
            xfull[...] = hx.numpy()
        else:
            xfull[...] = host_rows(xr, dim, 11)
            os.environ["SPECTAVI_GPUS"] = str(world)   # read once, at the library's first host-buffer call
        yfull[...] = hy.numpy()
        try:
            for _ in range(2):
                feature.nn_bruteforcel1k2(xfull, yfull)
            torch.cuda.synchronize()
            esteps = args.steps if world == 1 else max(3, min(args.steps, 5))
            la = dev.launch_count()
            t0 = time.perf_counter()
            for _ in range(esteps):
                ei, ed = feature.nn_bruteforcel1k2(xfull, yfull)
            e_s = time.perf_counter() - t0
            lb = dev.launch_count()
            e2e = {"value": pairs * esteps / e_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": int((xr + yr * world) * dim),
                   "d2h_bytes_per_step": int(yr * 2 * (8 + 4)), "ms_per_step": e_s / esteps * 1e3, "steps": esteps,
                   "kernel_launches_per_step": int((lb - la) // esteps),
                   "api": "feature.nn_bruteforcel1k2 (C ABI nn_bruteforcel1k2) on pageable numpy arrays" +
                          ("" if world == 1 else "; SPECTAVI_GPUS=%d: one process, one host thread per GPU, NCCL all-gather of the keys" % world)}
            gi_t = result_idx.cpu().numpy().view(np.uint64)
            c_abi = {"equals_device_resident_result": bool((ei == gi_t).all() and (ed == result_dist.cpu().numpy()).all()),
                     "result_hash": result_hash(ei, ed)}
            e2e.update(c_abi)
        except Exception as ex:
            e2e = None
            c_abi = {"error": repr(ex)[:300]}
        if e2e is None and torchrun_e2e is not None:   # keep an end-to-end figure if the C-ABI leg failed
            e2e = dict(torchrun_e2e, h2d_bytes_per_step=int((xn + yr) * dim), d2h_bytes_per_step=int(yr * 24))

    # ---- roofline
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    sad4_per_launch = float(xn) * yr * dim / 4           # algorithmic packed-SAD ops of one K1 launch
    achieved = sad4_per_launch / (k1_ms * 1e-3) / 1e9   # G sad4-ops / s
    roof = {"bound": "alu", "kernel": "l1_top2_kernel", "achieved": achieved, "peak": None, "unit": "G sad4-ops/s",
            "frac": None, "traffic": None, "k1_ms": k1_ms,
            "algorithmic": "dim/4 = %d VABSDIFF4.U8.ACC per pair x %d x %d pairs per launch" % (dim // 4, xn, yr),
            "hbm_floor": {"bytes": int((xn + yr) * dim + yr * 16), "achieved_GBps": ((xn + yr) * dim + yr * 16) / (k1_ms * 1e-3) / 1e9,
                          "peak_GBps": peaks.get("hbm_gbs", 6650.0),
                          "peak_source": "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback"}}
    sms = torch.cuda.get_device_properties(local).multi_processor_count
    mhz = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
    nominal = 64.0 * sms * mhz * 1e6 / 1e9
    roof["peak_nominal"] = nominal
    roof["frac_nominal"] = achieved / nominal
    roof["peak_nominal_source"] = "64 VABSDIFF4 per clk per SM (half-rate ALU pipe) x %d SMs x %.0f MHz" % (sms, mhz)
    peak_gops = None
    if not args.no_peak:
        tab = pipe_peak(local)
        if tab:
            sad = [v for k, v in tab.items() if k.startswith("sad4_acc(")][0]
            peak_gops = sad["gops"]
            roof["peak"] = sad["gops"]
            roof["frac"] = achieved / sad["gops"]
            roof["peak_source"] = ("spv_pipe_peak in this run (pipe_peak.cu): VABSDIFF4.U8.ACC issues %.0f G thread-ops/s on an idle "
                                   "GPU (burst, CUDA events) = %.1f per clk per SM at %d SMs x %.0f MHz (nominal 64: half-rate ALU pipe)"
                                   % (sad["gops"], sad["gops"] * 1e3 / (sms * mhz), sms, mhz))
            roof["pipe_table_gops"] = {k: round(v["gops"], 0) for k, v in tab.items()}
    try:
        prof = json.load(open(os.path.join(ROOT, "profiles", "k1_traffic.json")))
        if world == 1 and wname == "c3":      # the ncu capture is of this workload's launch
            roof["traffic"] = prof.get("dram_bytes_per_launch")
            roof["traffic_source"] = prof.get("source")
    except Exception:
        pass

    # ---- CPU baseline on a bounded sample; its results double as the parity check of the timed GPU output, at every N
    cpu = None
    parity = {"result_hash": result_hash(result_idx.cpu().numpy().view(np.uint64), result_dist.cpu().numpy())}
    if not args.no_cpu:
        cores = os.cpu_count() or 1
        xall = hx.numpy() if world == 1 else host_rows(xr, dim, 11)
        kind, gps, rows, secs, (oi, od) = cpu_sample(xall, hy.numpy(), args.cpu_seconds, cores)
        cpu = {"value": gps, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": "first %d of %d query rows x all %d train rows, %.1f s wall, extrapolated by rows (rows independent)" % (rows, yr, xr, secs)}
        oi, od = oi.reshape(-1, 2), od.reshape(-1, 2)
        gi = result_idx[:oi.shape[0]].cpu().numpy().view(np.uint64)
        gd = result_dist[:oi.shape[0]].cpu().numpy()
        parity["parity_rows_checked"] = int(oi.shape[0])
        parity["parity_bit_exact"] = bool((gi == oi).all() and (gd == od).all())
        cpu.update(parity)
        try:   # 1 thread: the reference test's default (spectavi/feature.py:292)
            kindf, fn = cpu_matcher()
            r1 = max(16, rows // cores)
            t0 = time.perf_counter(); fn(xall, np.ascontiguousarray(hy.numpy()[:r1]), 1); d1 = time.perf_counter() - t0
            cpu["one_thread"] = {"value": xr * r1 / d1 / 1e9, "unit": UNIT, "cores": 1, "sample": "%d query rows x all train rows" % r1}
        except Exception as ex:
            cpu["one_thread"] = {"error": repr(ex)[:200]}

    extras = {}
    if not args.no_extras:
        if world == 1:
            extras = side_measurements(device, peak_gops)
        else:
            if torchrun_e2e is not None:
                extras["torchrun_layout_e2e"] = torchrun_e2e
            extras["c_abi_mgpu"] = dict(c_abi, note="timed as the e2e figure of this line")
            if c5_sharded is not None:
                extras["c5_cascade_sharded"] = c5_sharded
            guarded(extras, "k1f_query_sharded", lambda: k1f_query_sharded(world, device))

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3) if args.warmup else 0, "ms_per_step": step_ms / args.steps,
            "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic uniform u8 (Philox, seeded per 64k-row block)",
            "config": config_of(wname, xr, yr, dim, world),
            "arm_notes": {"sample": "full workload every step", "timing": "CUDA events per step on the launching stream, summed, max over ranks"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(l1 - l0), "roofline": roof, "cpu_baseline": cpu}
    line.update(parity)
    if extras:
        line["extras"] = extras
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(group=cpu_group)
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
