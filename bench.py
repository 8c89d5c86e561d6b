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
`e2e`     : the same through the host-buffer API (nn_bruteforcel1k2 C ABI at N == 1,
            device.nn_bruteforcel1k2_sharded from pinned host buffers at N > 1),
            H2D of the descriptors and D2H of (idx, dist) inside the timed region
`roofline`: K1's packed-SAD rate against the VABSDIFF4 issue peak measured in the
            same run by spv_pipe_peak (the path is ALU-bound)
`cpu_baseline` / --impl reference: the reference's own BruteForceNnL1K2 (oracle/_ref,
            compiled from /root/reference through oracle/eigen_shim; else the C port)
            on all host cores over a bounded slice of query rows against the full
            train set (rows are independent, BruteForceNnL1K2.h:92-93).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

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
    ap.add_argument("--no-extras", action="store_true", help="skip the C2 (castle) and C5 (cascade) side measurements")
    return ap.parse_args()


def workload_of(args, world):
    name = args.workload or ("c3" if world == 1 else "c4")
    if name in WORKLOADS:
        return name, WORKLOADS[name]
    xr, yr, d = (int(v) for v in name.lower().split("x"))
    return name, (xr, yr, d)


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
    train set, sized to about budget_s seconds; returns (Gpairs/s, rows, seconds)."""
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
    import numpy as np
    wname, (xr, yr, dim) = workload_of(args, max(world, args.gpus))
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
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": max(world, args.gpus),
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "strong" if max(world, args.gpus) > 1 else "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic uniform u8",
            "config": {"workload": "%s: %d train x %d query x %d-D u8, brute-force L1 k=2" % (wname, xr, yr, dim),
                       "sample": sample},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------ GPU arm
def pipe_peak(dev_index):
    """P_sad4: VABSDIFF4.U8.ACC thread-ops/s measured on this GPU now (pipe_peak.cu)."""
    import ctypes as ct
    from spectavi_b200._libspectavi import clib
    out = (ct.c_double * 64)()
    names = (ct.c_char_p * 32)()
    lanes = (ct.c_double * 32)()
    clib.spv_pipe_peak.restype = ct.c_int
    clib.spv_pipe_peak.argtypes = [ct.c_int, ct.c_int, ct.c_void_p, ct.c_void_p, ct.c_void_p]
    n = clib.spv_pipe_peak(dev_index, 4000, out, names, lanes)
    if n <= 0:
        return None
    tab = {names[v].decode(): {"ops_per_clk_sm": out[2 * v], "gops": out[2 * v + 1]} for v in range(n)}
    return tab


def side_measurements(device):
    """Not part of the metric: BASELINE.json's other configs measured in the same run, for the record.
    c2: the castle pair's SIFT rows (tests/golden/castle_c2.npz) through the example's step 2 via the
    host-buffer API; c5: the cascade-of-hashes matcher at 1M x 1M against brute force (tools/bench_cascade.py)."""
    import numpy as np
    out = {}
    try:
        from spectavi_b200 import feature
        g = np.load(os.path.join(ROOT, "tests", "golden", "castle_c2.npz"))
        s1 = np.hstack([g["k1"], g["d1"].astype("float32")])
        s2 = np.hstack([g["k2"], g["d2"].astype("float32")])
        xu = (feature.normalize_to_ubyte_and_multiple_16_dim(s1) + 128).astype("uint8")
        yu = (feature.normalize_to_ubyte_and_multiple_16_dim(s2) + 128).astype("uint8")
        feature.nn_bruteforcel1k2(xu, yu)
        best = 1e9
        for _ in range(5):
            t0 = time.perf_counter(); gi, gd = feature.nn_bruteforcel1k2(xu, yu); best = min(best, time.perf_counter() - t0)
        t0 = time.perf_counter(); mx, my = feature.match_bruteforcel1k2(xu, yu, 1.75); tm = time.perf_counter() - t0
        cpu_ms = None
        try:
            from oracle import pyoracle as po
            cfn = po.ref_l1k2 if po.have_ref() else po.l1k2
            t0 = time.perf_counter(); cfn(xu, yu, os.cpu_count() or 1); cpu_ms = (time.perf_counter() - t0) * 1e3
        except Exception:
            pass
        out["c2_castle"] = {"rows": [int(xu.shape[0]), int(yu.shape[0])], "dim": int(xu.shape[1]),
                            "cpu_reference_ms(all host cores, real SIFT rows: the prune of BruteForceNnL1K2.h:118-121 bites)": cpu_ms,
                            "nn_bruteforcel1k2_ms_e2e": best * 1e3, "Gpairs_per_s_e2e": xu.shape[0] * yu.shape[0] / best / 1e9,
                            "match_ms_e2e": tm * 1e3, "matches": int(mx.shape[0]),
                            "equals_reference_output": bool((gi == g["idx"].astype(np.uint64)).all() and (gd == g["dist"]).all())}
    except Exception as ex:                                  # a side measurement must never cost the bench line
        out["c2_castle"] = {"error": repr(ex)[:200]}
    try:
        # the float matcher on the reference test's shape family (test/test_feature.py:67-81), p = 2, k = 2
        from spectavi_b200 import feature
        from oracle import pyoracle as po
        rng = np.random.RandomState(5)
        res = {}
        for rows in (1000, 20000):
            xf = rng.randn(rows, 132).astype("float32")
            yf = rng.randn(rows, 132).astype("float32")
            feature.nn_bruteforce(xf, yf, p=2., mu=0., k=2)
            best = 1e9
            for _ in range(3):
                t0 = time.perf_counter(); fi, fd = feature.nn_bruteforce(xf, yf, p=2., mu=0., k=2); best = min(best, time.perf_counter() - t0)
            ent = {"ms_e2e": best * 1e3, "Gpairs_per_s_e2e": rows * rows / best / 1e9}
            if rows == 1000:
                cfn = po.ref_bruteforce if po.have_ref() else po.bruteforce
                t0 = time.perf_counter(); oi, od = cfn(xf, yf, 2, 2., 0.); ent["cpu_reference_ms(1 thread: BruteForceNn is single-threaded)"] = (time.perf_counter() - t0) * 1e3
                ent["idx_equal"] = bool((fi == oi).all()); ent["dist_bit_exact"] = bool((fd.view(np.uint32) == od.view(np.uint32)).all())
            res["%dx%dx132" % (rows, rows)] = ent
        out["nn_bruteforce_f32_p2"] = res
    except Exception as ex:
        out["nn_bruteforce_f32_p2"] = {"error": repr(ex)[:200]}
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_cascade
        out["c5_cascade"] = bench_cascade.run(1_000_000, 128, 3, 1000)
    except Exception as ex:
        out["c5_cascade"] = {"error": repr(ex)[:200]}
    try:   # KMedians::find_nearest_neighbours at the castle pair's size
        import bench_kmedians
        out["kmedians_nn"] = bench_kmedians.run(10000, 132, 5)
    except Exception as ex:
        out["kmedians_nn"] = {"error": repr(ex)[:200]}
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
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
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

    # ---- e2e: host buffers in, host results out, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        esteps = args.steps
        h2d = (xn + yr) * dim
        d2h = yr * 2 * (8 + 4)
        if world == 1:
            # ordinary pageable numpy arrays, as a user's are (NOT views of the pinned tensors above)
            nx, ny = np.empty(hx.shape, np.uint8), np.empty(hy.shape, np.uint8)
            nx[...] = hx.numpy(); ny[...] = hy.numpy()
            for _ in range(2):
                feature.nn_bruteforcel1k2(nx, ny)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(esteps):
                ei, ed = feature.nn_bruteforcel1k2(nx, ny)
            e_s = time.perf_counter() - t0
        else:
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
            for _ in range(esteps):
                estep()
            barrier()
            e_s = time.perf_counter() - t0
            tt = torch.tensor([e_s], dtype=torch.float64, device=device)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            e_s = float(tt[0])
        e2e = {"value": pairs * esteps / e_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h if rank == 0 else 0), "ms_per_step": e_s / esteps * 1e3,
               "api": "feature.nn_bruteforcel1k2 (C ABI nn_bruteforcel1k2, host numpy buffers)" if world == 1 else
                      "device.nn_bruteforcel1k2_sharded from pinned host buffers, result to host on rank 0"}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- rank 0: roofline, cpu baseline, the JSON line
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
    if not args.no_peak:
        tab = pipe_peak(local)
        if tab:
            sad = [v for k, v in tab.items() if k.startswith("sad4_acc(")][0]
            roof["peak"] = sad["gops"]
            roof["frac"] = achieved / sad["gops"]
            sms = torch.cuda.get_device_properties(local).multi_processor_count
            mhz = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
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

    cpu = None
    if not args.no_cpu:
        cores = os.cpu_count() or 1
        xfull = hx.numpy() if world == 1 else host_rows(xr, dim, 11)
        kind, gps, rows, secs, (oi, od) = cpu_sample(xfull, hy.numpy(), args.cpu_seconds, cores)
        cpu = {"value": gps, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": "first %d of %d query rows x all %d train rows, %.1f s wall, extrapolated by rows (rows independent)" % (rows, yr, xr, secs)}
        if world == 1:
            # the CPU leg's results double as a parity check of the timed GPU output: every sampled query row
            oi, od = oi.reshape(-1, 2), od.reshape(-1, 2)
            gi = result_idx[:oi.shape[0]].cpu().numpy().view(np.uint64)
            gd = result_dist[:oi.shape[0]].cpu().numpy()
            cpu["parity_rows_checked"] = int(oi.shape[0])
            cpu["parity_bit_exact"] = bool((gi == oi).all() and (gd == od).all())

    extras = None
    if world == 1 and not args.no_extras:
        extras = side_measurements(device)

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3) if args.warmup else 0, "ms_per_step": step_ms / args.steps,
            "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic uniform u8 (Philox, seeded per 64k-row block)",
            "config": {"workload": "%s: %d train x %d query x %d-D u8, brute-force L1 k=2%s" % (
                           wname, xr, yr, dim, "" if world == 1 else ", train sharded %d-way + NCCL all-gather + top-2 merge" % world),
                       "l2": "L2 flushed between timed steps by a %d MiB fill outside the event pair" % (L2_FLUSH_BYTES >> 20),
                       "timing": "CUDA events per step on the launching stream, summed, max over ranks"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(l1 - l0), "roofline": roof, "cpu_baseline": cpu}
    if extras:
        line["extras"] = extras
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
