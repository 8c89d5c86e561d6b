"""Pinned host -> device copy rate against transfer size, and after a pause (the drop-in call always starts from an
idle link).  torch is plumbing here: cudaMemcpyAsync from cudaHostAlloc memory, CUDA events around it."""
import json, time, torch
out = {}
for mb in (1, 4, 10, 21, 64, 256):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device="cuda")
    h.fill_(1)
    d.copy_(h, non_blocking=True); torch.cuda.synchronize()
    res = {}
    for label, pause in (("back_to_back", 0.0), ("after_20ms_idle", 0.02)):
        best = 1e9; ts = []
        for _ in range(5):
            if pause: time.sleep(pause)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); d.copy_(h, non_blocking=True); e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        res[label + "_GBps"] = round(n / min(ts) / 1e6, 1)
        res[label + "_ms"] = round(min(ts), 3)
    # the same bytes as 1 MB pieces on 6 streams
    ss = [torch.cuda.Stream() for _ in range(6)]
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(0, n, 1 << 20):
        with torch.cuda.stream(ss[(i >> 20) % 6]):
            d[i:i + (1 << 20)].copy_(h[i:i + (1 << 20)], non_blocking=True)
    torch.cuda.synchronize(); res["six_streams_1MB_pieces_wall_ms"] = round((time.perf_counter() - t0) * 1e3, 3)
    out["%d MB" % mb] = res
print(json.dumps(out, indent=1))
