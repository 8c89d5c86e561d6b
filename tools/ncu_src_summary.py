#!/usr/bin/env python
"""Summarise an ncu --set full report's source page: warp-time by opcode class and
by non-SAD code segment.  usage: ncu_src_summary.py report.ncu-rep"""
import csv, subprocess, sys, io, re
rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; ix = {k: i for i, k in enumerate(h)}
seen = {}
for r in rows[hi + 1:]:
    if len(r) > 10 and r[0].startswith("0x"): seen.setdefault(r[0], r)
data = list(seen.values())
S = lambda r, k: int(r[ix[k]])
tot = sum(S(r, "# Samples") for r in data)
print("instructions %d, samples %d" % (len(data), tot))
keys = ["stall_selected", "stall_wait", "stall_math", "stall_not_selected", "stall_short_sb", "stall_long_sb",
        "stall_branch_resolving", "stall_no_inst", "stall_dispatch", "stall_barrier", "stall_mio", "stall_lg"]
def cls(src):
    if "VABSDIFF4" in src: return "SAD"
    if "LDS" in src: return "LDS"
    if "SYNCS" in src or "BRA" in src: return "BRA/SYNCS"
    return "other"
agg = {}
for r in data:
    c = cls(r[ix["Source"]]); a = agg.setdefault(c, {"n": 0, "exec": 0, "samples": 0, **{k: 0 for k in keys}})
    a["n"] += 1; a["exec"] += S(r, "Instructions Executed"); a["samples"] += S(r, "# Samples")
    for k in keys: a[k] += S(r, k)
for c, a in agg.items():
    print("%-10s n=%5d exec=%.3e samples=%8d (%.1f%%) " % (c, a["n"], a["exec"], a["samples"], 100.0 * a["samples"] / tot) +
          " ".join("%s=%.1f%%" % (k[6:], 100.0 * a[k] / tot) for k in keys if a[k] * 200 > tot))
base = int(data[0][0], 16); seg = []; cur = []
for r in data:
    if cls(r[ix["Source"]]) in ("SAD", "LDS"):
        if cur: seg.append(cur); cur = []
    else: cur.append(r)
if cur: seg.append(cur)
for sg in seg:
    t = sum(S(r, "# Samples") for r in sg)
    if t * 300 < tot: continue
    print("--- non-SAD segment @%x n=%d samples=%d (%.1f%%)" % (int(sg[0][0], 16) - base, len(sg), t, 100.0 * t / tot))
    for r in sorted(sg, key=lambda r: -S(r, "# Samples"))[:8]:
        print("     %5x %-56s %7d exec %s" % (int(r[0], 16) - base, r[ix["Source"]][:56], S(r, "# Samples"), r[ix["Instructions Executed"]]))
