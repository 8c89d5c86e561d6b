"""stdin: `cuobjdump -sass` text; stdout: one line per kernel with a histogram of the telling opcodes."""
import collections
import re
import subprocess
import sys

KEEP_FULL = ("UTMALDG", "UTMASTG", "LDTM", "STTM", "MUFU", "VABSDIFF4", "VABSDIFF")
KEEP_BASE = ("UTCIMMA", "UTCHMMA", "UTCQMMA", "UTCBAR", "SYNCS", "IDP", "FADD2", "FMUL2", "FFMA2", "FADD", "FMUL", "FFMA",
             "DADD", "DMUL", "DFMA", "LDS", "STS", "LDG", "STG", "LDGSTS", "ATOMS", "ATOMG", "ATOM", "RED", "STL", "LDL", "BAR",
             "SHFL", "VOTE", "VIMNMX", "VIMNMX3", "IMAD", "LOP3", "IADD3", "POPC", "REDUX")
ins = re.compile(r"^\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)")
kern, total, hist = None, 0, None
rows = []


def flush():
    if kern is not None:
        rows.append((kern, total, hist))


for line in sys.stdin:
    if "Function :" in line:
        flush()
        kern, total, hist = line.split("Function :")[1].strip(), 0, collections.Counter()
        continue
    m = ins.match(line)
    if not m or kern is None:
        continue
    op = m.group(1)
    base = op.split(".")[0]
    total += 1
    if base in KEEP_FULL:
        hist[".".join(op.split(".")[:3])] += 1
    elif base in KEEP_BASE:
        hist[base] += 1
flush()
names = [r[0] for r in rows]
try:
    dem = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
except OSError:
    dem = names
for (k, t, h), d in sorted(zip(rows, dem), key=lambda z: z[1]):
    d = re.sub(r"spv::\(anonymous namespace\)::", "", d)
    d = re.sub(r"\(.*", "", d)
    print("%s | %d | %s" % (d, t, " ".join("%s=%d" % kv for kv in sorted(h.items()))))
