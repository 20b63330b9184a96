"""SASS evidence for the pair kernel and the builders: per-kernel instruction histogram of libkrd_b200.so from
`cuobjdump -sass` (runs without a GPU).  Usage: python profiles/sass_histogram.py > profiles/r2_sass_histogram.txt
What to look for: DFMA/DADD (FP64 CUDA-core operator), DMMA (mma.sync.m8n8k4.f64 -> FP64 tensor path of the Gram mode),
UBLKCP (cp.async.bulk: TMA engine, 1-D bulk copies into the shared-memory ring), SYNCS (mbarrier), LDS.128/.64 (operand fetch),
MUFU.RSQ64H (seed of the own sqrt, p = 1/2)."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gappa_b200", "libkrd_b200.so")
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
KEEP = ("DFMA", "DADD", "DMUL", "DMMA", "UBLKCP", "SYNCS", "LDS", "STS", "LDG", "STG", "MUFU", "BAR", "BRA", "IMAD", "LOP3", "SEL", "ISETP", "MOV", "DSETP", "F2I", "I2F", "SHFL", "ATOM", "RED", "CALL")
cur, hist = None, collections.OrderedDict()
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1); hist[cur] = collections.Counter(); continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1)
        hist[cur]["_total"] += 1
        base = op.split(".")[0]
        key = op if base in ("LDS", "MUFU", "DMMA", "UBLKCP", "SYNCS") else base
        if base in KEEP:
            hist[cur][key] += 1
dem = subprocess.run(["cu++filt"], input="\n".join(hist.keys()), capture_output=True, text=True).stdout.splitlines()
print("instruction histogram per kernel (cuobjdump -sass %s)" % os.path.relpath(lib, ROOT))
for (name, h), d in zip(hist.items(), dem):
    d = re.sub(r"\(.*", "", d.replace("(int)", "").replace("(bool)", ""))
    if not any(k in d for k in ("k_pairs", "k_tile", "k_exact", "k_subtree", "k_gram", "k_epilogue", "k_sq_one", "k_km_assign", "k_peak", "k_unpack")):
        continue
    items = sorted(((k, v) for k, v in h.items() if k != "_total"), key=lambda kv: -kv[1])
    print("\n%s   [%d instructions]" % (d, h["_total"]))
    print("   " + "  ".join("%s=%d" % kv for kv in items[:18]))
