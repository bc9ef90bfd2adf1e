#!/usr/bin/env python3
"""Dynamic opcode histogram of one kernel from an ncu report's source page (per warp, i.e. per thread when not divergent).
usage: ncu_dyn_hist.py <report.ncu-rep> <kernel regex> [warps]"""
import collections, csv, io, re, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr = None
h = collections.Counter(); stall = collections.Counter(); total = 0; samples = 0
for r in rows:
    if r and r[0] == "Address":
        hdr = {n: i for i, n in enumerate(r)}; continue
    if hdr is None or len(r) < len(hdr): continue
    ins = re.sub(r"^@!?U?P\d+\s+", "", r[hdr["Source"]].strip())
    op = ins.split()[0]
    base = op.split(".")[0]
    key = ".".join(op.split(".")[:2]) if base in ("IMAD", "LDG", "STG", "LDS", "STS") else base
    n = int(r[hdr["Instructions Executed"]])
    h[key] += n; total += n
    s = int(r[hdr["# Samples"]]); stall[key] += s; samples += s
warps = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
print(f"total warp-instructions {total}  per warp {total / warps:.1f}")
for k, v in h.most_common(40):
    print(f"  {k:12s} {v / warps:9.1f}  {100.0 * v / total:5.1f}%   samples {100.0 * stall[k] / max(samples, 1):5.1f}%")
