#!/usr/bin/env python3
"""Per-source-line dynamic instruction counts of one kernel from an ncu report (needs -lineinfo and --import-source on).
usage: ncu_line_hist.py <report.ncu-rep> <kernel regex> <warps launched> [top]"""
import csv, io, subprocess, sys
rep, pat, warps = sys.argv[1], sys.argv[2], float(sys.argv[3])
top = int(sys.argv[4]) if len(sys.argv) > 4 else 60
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name", "regex:" + pat],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur_file = ""; hdr = None; out = []; total = 0; samples = 0
for r in rows:
    if len(r) >= 2 and r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = {n: i for i, n in enumerate(r)}; ie = r.index("Instructions Executed"); sm = r.index("# Samples"); continue
    if hdr is None or len(r) <= ie or not r[0].strip().isdigit(): continue
    if not r[ie].isdigit(): continue
    n = int(r[ie]); s = int(r[sm]) if r[sm].isdigit() else 0
    out.append((n, s, cur_file, int(r[0]), r[1].strip()[:100])); total += n; samples += s
out.sort(reverse=True)
print(f"total {total / warps:.1f} instructions per warp, {samples} samples")
for n, s, f, ln, src in out[:top]:
    print(f"{n / warps:8.1f} {100.0 * n / total:5.1f}%  smp {100.0 * s / max(samples, 1):5.1f}%  {f}:{ln}  {src}")
