#!/usr/bin/env python3
"""Static SASS opcode histogram per kernel of an object / shared library (cuobjdump -sass).

The key-switch kernels are fully unrolled and bound by instruction issue, so the static
count of a kernel's straight-line body tracks its dynamic instruction count: this is the offline proxy used
between GPU runs.   usage: sass_hist.py <file.o|.so> <kernel-name-regex> [--top N]
"""
import collections
import re
import subprocess
import sys


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def main():
    path, pat = sys.argv[1], re.compile(sys.argv[2])
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 25
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    funcs = collections.OrderedDict()
    cur = None
    for line in txt.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", line)
        if m and cur is not None:
            ins = m.group(1).strip()
            ins = re.sub(r"^@!?U?P\d+\s+", "", ins)
            funcs[cur].append(ins.split()[0])
    dm = demangle(list(funcs))
    for f, ops in funcs.items():
        name = dm.get(f, f)
        if not pat.search(name):
            continue
        h = collections.Counter()
        for op in ops:
            base = op.split(".")[0]
            if base in ("IMAD", "LDG", "STG", "LDS", "STS", "ISETP", "SEL", "IADD3", "LEA", "SHF", "LOP3", "DFMA", "DADD", "DMUL"):
                base = ".".join(op.split(".")[:2]) if base in ("IMAD", "LDG", "STG", "LDS", "STS") else base
            h[base] += 1
        tot = sum(h.values())
        print(f"== {name[:110]}  total {tot}")
        print("   " + "  ".join(f"{k}:{v}" for k, v in h.most_common(top)))


if __name__ == "__main__":
    main()
