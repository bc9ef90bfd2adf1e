"""Runs the reference's own executable, built unmodified against shim/openfhe.h (shim/_ref/ImageMatching, see shim/Makefile),
on a synthetic dataset file in the reference's text format (main.cpp:56-57, 216-230; generator = seeded restatement of
tools/gen_dataset.sh: rows in [-99, 99], matches in {1,2,3} at rows 0 and n-1, query all ones) and reports the latencies the
reference's own timers print (main.cpp:333-372), with every cc-> call synchronous (HYDIA_SYNC_EVERY_CALL=1).
  python tools/run_reference_shim.py --approach 5 --log2n 14 [--no-deferred]
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "shim", "_ref", "ImageMatching")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--approach", type=int, default=5)
    ap.add_argument("--log2n", type=int, default=14)
    ap.add_argument("--no-deferred", action="store_true")
    ap.add_argument("--chains", action="store_true",
                    help="HYDIA_DEFERRED=2: the compare chains of all blocks are recorded and run as one batch; the index outputs then "
                         "complete when the receiver reads them, so the latency to report is index + decrypt")
    ap.add_argument("--async-calls", action="store_true", help="leave HYDIA_SYNC_EVERY_CALL unset")
    a = ap.parse_args()
    n = 1 << a.log2n
    rng = np.random.default_rng(0)
    db = rng.integers(-99, 100, size=(n, 512))
    matches = [0, n - 1]
    for r in matches:
        db[r] = rng.integers(1, 4, size=512)
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "data.dat")
        with open(path, "w") as f:
            f.write(f"{n}\n" + " ".join(["1"] * 512) + "\n")
            np.savetxt(f, db, fmt="%d")
        for d in ("serial/db_diagonal", "serial/db_hers", "serial/db_blind", "serial/db_base"):
            os.makedirs(os.path.join(tmp, d))
        env = dict(os.environ)
        if not a.async_calls:
            env["HYDIA_SYNC_EVERY_CALL"] = "1"
        if a.no_deferred:
            env["HYDIA_DEFERRED"] = "0"
        elif a.chains:
            env["HYDIA_DEFERRED"] = "2"
        t0 = time.perf_counter()
        r = subprocess.run([EXE, path, str(a.approach)], cwd=tmp, capture_output=True, text=True, env=env)
        wall = time.perf_counter() - t0
    out = r.stdout
    t = dict(re.findall(r"Computing (\w+) scenario\.\.\. done \(([0-9.e+-]+)s\)", out))
    dec = dict(re.findall(r"Decrypting (\w+) results\.\.\. done \(([0-9.e+-]+)s\)", out))
    mem = re.search(r"Membership scenario:\s*(\w+)", out)
    idx = re.search(r"Index scenario:\s*\[([^\]]*)\]", out)
    got = [int(x) for x in idx.group(1).split()] if idx else None
    ok = r.returncode == 0 and mem and mem.group(1) == "true" and got == matches
    print(json.dumps({"executable": "reference src/main.cpp, unmodified, on shim/openfhe.h + libhydia_b200.so", "approach": a.approach, "n": n,
                      "deferred_layer": not a.no_deferred, "synchronous_calls": not a.async_calls,
                      "reference_timer_membership_s": float(t.get("membership", "nan")), "reference_timer_index_s": float(t.get("index", "nan")),
                      "reference_timer_index_plus_decrypt_s": float(t.get("index", "nan")) + float(dec.get("index", "nan")),
                      "batched_compare_chains": bool(a.chains),
                      "note": "membership runs first and includes one-time warm-up (allocator pool, tables)",
                      "membership": mem.group(1) if mem else None, "index": got, "results_ok": bool(ok), "process_wall_s": round(wall, 1)}), flush=True)
    if not ok:
        sys.stderr.write(out[-1500:] + r.stderr[-1500:])
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
