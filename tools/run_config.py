"""Runs BASELINE.json configs[2] / configs[3] (approaches 1, 3, 4) at the reference's full parameters on one B200 with the
product's own client (synthetic DB as tools/gen_dataset.sh: rows in [-99, 99], planted matches, query all ones), times the
Sender's membership + index scenarios on the device and checks the Receiver's decisions.  One JSON line per run.
  python tools/run_config.py --approach 4 --log2n 16
"""
import argparse
import json
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def synth_db(n, dim, seed, match_rows):
    rng = np.random.default_rng(seed)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    for r in match_rows:
        db[r] = rng.integers(1, 4, size=dim)
    return db


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--approach", type=int, required=True, choices=[1, 3, 4])
    ap.add_argument("--log2n", type=int, default=14)
    ap.add_argument("--ring-dim", type=int, default=0)
    ap.add_argument("--steps", type=int, default=1)
    ap.add_argument("--profile", action="store_true", help="extra pass with per-kernel-family device times and phase wall clocks")
    a = ap.parse_args()
    import image_matching_b200 as hx
    from image_matching_b200 import OpenFHEWrapper
    from image_matching_b200.config import CHUNK_LEN, VECTOR_DIM

    n, dim = 1 << a.log2n, VECTOR_DIM
    depth = OpenFHEWrapper.computeRequiredDepth(a.approach)
    cc = hx.CryptoContext.generate(depth=depth, ring_dim=a.ring_dim)
    batch = cc.GetBatchSize()
    cl = hx.Client(cc, seed=1)
    t0 = time.perf_counter()
    cl.EvalMultKeyGen()
    cl.EvalRotateKeyGen([s * (1 << k) for k in range(int(math.log2(batch))) for s in (1, -1)])   # binaryRotate / EvalSum set
    matches = [0, n - 1]
    db = synth_db(n, dim, 100, matches)
    q = np.ones(dim)
    if a.approach == 4:
        nmat = cl.enroll_hers(db)
        snd = hx.HersSender(cc, None, n)
        qc = cl.encryptQueryHers(q)
        decode = cl.decryptIndex
        work = f"{nmat} matrices x {dim} database ciphertexts, query = {dim} ciphertexts"
    elif a.approach == 1:
        rows = hx.base_pack(db, batch)
        dbc = [cl.encrypt(r) for r in rows]
        snd = hx.BaseSender(cc, None, n, db_loader=lambda i: dbc[i])
        qc = cl.encryptQueryBase(q)
        decode = cl.decryptIndex
        work = f"{len(dbc)} database ciphertexts ({batch // dim} vectors each)"
    else:
        pk = hx.blind_pack(db, batch, CHUNK_LEN)
        dbc = [[cl.encrypt(pk[m, c]) for c in range(pk.shape[1])] for m in range(pk.shape[0])]
        snd = hx.BlindSender(cc, None, n, db_loader=lambda m, j: dbc[m][j // CHUNK_LEN])
        qc = cl.encryptQueryBlind(q, CHUNK_LEN)
        decode = lambda cts: sorted(cl.decryptIndexBlind(cts, CHUNK_LEN))
        work = f"{len(dbc)} matrices x {pk.shape[1]} database ciphertexts"
    setup_s = time.perf_counter() - t0
    mem = snd.membershipScenario(qc)          # warm-up (allocator pool, masks, tables)
    cc.sync()
    l0 = cc.launches()
    cc.timer_start()
    w0 = time.perf_counter()
    for _ in range(a.steps):
        mem = snd.membershipScenario(qc)
        idx = snd.indexScenario(qc)
    cc.sync()
    ms = cc.timer_stop() / a.steps
    wall = (time.perf_counter() - w0) / a.steps
    if a.profile:
        cc.prof_reset(); cc.prof_enable(None)
        p0 = time.perf_counter(); sc = snd.computeSimilarity(qc); cc.sync(); p1 = time.perf_counter()
        from image_matching_b200.config import MATCH_THRESHOLD, COMP_DEPTH
        [OpenFHEWrapper.chebyshevCompare(cc, s, MATCH_THRESHOLD, COMP_DEPTH) for s in sc]; cc.sync(); p2 = time.perf_counter()
        fam = {k: (v["launches"], round(v["ms"], 2)) for k, v in cc.prof_read().items() if v["launches"]}
        cc.prof_enable([])
        print(json.dumps({"similarity_s": p1 - p0, "compare_s": p2 - p1, "outputs": len(sc), "families": fam}), flush=True)
    got = decode(idx)
    ok = (got == matches) and cl.decryptMembership(mem)
    print(json.dumps({"approach": a.approach, "n": n, "ring_dim": cc.N, "limbs": f"{cc.L}+{cc.K}", "depth": depth, "workload": work,
                      "seconds_membership_plus_index_device": ms / 1e3, "seconds_wall": wall, "gpu_launches_per_step": (cc.launches() - l0) // a.steps,
                      "setup_seconds": setup_s, "index": got[:8], "results_ok": bool(ok)}), flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
