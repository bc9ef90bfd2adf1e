#!/usr/bin/env python3
"""Race probe at bench.py's launch shape (N = 2^15, 512 diagonals, 2 blocks): the same query through membership -> index ->
membership back to back, many times, under every combination of the engine's overlap options; every result is compared with the
result of the most conservative configuration (one stream, no rotation lane, no PDL, unfused ModDown launch sequence).
Prints, per configuration, how many calls differed and where the first differing words sit.  GPU only; no oracle needed."""
import itertools
import sys
import os
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import image_matching_b200 as hx


def main():
    repeats = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    ndiag, blocks = 512, 2
    ring = int(os.environ.get("PROBE_RING", "0"))   # 0 = the reference's ring (2^15); 4096 exercises the small-ring kernels
    cc = hx.CryptoContext.generate(depth=11, ring_dim=ring)
    cl = hx.Client(cc, seed=3)
    cl.EvalMultKeyGen(); cl.EvalSumKeyGen(); cl.EvalRotateKeyGen(list(range(1, ndiag)))
    rng = np.random.default_rng(5)
    rows = [rng.uniform(-0.05, 0.05, cc.GetBatchSize()) for _ in range(16)]
    for b in range(blocks):
        for d in range(ndiag):
            cl.enroll_row(b, d, rows[(d * 7 + b) % 16])
    q0 = cl.encrypt(rng.uniform(-0.05, 0.05, cc.GetBatchSize()))
    q = cc.upload(q0.to_numpy(), 0, 1)   # an UPLOADED query is 'settled': only then does the engine use the rotation lane

    def setopt(lane, fuse, streams, pdl):
        cc.set_option("rot_lane", lane); cc.set_option("ks_moddown_w", fuse); cc.set_option("rot_streams", streams); cc.set_option("pdl", pdl)

    setopt(0, 0, 1, 0)   # the most conservative configuration: r1 ModDown launch sequence, one stream, no lane, no PDL
    ref_s = [c.to_numpy() for c in cc.score_diag(q, ndiag)]
    ref_i = [c.to_numpy() for c in cc.index(q, ndiag)]
    ref_m = cc.membership(q, ndiag).to_numpy()
    for _ in range(3):
        assert all(np.array_equal(a, c.to_numpy()) for a, c in zip(ref_s, cc.score_diag(q, ndiag))), "conservative configuration is not deterministic"

    def where(a, b):
        d = np.argwhere(a != b)
        if not len(d):
            return "equal"
        comps = sorted(set(map(tuple, d[:, :2].tolist())))
        pos = d[:, 2]
        return f"{len(d)} words differ; (comp,limb) {comps[:6]}{'...' if len(comps) > 6 else ''}; positions {pos.min()}..{pos.max()}, tiles {sorted(set((pos >> 11).tolist()))[:8]}"

    pools = [int(x) for x in os.environ.get("PROBE_POOL", "7").split(",")]
    lanes = [int(x) for x in os.environ.get("PROBE_LANE", "0,1").split(",")]
    only = os.environ.get("PROBE_ONLY")   # "moddown_w,streams,pdl" to run a single configuration
    for pool, lane, fuse, streams, pdl in itertools.product(pools, lanes, (0, 1), (1, 2), (0, 1)):
        if only and only != f"{fuse},{streams},{pdl}":
            continue
        cc.set_option("pool_reuse", pool)
        setopt(lane, fuse, streams, pdl)
        bad_s = bad_chain = 0
        first = None
        details = []
        for it in range(repeats):
            s = cc.score_diag(q, ndiag)
            mode = os.environ.get("PROBE_MODE", "")
            m1, i1 = cc.membership(q, ndiag), cc.index(q, ndiag)
            early = clones = None
            if mode == "early":      # download the index results before the next scenario is issued
                early = [c.to_numpy() for c in i1]
            elif mode == "clone":    # device-side copies taken right after the index scenario, compared at the end
                clones = [cc._ct(cc._lib.hx_ct_clone(c._h)) for c in i1]
            m2 = cc.membership(q, ndiag)
            if early is not None:
                for g, (a, e) in enumerate(zip(ref_i, early)):
                    if not np.array_equal(a, e) and len(details) < 6:
                        details.append(f"it {it} EARLY download of i1[{g}] already wrong: " + where(a, e))
            if clones is not None:
                for g, (a, c, o) in enumerate(zip(ref_i, clones, i1)):
                    cn, on = c.to_numpy(), o.to_numpy()
                    if (not np.array_equal(a, cn) or not np.array_equal(a, on)) and len(details) < 6:
                        details.append(f"it {it} i1[{g}]: clone {'ok' if np.array_equal(a, cn) else 'WRONG'}, original {'ok' if np.array_equal(a, on) else 'WRONG'}")
            for g, (a, c) in enumerate(zip(ref_s, s)):
                x = c.to_numpy()
                if not np.array_equal(a, x):
                    bad_s += 1
                    first = first or f"score block {g}: " + where(a, x)
            ok = True
            for name, got, want in [("m1", m1.to_numpy(), ref_m), ("m2", m2.to_numpy(), ref_m)] + [(f"i1[{g}]", c.to_numpy(), a) for g, (a, c) in enumerate(zip(ref_i, i1))]:
                if not np.array_equal(got, want):
                    ok = False
                    if len(details) < 6:
                        details.append(f"it {it} {name}: " + where(want, got))
            bad_chain += 0 if ok else 1
        print(f"pool={pool} lane={lane} moddown_w={fuse} streams={streams} pdl={pdl}: score mismatches {bad_s}/{repeats * blocks}, scenario-chain mismatches {bad_chain}/{repeats}"
              + (f"  first: {first}" if first else ""), flush=True)
        for d in details:
            print("      " + d, flush=True)


if __name__ == "__main__":
    main()
