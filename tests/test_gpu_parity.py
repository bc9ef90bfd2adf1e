"""GPU parity tests: every CUDA op through the C ABI against the CPU oracle on the same keys and
inputs, BIT-EXACT (integer residues), plus decrypted-level checks with stated tolerances."""
import os

import numpy as np
import pytest

from oracle import ckks_oracle as O
from tests.helpers import make_gpu_context, make_oracle, same, up

pytestmark = pytest.mark.gpu

ROTS = [1, 2, 3, 7, 100, 511, -1, -2, -64, 512, 1024]


@pytest.fixture(scope="module")
def env():
    orc = make_oracle(ring_dim=4096, depth=11, rotations=ROTS + [2 ** i for i in range(11)])
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(7)
    a, b = rng.uniform(-1, 1, orc.P.slots), rng.uniform(-1, 1, orc.P.slots)
    ca, cb = orc.encrypt(a), orc.encrypt(b)
    return dict(orc=orc, cc=cc, a=a, b=b, ca=ca, cb=cb, ga=up(cc, ca), gb=up(cc, cb))


@pytest.mark.parametrize("logn", [10, 11, 12, 13, 14, 15, 16])
def test_ntt_bit_exact(logn):
    """Row P2: forward and inverse NTT of random limbs over Q and P primes, every ring size the
    approaches use (2^15, 2^16) and the small test rings; also iNTT(NTT(x)) == x."""
    N = 1 << logn
    P = O.gen_params(3, ring_dim=N)
    orc = O.Oracle(P)
    import image_matching_b200 as hx
    cc = hx.CryptoContext(P.N, P.q, P.p, P.psi_q, P.psi_p, P.dnum)
    rng = np.random.default_rng(logn)
    primes = list(range(P.L + P.K))
    x = np.stack([rng.integers(0, orc.mods[i], size=N, dtype=np.uint64) for i in primes])
    # edge values: 0, 1, q-1
    x[:, 0] = 0; x[:, 1] = 1; x[:, 2] = np.array([m - 1 for m in orc.mods], dtype=np.uint64)
    want = x.copy(); orc.ntt(want, primes)
    got = cc.ntt(x, primes)
    assert np.array_equal(got, want)
    back = cc.ntt(got, primes, inverse=True)
    assert np.array_equal(back, x)
    wanti = x.copy(); orc.ntt(wanti, primes, inverse=True)
    assert np.array_equal(cc.ntt(x, primes, inverse=True), wanti)
    # batch with a repeating prime pattern longer than one period
    xb = np.concatenate([x, x])
    wb = np.concatenate([want, want])
    assert np.array_equal(cc.ntt(xb, primes), wb)


def test_elementwise_ops_bit_exact(env):
    orc, cc, ca, cb, ga, gb = (env[k] for k in ("orc", "cc", "ca", "cb", "ga", "gb"))
    assert same(cc.EvalAdd(ga, gb), orc.add(ca, cb))
    assert same(cc.EvalSub(ga, gb), orc.sub(ca, cb))
    for c in (0.25, -0.37, 1.0, 2.4609375, -3.28125, 1e-9, 0.0):
        assert same(cc.EvalMult(ga, c), orc.mult_const(ca, c))
        assert same(cc.EvalAdd(ga, c), orc.add_const(ca, c))
    # constants on a degree-2 ciphertext use scale^2 (GetElementForEvalAddOrSub)
    m2, g2 = orc.mult_const(ca, 0.5), cc.EvalMult(ga, 0.5)
    assert same(cc.EvalAdd(g2, -1.0), orc.add_const(m2, -1.0))
    assert same(cc.LevelReduce(ga, 3), orc.level_reduce(ca, 3))
    # level alignment by dropping limbs (FIXEDMANUAL)
    assert same(cc.EvalAdd(cc.LevelReduce(ga, 2), gb), orc.add(orc.level_reduce(ca, 2), cb))
    assert same(cc.EvalAddMany([ga, gb, ga, gb, ga]), orc.add_many([ca, cb, ca, cb, ca]))


def test_tensor_relin_rescale_bit_exact(env):
    orc, cc, ca, cb, ga, gb = (env[k] for k in ("orc", "cc", "ca", "cb", "ga", "gb"))
    t, gt = orc.mult_norelin(ca, cb), cc.EvalMultNoRelin(ga, gb)
    assert same(gt, t)                                        # row P6
    assert same(cc.EvalAdd(gt, gt), orc.add(t, t))            # 3-component add
    r, gr = orc.relinearize(t), cc.Relinearize(gt)
    assert same(gr, r)                                        # rows P4 + P5 (hybrid key switch, l = 12)
    s, gs = orc.rescale(r), cc.Rescale(gr)
    assert same(gs, s)                                        # row P7
    assert same(cc.Rescale(gt), orc.rescale(t))               # rescale of a 3-component ciphertext
    assert same(cc.EvalMult(ga, gb), r) and same(cc.EvalSquare(ga), orc.square(ca))
    assert same(cc.Relinearize(ga), ca)                       # no-op on 2 components
    dec = orc.decrypt(O.Ct(list(gs.to_numpy()), 1, 1))
    assert np.abs(dec - env["a"] * env["b"]).max() <= 1e-6


@pytest.mark.parametrize("drop", [0, 1, 3, 4, 7, 8, 10, 11])
def test_key_switch_every_level_bit_exact(env, drop):
    """Hybrid key switch at every digit shape: l = 12..1 limbs -> 3, 2 and 1 digits, ragged last digit."""
    orc, cc, ca, cb, ga, gb = (env[k] for k in ("orc", "cc", "ca", "cb", "ga", "gb"))
    la, lb = orc.level_reduce(ca, drop), orc.level_reduce(cb, drop)
    xa, xb = cc.LevelReduce(ga, drop), cc.LevelReduce(gb, drop)
    assert same(cc.EvalMult(xa, xb), orc.mult(la, lb))
    assert same(cc.EvalRotate(xa, 7), orc.rotate(la, 7))
    if la.nl >= 2:
        assert same(cc.Rescale(cc.EvalMult(xa, xb)), orc.rescale(orc.mult(la, lb)))


def test_rotations_bit_exact(env):
    orc, cc, ca, ga = env["orc"], env["cc"], env["ca"], env["ga"]
    for r in ROTS:
        assert same(cc.EvalRotate(ga, r), orc.rotate(ca, r)), r            # rows P3 + P5
    dg, gdg = orc.fast_rotation_precompute(ca), cc.EvalFastRotationPrecompute(ga)
    for r in (0, 1, 3, 511, -64):
        assert same(cc.EvalFastRotation(ga, r, 2 * orc.N, gdg), orc.fast_rotation(ca, r, dg)), r
    low, glow = orc.level_reduce(ca, 11), cc.LevelReduce(ga, 11)
    assert same(cc.EvalSum(glow, orc.P.slots), orc.eval_sum(low, orc.P.slots))
    cb, gb = env["cb"], env["gb"]
    both = cc.stack([ga, gb])                                              # batched ciphertext: same op on every entry
    r0, r1 = cc.unstack(cc.EvalRotate(both, 3))
    assert same(r0, orc.rotate(ca, 3)) and same(r1, orc.rotate(cb, 3))
    m0, m1 = cc.unstack(cc.Rescale(cc.EvalMult(ga, both)))                 # single ciphertext broadcast over the batch
    assert same(m0, orc.rescale(orc.mult(ca, ca))) and same(m1, orc.rescale(orc.mult(ca, cb)))
    from image_matching_b200._lib import HxError
    with pytest.raises(HxError):
        both.to_numpy()                                                    # entries are downloaded one by one
    with pytest.raises(HxError, match="not generated"):                    # OpenFHE throws on a missing key
        cc.EvalRotate(ga, 5)
    with pytest.raises(HxError):
        cc.Rescale(cc.LevelReduce(ga, 11))                                 # no limb left to drop


def test_chebyshev_compare_bit_exact(env):
    """Rows R4 + P9: the whole chebyshevCompare op DAG (Chebyshev-59 PS, f4 PS, +1) bit for bit,
    through both the fused hx_compare and the reference-style per-call sequence."""
    orc, cc = env["orc"], env["cc"]
    x = np.random.default_rng(3).uniform(-0.3, 0.3, orc.P.slots)
    x[:6] = [0.93, 0.47, 0.4542, 0.4540, 0.43, -0.9]
    ct = orc.level_reduce(orc.encrypt(x), 1)
    g = up(cc, ct)
    want = orc.chebyshev_compare(ct, 0.44, 10)
    got = cc.compare(g, 0.44, 10)
    assert same(got, want)
    import image_matching_b200 as hx
    assert same(hx.OpenFHEWrapper.chebyshevCompare(cc, g, 0.44, 10), want)
    dec = orc.decrypt(want)
    ref = O.chebyshev_compare(O.PlainEval(), O.PlainCt(x)).v
    assert np.abs(dec - ref).max() <= 1e-4
    assert list(dec[:6] >= 1.0) == [True, True, True, False, False, False]


def test_hydia_fused_path_bit_exact_small_ring(golden_dir):
    """Rows R1-R6: fused score_diag / index / membership vs the oracle's per-op restatement of
    sender_diag.cpp on a ragged database (1300 vectors -> 3 square blocks in one 2048-slot cipher
    + a second, mostly empty cipher block), and vs the reference-style per-op DiagonalSender."""
    import image_matching_b200 as hx
    rots = list(range(1, 512)) + [2 ** i for i in range(11)]
    orc = make_oracle(ring_dim=4096, depth=11, rotations=rots, seed=5)
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(11)
    n = 2300
    db = rng.integers(-99, 100, size=(n, 512)).astype(np.float64)
    q = np.ones(512)
    db[0] = rng.integers(1, 4, size=512); db[n - 1] = rng.integers(1, 4, size=512)
    rows = O.diagonal_pack(db, orc.P.slots)
    assert rows.shape[0] == 1024
    dbc = [orc.encrypt(r) for r in rows]
    for i, c in enumerate(dbc):
        cc.db_register(i // 512, i % 512, np.stack(c.c))
    qc = O.encrypt_query_diag(orc, q)
    gq = up(cc, qc)
    want = O.diag_compute_similarity(orc, qc, dbc)
    got = cc.score_diag(gq)
    assert len(got) == 2 and all(same(g, w) for g, w in zip(got, want))
    widx = [O.chebyshev_compare(orc, s) for s in want]
    gidx = cc.index(gq)
    assert all(same(g, w) for g, w in zip(gidx, widx))
    wm = orc.eval_sum(orc.add_many(widx), orc.P.slots)
    assert same(cc.membership(gq), wm)
    # the two blocks above went through ONE batched relinearise / rescale / compare chain; per-block chain must agree
    cc.set_option("batch_blocks", 0)
    assert all(same(g, w) for g, w in zip(cc.index(gq), widx))
    cc.set_option("batch_blocks", 1)
    # one stream vs two for the 32-rotation batches: same polynomials
    for ns in (1, 2):
        cc.set_option("rot_streams", ns)
        assert all(same(g, w) for g, w in zip(cc.score_diag(gq), want))
    # rotation lane (the uploaded query's rotations issued on low-priority side streams, so a previous call's compare chain
    # overlaps them) vs everything ordered on the engine stream: back-to-back scenarios, same polynomials
    for lane in (0, 1):
        cc.set_option("rot_lane", lane)
        m1, i1, m2 = cc.membership(gq), cc.index(gq), cc.membership(gq)
        assert same(m1, wm) and same(m2, wm) and all(same(g, w) for g, w in zip(i1, widx))
    # decisions (receiver_hers.cpp:26-54) and score accuracy (main_accuracy.cpp:360)
    assert O.decrypt_index(orc, widx) == [0, n - 1] and O.decrypt_membership(orc, wm)
    cos = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    s = np.concatenate([orc.decrypt(O.Ct(list(g.to_numpy()), 1, 1)) for g in got])
    assert np.abs(s[:n] - cos).max() <= 1e-4
    # reference-style call sequence (511 EvalFastRotation, 512 EvalMultNoRelin, 511 EvalAdd ...), first block only
    gdb = [up(cc, c) for c in dbc[:512]]
    snd = hx.DiagonalSender(cc, None, 2048, fused=False, db_loader=lambda i: gdb[i])
    per_op = snd.computeSimilarity([gq])
    assert len(per_op) == 1 and same(per_op[0], want[0])
    # the same path in two steps (what shim/openfhe.h's deferred layer calls): database ciphertexts handed over as device
    # handles, rotations once, blocks scored against the rotation set in any order; stored ciphertexts read back unchanged
    cc.db_register_ct(2, 7, gdb[7])
    assert np.array_equal(cc.db_get(2, 7).to_numpy(), gdb[7].to_numpy()) and same(cc.db_get(1, 3), dbc[512 + 3])
    scores, rotated = cc.score_blocks(gq, 512, [1, 0], rotations=[0, 1, 300])
    assert same(scores[0], want[1]) and same(scores[1], want[0])
    assert all(same(r, orc.rotate(qc, k)) for r, k in zip(rotated, (0, 1, 300)))


def test_full_ring_ops_bit_exact():
    """N = 2^15, 12 + 4 limbs (the configuration of configs[0..1]): key switch, rescale, hoisted
    rotation and tensor product bit-exact vs the oracle at full size."""
    orc = make_oracle(ring_dim=None, depth=11, rotations=[1, 300, -8])
    assert orc.N == 32768
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(9)
    a, b = rng.uniform(-1, 1, orc.P.slots), rng.uniform(-1, 1, orc.P.slots)
    ca, cb = orc.encrypt(a), orc.encrypt(b)
    ga, gb = up(cc, ca), up(cc, cb)
    t = orc.mult_norelin(ca, cb)
    r = orc.relinearize(t)
    gr = cc.Relinearize(cc.EvalMultNoRelin(ga, gb))
    assert same(gr, r)
    assert same(cc.Rescale(gr), orc.rescale(r))
    dg, gdg = orc.fast_rotation_precompute(ca), cc.EvalFastRotationPrecompute(ga)
    for rot in (1, 300):
        assert same(cc.EvalFastRotation(ga, rot, 2 * orc.N, gdg), orc.fast_rotation(ca, rot, dg))
    assert same(cc.EvalRotate(ga, -8), orc.rotate(ca, -8))


@pytest.mark.parametrize("ring", [4096, 8192, 16384, 65536])
def test_fused_pipeline_equals_unfused_kernels(ring):
    """The fused key-switch / ModUp / rescale pipeline (hx_fused.cu) against the unfused one-op-per-kernel path
    (hx_kernels.cu) on the product's own keys, at every tile geometry (N = 2^12 .. 2^16), all digit shapes."""
    import image_matching_b200 as hx
    cc = hx.CryptoContext.generate(depth=11, ring_dim=ring)
    cl = hx.Client(cc, seed=ring)
    cl.EvalMultKeyGen(); cl.EvalRotateKeyGen([1, -3, 37])
    rng = np.random.default_rng(ring)
    ca, cb = cl.encrypt(rng.uniform(-1, 1, ring // 2)), cl.encrypt(rng.uniform(-1, 1, ring // 2))

    def run():
        outs = []
        for drop in (0, 3, 7, 8, 10, 11):
            xa, xb = cc.LevelReduce(ca, drop), cc.LevelReduce(cb, drop)
            m = cc.EvalMult(xa, xb)
            outs += [m, cc.EvalRotate(xa, 37), cc.EvalRotate(m, -3)]
            if drop < 11:
                outs += [cc.Rescale(m), cc.Rescale(cc.EvalMultNoRelin(xa, xb))]
        dg = cc.EvalFastRotationPrecompute(ca)
        outs += [cc.EvalFastRotation(ca, 1, 2 * ring, dg), cc.EvalFastRotation(ca, -3, 2 * ring, dg)]
        return [o.to_numpy() for o in outs]

    fused = run()
    cc.set_option("force_unfused", 1)
    plain = run()
    cc.set_option("force_unfused", 0)
    assert len(fused) == len(plain)
    for f, p in zip(fused, plain):
        assert np.array_equal(f, p)


def test_hers_sender_bit_exact_small_ring():
    """Row R7: HersSender (approach 4) -- fused score_hers + batched compare vs the oracle's restatement of
    sender_hers.cpp, on a ragged database spanning two matrices (2048 slots each), 40 coordinates so the
    32-wide key-switch batches see a full and a partial chunk; and vs the reference-style per-op call sequence."""
    import image_matching_b200 as hx
    dim, n = 40, 2500
    orc = make_oracle(ring_dim=4096, depth=11, rotations=[2 ** i for i in range(11)], seed=7)
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(21)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    q = rng.integers(1, 4, size=dim).astype(np.float64)
    c0 = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    db[c0 > 0.3] *= -1.0                      # keep every unplanted cosine well below the 0.44 threshold
    db[3] = q; db[n - 1] = 2 * q
    packed = O.hers_pack(db, orc.P.slots)
    assert packed.shape == (2, dim, orc.P.slots)
    assert np.array_equal(packed, hx.hers_pack(db, orc.P.slots))
    dbc = [[orc.encrypt(packed[m, j]) for j in range(dim)] for m in range(2)]
    for m in range(2):
        for j in range(dim):
            cc.db_register(m, j, np.stack(dbc[m][j].c))
    qc = O.encrypt_query_hers(orc, q)
    gq = [up(cc, c) for c in qc]
    want = O.hers_compute_similarity(orc, qc, dbc)
    snd = hx.HersSender(cc, None, n)
    got = snd.computeSimilarity(gq)
    assert len(got) == 2 and all(same(g, w) for g, w in zip(got, want))
    widx = [O.chebyshev_compare(orc, s) for s in want]
    gidx = snd.indexScenario(gq)
    assert all(same(g, w) for g, w in zip(gidx, widx))
    wm = orc.eval_sum(orc.add_many(widx), orc.P.slots)
    assert same(snd.membershipScenario(gq), wm)
    assert O.decrypt_index(orc, widx) == [3, n - 1] and O.decrypt_membership(orc, wm)
    cos = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    s = np.concatenate([orc.decrypt(w) for w in want])
    assert np.abs(s[:n] - cos).max() <= 1e-4
    # reference-style call sequence on the first matrix
    gdb = [up(cc, c) for c in dbc[0]]
    per_op = hx.HersSender(cc, None, orc.P.slots, fused=False, db_loader=lambda m, i: gdb[i]).computeSimilarity(gq)
    assert len(per_op) == 1 and same(per_op[0], want[0])


def _pm_pow2(slots):
    return [s * (1 << k) for k in range(int(np.log2(slots))) for s in (1, -1)]


def test_base_sender_bit_exact_small_ring():
    """Row R9 (+R10): BaseSender (approach 1) through the per-op C ABI -- EvalInnerProduct, EvalMult(ct, Plaintext) masks,
    binaryRotate, mergeCiphers -- bit-exact vs the oracle's restatement, then compare + decisions."""
    import image_matching_b200 as hx
    dim, n = 64, 80
    orc = make_oracle(ring_dim=4096, depth=13, rotations=_pm_pow2(2048), seed=5)
    cc = make_gpu_context(orc)
    cc.encoder = orc.encode                    # both sides multiply by the same encoded masks
    rng = np.random.default_rng(6)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    q = rng.integers(1, 4, size=dim).astype(np.float64)
    c0 = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    db[c0 > 0.3] *= -1.0
    db[41] = 2 * q
    packed = O.base_pack(db, orc.P.slots)
    assert np.array_equal(packed, hx.base_pack(db, orc.P.slots))
    dbc = [orc.encrypt(r) for r in packed]
    gdb = [up(cc, c) for c in dbc]
    qc = O.encrypt_query_base(orc, q)
    gq = [up(cc, qc[0])]
    want = O.base_compute_similarity(orc, qc, dbc, dim)
    snd = hx.BaseSender(cc, None, n, db_loader=lambda i: gdb[i], dim=dim)
    got = snd.computeSimilarity(gq)
    assert len(got) == len(want) == 1 and same(got[0], want[0])
    per_op = hx.BaseSender(cc, None, n, db_loader=lambda i: gdb[i], dim=dim, batched=False).computeSimilarity(gq)
    assert same(per_op[0], want[0])                 # one ciphertext at a time, as the reference loops
    widx = [O.chebyshev_compare(orc, w) for w in want]
    gidx = snd.indexScenario(gq)
    assert all(same(g, w) for g, w in zip(gidx, widx))
    assert O.decrypt_index(orc, widx) == [41]
    assert same(snd.membershipScenario(gq), orc.eval_sum(orc.add_many(widx), orc.P.slots))


def test_blind_sender_bit_exact_small_ring():
    """Row R8: BlindSender (approach 3) -- chunked EvalMultNoRelin, rotate-and-sum, compressCiphers -- bit-exact vs the oracle."""
    import image_matching_b200 as hx
    dim, chunk, n = 64, 16, 200
    orc = make_oracle(ring_dim=4096, depth=12, rotations=_pm_pow2(2048), seed=8)
    cc = make_gpu_context(orc)
    cc.encoder = orc.encode
    rng = np.random.default_rng(9)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    q = rng.integers(1, 4, size=dim).astype(np.float64)
    c0 = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    db[c0 > 0.3] *= -1.0
    db[5] = q; db[150] = 3 * q
    packed = O.blind_pack(db, orc.P.slots, chunk)
    assert np.array_equal(packed, hx.blind_pack(db, orc.P.slots, chunk))
    dbc = [[orc.encrypt(packed[m, c]) for c in range(packed.shape[1])] for m in range(packed.shape[0])]
    gdb = [[up(cc, c) for c in row] for row in dbc]
    qc = O.encrypt_query_blind(orc, q, chunk)
    gq = [up(cc, c) for c in qc]
    want = O.blind_compute_similarity(orc, qc, dbc, chunk)
    snd = hx.BlindSender(cc, None, n, db_loader=lambda m, j: gdb[m][j // chunk], dim=dim, chunk=chunk)
    got = snd.computeSimilarity(gq)
    assert len(got) == len(want) == 1 and same(got[0], want[0])
    widx = [O.chebyshev_compare(orc, w) for w in want]
    assert all(same(g, w) for g, w in zip(snd.indexScenario(gq), widx))
    per_op = hx.BlindSender(cc, None, n, db_loader=lambda m, j: gdb[m][j // chunk], dim=dim, chunk=chunk, batched=False).computeSimilarity(gq)
    assert same(per_op[0], want[0])
    assert sorted(O.blind_decrypt_index(orc, widx, chunk)) == [5, 150]


def test_client_masks_and_roles_for_approaches_1_and_3():
    """The product's own client (hx_client_encode = MakeCKKSPackedPlaintext, encryptQueryBase/Blind, decryptIndexBlind):
    BaseSender and BlindSender end to end on the product's keys, decisions checked against the planted matches."""
    import image_matching_b200 as hx
    dim, chunk, n = 64, 16, 200
    cc = hx.CryptoContext.generate(depth=13, ring_dim=4096)
    cl = hx.Client(cc, seed=3)
    cl.EvalMultKeyGen(); cl.EvalRotateKeyGen(_pm_pow2(cc.GetBatchSize()))
    rng = np.random.default_rng(2)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    q = rng.integers(1, 4, size=dim).astype(np.float64)
    c0 = hx.plaintext_normalize(db) @ hx.plaintext_normalize(q)
    db[c0 > 0.3] *= -1.0
    db[7] = q; db[199] = 2 * q
    bdb = [cl.encrypt(r) for r in hx.base_pack(db, cc.GetBatchSize())]
    base = hx.BaseSender(cc, None, n, db_loader=lambda i: bdb[i], dim=dim)
    assert cl.decryptIndex(base.indexScenario(cl.encryptQueryBase(q))) == [7, 199]
    pk = hx.blind_pack(db, cc.GetBatchSize(), chunk)
    kdb = [[cl.encrypt(pk[m, c]) for c in range(pk.shape[1])] for m in range(pk.shape[0])]
    blind = hx.BlindSender(cc, None, n, db_loader=lambda m, j: kdb[m][j // chunk], dim=dim, chunk=chunk)
    assert sorted(cl.decryptIndexBlind(blind.indexScenario(cl.encryptQueryBlind(q, chunk)), chunk)) == [7, 199]
    assert cl.decryptMembership(blind.membershipScenario(cl.encryptQueryBlind(q, chunk)))


def test_hydia_fused_path_bit_exact_full_ring():
    """configs[1] geometry (N = 2^15, 12 + 4 limbs, the tile geometry and launch shapes bench.py runs): 63 hoisted rotations in
    two key-switch batches, the diagonal MAC, relinearise, rescale and the whole compare chain, bit-exact vs the oracle."""
    ndiag = 64
    orc = make_oracle(ring_dim=None, depth=11, rotations=list(range(1, ndiag)) + [2 ** i for i in range(14)], seed=9)
    assert orc.N == 32768 and orc.P.L == 12 and orc.P.K == 4
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(31)
    dbc = [orc.encrypt(rng.uniform(-0.1, 0.1, orc.P.slots)) for _ in range(ndiag)]
    for d, c in enumerate(dbc):
        cc.db_register(0, d, np.stack(c.c))
    qc = orc.encrypt(rng.uniform(-0.1, 0.1, orc.P.slots))
    gq = up(cc, qc)
    want = O.diag_compute_similarity(orc, qc, dbc, ndiag)
    got = cc.score_diag(gq, ndiag)
    assert len(got) == 1 and same(got[0], want[0])
    widx = [O.chebyshev_compare(orc, s) for s in want]
    assert same(cc.index(gq, ndiag)[0], widx[0])
    assert same(cc.membership(gq, ndiag), orc.eval_sum(orc.add_many(widx), orc.P.slots))


def test_error_behaviour_matches_openfhe_exceptions():
    """OpenFHE throws on a missing evaluation key, on an exhausted modulus chain and on mismatched noiseScaleDeg; the
    reference never catches. The C ABI reports the same situations as errors instead of computing garbage."""
    import image_matching_b200 as hx
    from image_matching_b200._lib import HxError
    orc = make_oracle(ring_dim=4096, depth=2, rotations=[1], seed=2)
    cc = make_gpu_context(orc)
    ct = up(cc, orc.encrypt(np.linspace(-1, 1, orc.P.slots)))
    with pytest.raises(HxError, match="Galois"):
        cc.EvalRotate(ct, 5)                                   # EvalAutomorphismKey not generated
    with pytest.raises(HxError):
        cc.score_diag(ct, 8)                                   # no database registered
    low = cc.Rescale(cc.Rescale(cc.EvalMult(cc.EvalMult(ct, 0.5), 0.5)))
    assert low.nlimbs == 1
    with pytest.raises(HxError, match="limb"):
        cc.Rescale(cc.EvalMult(low, 0.5))                      # level exhausted
    with pytest.raises(HxError, match="noiseScaleDeg"):
        cc.EvalAdd(ct, cc.EvalMult(ct, 0.5))                   # FIXEDMANUAL: operands must be rescaled by the caller
    fresh = hx.CryptoContext(orc.P.N, orc.P.q, orc.P.p, orc.P.psi_q, orc.P.psi_p, orc.P.dnum, orc.P.scale_bits, 0)
    sq = fresh.EvalMultNoRelin(fresh.upload(np.stack(orc.encrypt([0.5]).c), 0, 1), fresh.upload(np.stack(orc.encrypt([0.5]).c), 0, 1))
    with pytest.raises(HxError, match="EvalMultKey"):
        fresh.Relinearize(sq)                                  # EvalMultKey not generated
    # the two-step fused entries: unknown block, ciphertext that is not fresh, rotation set index out of range
    with pytest.raises(HxError):
        cc.db_get(0, 0)
    cc.db_register_ct(0, 0, ct)
    with pytest.raises(HxError):
        cc.db_register_ct(0, 1, low)
    with pytest.raises(HxError, match="not registered"):
        cc.score_blocks(ct, 1, [7])
    with pytest.raises(HxError):
        cc.score_blocks(ct, 1, [0], rotations=[1])
    with pytest.raises(HxError):
        cc.score_blocks(ct, 3, [0])                            # rotation keys 2 is missing: the hoisted batch must refuse
