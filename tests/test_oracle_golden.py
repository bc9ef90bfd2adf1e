"""CPU tests: the oracle against every known answer the reference holds for the path
(SURVEY.md §8(c) items 1-4). No GPU, no product code."""
import os

import numpy as np
import pytest

from oracle import ckks_oracle as O


# depth from computeRequiredDepth (/root/reference/src/openFHE_wrapper.cpp:6-44, COMP_DEPTH 10,
# ALPHA_DEPTH 2), (N, L) pinned by the ciphertext byte counts 2*L*N*8 behind /root/reference/tools/figures/
# 15IndexTotals.csv:2,5 and 15MembershipTotals.csv:2,5 (golden item 4 of SURVEY 8(c)).
APPENDIX_B = {   # approach: (depth, N, L, dnum, alpha, fresh ciphertext bytes, qBound of the ring-dimension estimate)
    1: (13, 65536, 14, 3, 5, 14680064, 945),
    2: (18, 65536, 19, 3, 7, 19922944, 1230),
    3: (12, 65536, 13, 3, 5, 13631488, 900),
    4: (11, 32768, 12, 3, 4, 6291456, 795),
    5: (11, 32768, 12, 3, 4, 6291456, 795),
}


@pytest.mark.parametrize("approach", [1, 2, 3, 4, 5])
def test_parameter_recipe_matches_reference_sizes(approach):
    depth, N, L, dnum, alpha, ct_bytes, qbound = APPENDIX_B[approach]
    assert O.estimate_qbound(depth) == qbound
    P = O.gen_params(depth)
    assert (P.N, P.L, P.dnum, P.alpha) == (N, L, dnum, alpha)
    assert 2 * P.L * P.N * 8 == ct_bytes
    # number of P limbs: PrecomputeCRTTables takes ceil(bits(largest digit product) / 60) on the GENERATED primes, while the
    # ring-dimension estimate above charges one 60-bit prime more whenever the nominal digit size is a multiple of 60
    # (addOne): approaches 1 and 3 have a 60 + 4*45 = 240-bit nominal digit whose real product is 239.99.. bits -> K = 4
    # (SURVEY App. B lists the estimate's 5; no reference-held fixture pins K)
    digit_bits = max(O._prod(P.q[j * P.alpha:(j + 1) * P.alpha]).bit_length() for j in range(P.dnum))
    assert P.K == -(-digit_bits // 60) == {1: 4, 2: 6, 3: 4, 4: 4, 5: 4}[approach]
    assert all(x % (2 * P.N) == 1 for x in P.q + P.p)
    assert len(set(P.q + P.p)) == P.L + P.K
    assert P.q[0].bit_length() == 60 and all(x.bit_length() == 60 for x in P.p)
    assert all(x.bit_length() in (45, 46) for x in P.q[1:])
    if approach == 5:
        assert P.q[0] == 0xFFFFFFFFFFC0001              # largest 60-bit prime = 1 mod 2^16
        assert [x.bit_length() for x in P.q[1:]] == [46, 45] * 5 + [46]
        assert 2 * P.dnum * (P.L + P.K) * P.N * 8 == 25165824   # one evaluation key 25.17 MB
    for q, psi in zip(P.q + P.p, P.psi_q + P.psi_p):
        assert pow(psi, P.N, q) == q - 1


def test_ps_degrees():
    assert O.compute_degrees_ps(59) == (4, 4)      # Chebyshev degree 59 (COMP_DEPTH 10 -> DEPTH_TO_DEGREE[6])
    assert O.compute_degrees_ps(9) == (4, 2)       # f4, degree 9


def test_compare_matches_signapprox_golden(golden_dir):
    """tools/figures/signApprox.csv `combined` = chebyshevCompare(0.44, 10) as checked in by the authors."""
    g = np.load(os.path.join(golden_dir, "signapprox.npz"))
    out = O.chebyshev_compare(O.PlainEval(), O.PlainCt(g["input"]), 0.44, 10)
    assert out.level == 10 and out.deg == 1            # COMP_DEPTH levels consumed
    assert np.abs(out.v - g["combined"]).max() <= 1e-4  # 5.9e-5 measured
    # effective decision threshold (>= 1.0, receiver_hers.cpp:30) sits between 0.4540 and 0.4542
    x = g["input"]
    assert np.all(out.v[x >= 0.4542][x[x >= 0.4542] <= 0.6] >= 1.0) and np.all(out.v[x <= 0.4540] < 1.0)


def test_chebyshev_division_reconstructs():
    rng = np.random.default_rng(0)
    f = list(rng.normal(size=28)); g = [0.0] * 16 + [1.0]
    q, r = O.long_div_cheb(f, g)
    # evaluate both sides on a grid in the Chebyshev basis (q's free term is stored doubled)
    xs = np.linspace(-1, 1, 33)
    T = lambda c: np.polynomial.chebyshev.chebval(xs, c)
    qq = list(q); qq[0] /= 2
    assert np.allclose(T(f), T(qq) * T(g) + T(r), atol=1e-9)


@pytest.fixture(scope="module")
def small():
    P = O.gen_params(11, ring_dim=2048)
    orc = O.Oracle(P)
    orc.keygen(3); orc.mult_keygen(); orc.rot_keygen([1, 2, 5, -1, -4, 512] + [2 ** i for i in range(10)])
    return orc


def test_encrypted_ops_small_ring(small):
    orc = small
    rng = np.random.default_rng(1)
    a, b = rng.uniform(-1, 1, orc.P.slots), rng.uniform(-1, 1, orc.P.slots)
    ca, cb = orc.encrypt(a), orc.encrypt(b)
    assert np.abs(orc.decrypt(ca) - a).max() < 1e-7
    assert np.abs(orc.decrypt(orc.rescale(orc.mult(ca, cb))) - a * b).max() < 1e-7
    for r in (1, 5, -1, -4, 512):
        assert np.abs(orc.decrypt(orc.rotate(ca, r)) - np.roll(a, -r)).max() < 1e-7
    dg = orc.fast_rotation_precompute(ca)
    fr, pr = orc.fast_rotation(ca, 2, dg), orc.rotate(ca, 2)
    assert all(np.array_equal(x, y) for x, y in zip(fr.c, pr.c))      # hoisted == plain, bit for bit
    s = orc.eval_sum(orc.level_reduce(ca, 10), orc.P.slots)
    assert abs(orc.decrypt(s)[0] - a.sum()) < 1e-5
    assert np.abs(orc.decrypt(orc.add_const(orc.mult_const(ca, -0.37), 0.25)) - (a * -0.37 + 0.25)).max() < 1e-7


def test_encrypted_compare_small_ring(small):
    orc = small
    x = np.random.default_rng(2).uniform(-0.3, 0.3, orc.P.slots)
    x[:4] = [0.93, 0.47, 0.43, -0.9]
    ct = orc.level_reduce(orc.encrypt(x), 1)
    out = orc.chebyshev_compare(ct)
    assert out.nl == 1 and out.level == 11
    ref = O.chebyshev_compare(O.PlainEval(), O.PlainCt(x)).v
    dec = orc.decrypt(out)
    assert np.abs(dec - ref).max() < 1e-4
    assert dec[0] >= 1.0 and dec[1] >= 1.0 and dec[2] < 1.0 and dec[3] < 1.0


@pytest.mark.slow
def test_hydia_2_10_known_answer(golden_dir):
    """run_artifact.sh:7-27 `./ImageMatching ../test/2_10.dat 5`: membership true, index [0]
    (tools/gen_dataset.sh:66-72), scores within 1e-4 of plaintext cosine (main_accuracy.cpp:360).
    Full parameters (N = 2^15, 12+4 limbs); ~2 minutes on 8 cores."""
    g = np.load(os.path.join(golden_dir, "dat_2_10.npz"))
    n, q, db = int(g["n"]), g["query"].astype(np.float64), g["db"].astype(np.float64)
    P = O.gen_params(11)
    orc = O.Oracle(P)
    orc.keygen(1); orc.mult_keygen()
    orc.rot_keygen(list(range(1, 512)) + [2 ** i for i in range(14)])   # rotations the path uses
    rows = O.diagonal_pack(db, P.slots)
    assert rows.shape == (512, 16384)
    dbc = [orc.encrypt(r) for r in rows]
    qc = O.encrypt_query_diag(orc, q)
    scores = O.diag_compute_similarity(orc, qc, dbc)
    cos = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    s = orc.decrypt(scores[0])
    assert np.abs(s[:n] - cos).max() <= 1e-4 and np.abs(s[n:]).max() <= 1e-4
    assert abs(cos[0] - 0.92680) < 1e-5
    idx = [O.chebyshev_compare(orc, x) for x in scores]
    assert O.decrypt_index(orc, idx) == [0]
    m = orc.eval_sum(orc.add_many(idx), P.slots)
    assert O.decrypt_membership(orc, m) is True


def test_hers_restatement_small_ring():
    """Row R7 on the CPU oracle: HersSender scores equal the plaintext cosines to 1e-4 (the tolerance of
    /root/reference/src/main_accuracy.cpp:354-364) and the receiver decisions are the planted matches."""
    dim, n = 8, 300
    orc = O.Oracle(O.gen_params(11, ring_dim=4096))
    orc.keygen(3)
    orc.mult_keygen()
    orc.rot_keygen([2 ** i for i in range(11)])
    rng = np.random.default_rng(4)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    q = rng.integers(1, 4, size=dim).astype(np.float64)
    c0 = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    db[c0 > 0.3] *= -1.0                      # keep every unplanted cosine well below the 0.44 threshold
    db[17] = 3 * q
    packed = O.hers_pack(db, orc.P.slots)
    dbc = [[orc.encrypt(packed[m, j]) for j in range(dim)] for m in range(packed.shape[0])]
    qc = O.encrypt_query_hers(orc, q)
    scores = O.hers_compute_similarity(orc, qc, dbc)
    cos = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    assert np.abs(orc.decrypt(scores[0])[:n] - cos).max() <= 1e-4
    idx = O.hers_index_scenario(orc, qc, dbc)
    assert O.decrypt_index(orc, idx) == [17]
    assert O.decrypt_membership(orc, orc.eval_sum(orc.add_many(idx), orc.P.slots))


def _pm_pow2_rotations(slots):
    return [s * (1 << k) for k in range(int(np.log2(slots))) for s in (1, -1)]


def test_base_restatement_small_ring():
    """Row R9 on the CPU oracle: BaseSender (EvalInnerProduct + mergeCiphers) gives the plaintext cosines, packed
    consecutively, to 1e-4; index decisions are the planted matches."""
    dim, n = 64, 80
    orc = O.Oracle(O.gen_params(13, ring_dim=4096))
    orc.keygen(5)
    orc.mult_keygen()
    orc.rot_keygen(_pm_pow2_rotations(orc.P.slots))
    rng = np.random.default_rng(6)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    q = rng.integers(1, 4, size=dim).astype(np.float64)
    c0 = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    db[c0 > 0.3] *= -1.0
    db[41] = 2 * q
    dbc = [orc.encrypt(r) for r in O.base_pack(db, orc.P.slots)]
    assert len(dbc) == 3
    qc = O.encrypt_query_base(orc, q)
    merged = O.base_compute_similarity(orc, qc, dbc, dim)
    assert len(merged) == 1
    cos = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    assert np.abs(orc.decrypt(merged[0])[:n] - cos).max() <= 1e-4
    idx = [O.chebyshev_compare(orc, m) for m in merged]
    assert O.decrypt_index(orc, idx) == [41]


def test_blind_restatement_small_ring():
    """Row R8 on the CPU oracle: BlindSender (chunked products, rotate-and-sum, compressCiphers) + BlindReceiver::decryptIndex."""
    dim, chunk, n = 64, 16, 200
    orc = O.Oracle(O.gen_params(12, ring_dim=4096))
    orc.keygen(8)
    orc.mult_keygen()
    orc.rot_keygen(_pm_pow2_rotations(orc.P.slots))
    rng = np.random.default_rng(9)
    db = rng.integers(-99, 100, size=(n, dim)).astype(np.float64)
    q = rng.integers(1, 4, size=dim).astype(np.float64)
    c0 = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    db[c0 > 0.3] *= -1.0
    db[5] = q; db[150] = 3 * q
    packed = O.blind_pack(db, orc.P.slots, chunk)
    assert packed.shape == (2, 4, orc.P.slots)
    dbc = [[orc.encrypt(packed[m, c]) for c in range(packed.shape[1])] for m in range(packed.shape[0])]
    qc = O.encrypt_query_blind(orc, q, chunk)
    comp = O.blind_compute_similarity(orc, qc, dbc, chunk)
    assert len(comp) == 1
    # slot j of the compressed cipher holds the score of vector j // chunk + (j % chunk) * (slots // chunk)
    cos = O.plaintext_normalize(db) @ O.plaintext_normalize(q)
    dec = orc.decrypt(comp[0])
    per = orc.P.slots // chunk
    for v in (5, 150, 17):
        m, i = divmod(v, per)
        assert abs(dec[i * chunk + m] - cos[v]) <= 1e-4
    idx = [O.chebyshev_compare(orc, c) for c in comp]
    assert sorted(O.blind_decrypt_index(orc, idx, chunk)) == [5, 150]
