"""GPU parity at the shapes that are TIMED and at the parameter sets the reference uses:

* bench.py's exact launch shape -- N = 2^15, 12 + 4 limbs, 512 diagonals (16 key-switch batches of 32 rotations, q0 fold every 64
  terms), 2 database blocks, rot_lane 1, the warp-task ModDown kernel -- bit-exact against the oracle for score_diag / index / membership
  (/root/reference/src/sender/sender_diag.cpp:12-63), the GPU side repeated >= 20 times back to back (membership -> index ->
  membership) against one oracle answer: the regression test for cross-stream / stale-line races in the hoisted-rotation batch.
* N = 2^16 with the limb counts of approach 3 (depth 12: 13 + 4) and approach 1 (depth 13: 14 + 4), and approach 2's 19 + 6
  (7-limb digits): key switch, rescale, hoisted rotation, tensor product, EvalInnerProduct, EvalMult(plaintext) bit-exact
  (/root/reference/src/sender/sender_base.cpp:84-98, src/sender/sender_blind.cpp:59-83).
"""
import numpy as np
import pytest

from oracle import ckks_oracle as O
from tests.helpers import make_gpu_context, make_oracle, same, up

pytestmark = pytest.mark.gpu


def test_bench_shape_bit_exact_repeated():
    ndiag, blocks, repeats = 512, 2, 20
    orc = make_oracle(ring_dim=None, depth=11, rotations=list(range(1, ndiag)) + [2 ** i for i in range(14)], seed=21)
    assert (orc.N, orc.P.L, orc.P.K) == (32768, 12, 4)
    cc = make_gpu_context(orc)
    cc.set_option("rot_lane", 1)
    cc.set_option("ks_moddown_w", 1)
    rng = np.random.default_rng(77)
    # the database blocks reuse 64 distinct ciphertexts each (the arithmetic per diagonal does not care, encryption time does)
    base = [orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots)) for _ in range(64 * blocks)]
    dbc = [base[b * 64 + (d * 7 + b) % 64] for b in range(blocks) for d in range(ndiag)]
    for i, c in enumerate(dbc):
        cc.db_register(i // ndiag, i % ndiag, np.stack(c.c))
    qc = orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots))
    gq = up(cc, qc)
    want = O.diag_compute_similarity(orc, qc, dbc, ndiag)
    assert len(want) == blocks
    widx = [O.chebyshev_compare(orc, s) for s in want]
    wm = orc.eval_sum(orc.add_many(widx), orc.P.slots)
    got = cc.score_diag(gq, ndiag)
    assert len(got) == blocks and all(same(g, w) for g, w in zip(got, want))
    bad = []
    for it in range(repeats):
        m1, i1, m2 = cc.membership(gq, ndiag), cc.index(gq, ndiag), cc.membership(gq, ndiag)   # no host sync in between
        if not (same(m1, wm) and same(m2, wm) and len(i1) == blocks and all(same(g, w) for g, w in zip(i1, widx))):
            bad.append(it)
    assert not bad, f"mismatch against the oracle in repeats {bad} of {repeats}"
    # and with everything ordered on the engine stream / the r1 ModDown launch sequence (k_inv_cols + k_conv_fwd_col)
    for lane, mw in ((0, 1), (1, 0), (0, 0)):
        cc.set_option("rot_lane", lane)
        cc.set_option("ks_moddown_w", mw)
        for _ in range(3):
            assert same(cc.membership(gq, ndiag), wm) and all(same(g, w) for g, w in zip(cc.index(gq, ndiag), widx)), (lane, mw)


@pytest.mark.parametrize("depth,L,K", [(12, 13, 4), (13, 14, 4), (18, 19, 6)])
def test_ring_2p16_reference_limb_counts_bit_exact(depth, L, K):
    rots = [1, 2, 5, 64, 511, 512, -1, -3, -128] + [2 ** i for i in range(9)]
    orc = make_oracle(ring_dim=None, depth=depth, rotations=rots, seed=depth)
    assert (orc.N, orc.P.L, orc.P.K) == (65536, L, K)
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(depth)
    a, b = rng.uniform(-1, 1, orc.P.slots), rng.uniform(-1, 1, orc.P.slots)
    ca, cb = orc.encrypt(a), orc.encrypt(b)
    ga, gb = up(cc, ca), up(cc, cb)
    # tensor product, relinearisation (key switch with the mult key at full level), rescale
    t_w, t_g = orc.mult_norelin(ca, cb), cc.EvalMultNoRelin(ga, gb)
    assert same(t_g, t_w)
    r_w, r_g = orc.relinearize(t_w), cc.Relinearize(t_g)
    assert same(r_g, r_w)
    s_w, s_g = orc.rescale(r_w), cc.Rescale(r_g)
    assert same(s_g, s_w)
    # rotations (key switch + automorphism) at full level and one level down (shorter last digit)
    for r in (1, 5, 511, -3, 512):
        assert same(cc.EvalRotate(ga, r), orc.rotate(ca, r)), r
        assert same(cc.EvalRotate(s_g, r), orc.rotate(s_w, r)), ("lower level", r)
    # hoisted rotations (EvalFastRotationPrecompute + EvalFastRotation)
    dw, dg = orc.fast_rotation_precompute(ca), cc.EvalFastRotationPrecompute(ga)
    for r in (1, 2, 64, -1, -128):
        assert same(cc.EvalFastRotation(ga, r, 2 * orc.N, dg), orc.fast_rotation(ca, r, dw)), r
    # EvalInnerProduct(q, db, 512) + rescale (sender_base.cpp:93-95)
    ip_w = orc.rescale(orc.inner_product(ca, cb, 512))
    ip_g = cc.Rescale(cc.EvalInnerProduct(ga, gb, 512))
    assert same(ip_g, ip_w)
    # EvalMult(ct, Plaintext) + rescale (mask multiplications of mergeSingleCipher / compressCiphers, openFHE_wrapper.cpp:235,289)
    mask = np.zeros(orc.P.slots); mask[::128] = 1.0
    pt = orc.encode(mask, s_w.nl)
    mp_w = orc.rescale(orc.mult_plain(s_w, pt))
    mp_g = cc.Rescale(cc.EvalMult(s_g, pt))
    assert same(mp_g, mp_w)
    # decrypted-level sanity with the tolerance the reference states (main_accuracy.cpp:354-364)
    assert np.abs(orc.decrypt(s_w) - a * b).max() <= 1e-4


def test_ragged_diagonal_count_and_odd_block_count_bit_exact():
    """N = 2^15 with a diagonal count that is not a multiple of the 32-rotation batch (37 = 32 + a batch of 4 rotations after
    rotation 0, the small-batch launch shapes of every key-switch kernel) and three database blocks (the batched compare chain on an
    odd batch): score_diag / index / membership against the oracle (sender_diag.cpp:12-63), with two rotation-batch sizes."""
    ndiag, blocks = 37, 3
    orc = make_oracle(ring_dim=None, depth=11, rotations=list(range(1, ndiag)) + [2 ** i for i in range(14)], seed=23)
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(5)
    base = [orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots)) for _ in range(16)]
    dbc = [base[(d * 5 + 3 * b) % 16] for b in range(blocks) for d in range(ndiag)]
    for i, c in enumerate(dbc):
        cc.db_register(i // ndiag, i % ndiag, np.stack(c.c))
    qc = orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots))
    gq = up(cc, qc)
    want = O.diag_compute_similarity(orc, qc, dbc, ndiag)
    widx = [O.chebyshev_compare(orc, s) for s in want]
    wm = orc.eval_sum(orc.add_many(widx), orc.P.slots)
    for chunk in (32, 5):
        cc.set_option("rot_chunk", chunk)
        got = cc.score_diag(gq, ndiag)
        assert len(got) == blocks and all(same(g, w) for g, w in zip(got, want)), chunk
        assert all(same(g, w) for g, w in zip(cc.index(gq, ndiag), widx)), chunk
        assert same(cc.membership(gq, ndiag), wm), chunk
