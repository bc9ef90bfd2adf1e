"""hx_rescale_noalloc (include/hydia_b200.h): the allocation-free variant of Rescale against hx_rescale, through the C ABI.

No counterpart in the reference; the existing entry point hx_rescale -- itself
bit-exact against the oracle's DropLastElementAndScale in tests/test_gpu_parity.py -- is the checker here.  Integer residues: the
bar is bit-exact, like every other parity test of this repository.  Collected last on purpose
(file name).

* two calls in a row on one context at two DIFFERENT levels (the second rescales the first one's output): the shared static
  scratch keeps the first call's conversion table and limbs, so anything stale would show in the second result;
* a 3-component (not yet relinearised) ciphertext, and the bench ring (N = 2^15, 12 limbs);
* sizes beyond HX_NOALLOC_MAX_WORDS, a limb count of 1 and null pointers are rejected with an error, nothing is launched.
"""
import numpy as np
import pytest

from tests.helpers import make_gpu_context, make_oracle, up

pytestmark = pytest.mark.gpu

MAX_WORDS = 1 << 21   # HX_NOALLOC_MAX_WORDS


def _dev_buffer(words):
    import torch
    buf = torch.full((words,), -1, dtype=torch.int64, device="cuda:0")   # caller-owned output, poisoned
    torch.cuda.synchronize()
    return buf


def _noalloc_rescale(cc, ct):
    """Rescale `ct` through hx_rescale_noalloc into a buffer this test owns; returns the residues [ncomp][nl-1][N] as numpy."""
    ncomp, nl, _, _ = ct.info()
    in_ptr, words, nb = cc.device_view(ct)
    assert nb == 1 and words == ncomp * nl * cc.N
    out = _dev_buffer(ncomp * (nl - 1) * cc.N)
    cc.rescale_noalloc(in_ptr, ncomp, nl, out.data_ptr())
    cc.sync()
    return out.cpu().numpy().view(np.uint64).reshape(ncomp, nl - 1, cc.N), out


@pytest.mark.parametrize("ring_dim,depth", [(4096, 5), (None, 11)])
def test_noalloc_rescale_matches_rescale_twice_in_a_row(ring_dim, depth):
    orc = make_oracle(ring_dim=ring_dim, depth=depth, seed=5)
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(11)
    a, b = rng.uniform(-1, 1, orc.P.slots), rng.uniform(-1, 1, orc.P.slots)
    ga, gb = up(cc, orc.encrypt(a)), up(cc, orc.encrypt(b))
    prod = cc.EvalMult(ga, gb)                      # relinearised, 2 components, full level, degree 2
    want1 = cc.Rescale(prod)                        # existing entry point
    got1, keep1 = _noalloc_rescale(cc, prod)
    assert np.array_equal(got1, want1.to_numpy())
    # second call straight after the first, one level down, on the first call's own output
    sq = cc.EvalMult(want1, want1)
    want2 = cc.Rescale(sq)
    got2, _ = _noalloc_rescale(cc, sq)
    assert np.array_equal(got2, want2.to_numpy())
    # the same two calls back to back with NO host synchronisation or other kernel in between
    import torch
    n1, l1, _, _ = prod.info()
    n2, l2, _, _ = sq.info()
    o1, o2 = _dev_buffer(n1 * (l1 - 1) * cc.N), _dev_buffer(n2 * (l2 - 1) * cc.N)
    p1, p2 = cc.device_view(prod)[0], cc.device_view(sq)[0]
    cc.rescale_noalloc(p1, n1, l1, o1.data_ptr())
    cc.rescale_noalloc(p2, n2, l2, o2.data_ptr())
    cc.rescale_noalloc(p1, n1, l1, o1.data_ptr())
    cc.sync()
    assert np.array_equal(o1.cpu().numpy().view(np.uint64).reshape(got1.shape), want1.to_numpy())
    assert np.array_equal(o2.cpu().numpy().view(np.uint64).reshape(got2.shape), want2.to_numpy())
    # a 3-component ciphertext (EvalMultNoRelin output)
    t3 = cc.EvalMultNoRelin(ga, gb)
    want3 = cc.Rescale(t3)
    got3, _ = _noalloc_rescale(cc, t3)
    assert got3.shape[0] == 3 and np.array_equal(got3, want3.to_numpy())
    # decrypted-level sanity of what came out of the allocation-free call (tolerance of main_accuracy.cpp:354-364)
    assert np.abs(orc.decrypt(orc.rescale(orc.mult(orc.encrypt(a), orc.encrypt(b)))) - a * b).max() <= 1e-4
    del keep1, torch


def test_noalloc_rejects_oversize_and_bad_arguments():
    orc = make_oracle(ring_dim=4096, depth=5, seed=6)
    cc = make_gpu_context(orc)
    ga = up(cc, orc.encrypt(np.zeros(orc.P.slots)))
    ncomp, nl, _, _ = ga.info()
    in_ptr = cc.device_view(ga)[0]
    out = _dev_buffer(ncomp * nl * cc.N)
    launches = cc.launches()
    with pytest.raises(RuntimeError, match="no limb left|limb count"):
        cc.rescale_noalloc(in_ptr, ncomp, 1, out.data_ptr())
    with pytest.raises(RuntimeError, match="limb count"):
        cc.rescale_noalloc(in_ptr, ncomp, orc.P.L + 1, out.data_ptr())
    with pytest.raises(RuntimeError, match="components"):
        cc.rescale_noalloc(in_ptr, 4, nl, out.data_ptr())
    with pytest.raises(RuntimeError, match="null"):
        cc.rescale_noalloc(0, ncomp, nl, out.data_ptr())
    assert cc.launches() == launches
    assert ncomp * nl * cc.N <= MAX_WORDS


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    # N = 2^16 with 19 limbs (approach 2's parameters): 3 x 19 x 65536 words > 2^21 -- rejected with the limit in the message
    orc = make_oracle(ring_dim=None, depth=18, seed=7, mult_key=False)
    assert orc.N == 65536
    cc = make_gpu_context(orc)
    ga = up(cc, orc.encrypt(np.zeros(orc.P.slots)))
    ncomp, nl, _, _ = ga.info()
    assert ncomp * nl * cc.N > MAX_WORDS
    out = _dev_buffer(16)
    with pytest.raises(RuntimeError, match="exceeds the static scratch"):
        cc.rescale_noalloc(cc.device_view(ga)[0], ncomp, nl, out.data_ptr())
