"""Diagonal sharding of the multi-GPU sender (image_matching_b200/sharded.py) on ONE device: two contexts stand for two
ranks, each holding its diagonal range of every block and computing only its own hoisted rotations; their partial accumulators,
added as raw 64-bit words and folded (what the reduce-scatter + hx_ct_fold_mod do), relinearised and rescaled, must equal the
oracle's computeSimilarity (/root/reference/src/sender/sender_diag.cpp:12-33, 66-83) bit for bit."""
import numpy as np
import pytest

from oracle import ckks_oracle as O
from tests.helpers import make_gpu_context, make_oracle, same, up

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world", [2, 3])
def test_diagonal_sharded_partial_sums_bit_exact(world):
    from image_matching_b200.sharded import block_range, diag_range
    ndiag, blocks = 16, 3
    orc = make_oracle(ring_dim=4096, depth=11, rotations=list(range(1, ndiag)) + [2 ** i for i in range(11)], seed=4)
    rng = np.random.default_rng(3)
    dbc = [orc.encrypt(rng.uniform(-0.2, 0.2, orc.P.slots)) for _ in range(blocks * ndiag)]
    qc = orc.encrypt(rng.uniform(-0.2, 0.2, orc.P.slots))
    want = O.diag_compute_similarity(orc, qc, dbc, ndiag)
    ranks, parts = [], []
    for r in range(world):
        cc = make_gpu_context(orc)
        d0, d1 = diag_range(r, world, ndiag)
        cc.set_option("db_diag0", d0)
        cc.set_option("db_diag_count", d1 - d0)
        for g in range(blocks):
            for d in range(d0, d1):
                cc.db_register(g, d, np.stack(dbc[g * ndiag + d].c))
        with pytest.raises(Exception):
            cc.db_register(0, d1 % ndiag if d1 < ndiag else d0 - 1, np.stack(dbc[0].c))   # outside the resident range
        acc = cc.diag_mac_partial(up(cc, qc), d0, d1, range(blocks))
        ptr, words, nb = cc.device_view(acc)
        assert nb == blocks and words == blocks * 3 * orc.P.L * orc.N
        parts.append(np.stack([c.to_numpy() for c in cc.unstack(acc)]))               # [blocks][3][L][N]
        ranks.append(cc)
    total = np.zeros_like(parts[0])
    for p in parts:
        total = total + p                                                              # plain 64-bit sums: world * q < 2^63
    # every rank relinearises / rescales its share of the blocks from the summed words
    for r, cc in enumerate(ranks):
        lo, hi = block_range(r, world, blocks)
        if hi == lo:
            continue
        mine = cc.alloc(3, orc.P.L, 0, 2, hi - lo)
        staged = cc.stack([cc.upload(total[g], 0, 2) for g in range(lo, hi)])          # raw (un-reduced) words go in unchanged
        cc.fold_mod(staged)
        got = cc.unstack(cc.Rescale(cc.Relinearize(staged)))
        assert len(got) == hi - lo and all(same(gc, want[lo + i]) for i, gc in enumerate(got))
        del mine
