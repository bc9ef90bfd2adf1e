"""Shared test plumbing: an oracle context and a GPU context holding the SAME parameters, keys and
ciphertexts (the oracle plays OpenFHE's client: it generates keys and encrypts, and everything is
uploaded through the C ABI exactly as the reference-side shim would)."""
from __future__ import annotations

import numpy as np

from oracle import ckks_oracle as O


def make_oracle(ring_dim=4096, depth=11, rotations=(), seed=1, mult_key=True):
    P = O.gen_params(depth, ring_dim=ring_dim)
    orc = O.Oracle(P)
    orc.keygen(seed)
    if mult_key:
        orc.mult_keygen()
    if rotations:
        orc.rot_keygen(list(rotations))
    return orc


def make_gpu_context(orc, device=0):
    import image_matching_b200 as hx
    P = orc.P
    cc = hx.CryptoContext(P.N, P.q, P.p, P.psi_q, P.psi_p, P.dnum, P.scale_bits, device)
    if orc.mult_key is not None:
        cc.load_mult_key(orc.mult_key[0], orc.mult_key[1])
    for k, (a, b) in orc.rot_keys.items():
        cc.load_auto_key(k, a, b)
    return cc


def up(cc, ct: O.Ct):
    return cc.upload(np.stack(ct.c), ct.level, ct.deg)


def same(gpu_ct, ct: O.Ct) -> bool:
    ncomp, nl, level, deg = gpu_ct.info()
    if (ncomp, nl, level, deg) != (len(ct.c), ct.nl, ct.level, ct.deg):
        return False
    return bool(np.array_equal(gpu_ct.to_numpy(), np.stack(ct.c)))
