"""CPU tests of the product's host logic and C-ABI surface (no compute calls: no GPU here)."""
import os
import re

import numpy as np
import pytest

from oracle import ckks_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from image_matching_b200 import _lib
    lib = _lib.load()
    header = open(os.path.join(ROOT, "include", "hydia_b200.h")).read()
    declared = set(re.findall(r"\b(hx_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/hydia_b200.h but not exported"
    assert declared == set(_lib.SIGNATURES), "ctypes signature table out of sync with the header"


def test_no_cpu_fallback_without_device():
    import image_matching_b200 as hx
    from image_matching_b200._lib import HxError, load
    if load().hx_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(HxError, match="no CUDA device"):
        hx.CryptoContext.generate(depth=11, ring_dim=2048)


def test_product_parameter_generation_equals_oracle():
    """hx_params_generate (product, C++) vs oracle gen_params (independent restatement)."""
    import ctypes as C
    from image_matching_b200 import _lib
    lib = _lib.load()
    for depth, ring in ((11, 0), (11, 4096), (3, 2048), (13, 0), (12, 0), (18, 0)):   # 0: the recipe picks the ring
        N, L, K, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        q, p, sq, sp = (np.zeros(40, dtype=np.uint64) for _ in range(4))
        u = lambda a: a.ctypes.data_as(_lib.U64P)
        assert lib.hx_params_generate(depth, 45, 60, ring, 0, C.byref(N), C.byref(L), C.byref(K), C.byref(d), u(q), u(p), u(sq), u(sp)) == 0
        P = O.gen_params(depth, ring_dim=ring or None)
        assert (N.value, L.value, K.value, d.value) == (P.N, P.L, P.K, P.dnum)
        assert q[:P.L].tolist() == P.q and p[:P.K].tolist() == P.p
        assert sq[:P.L].tolist() == P.psi_q and sp[:P.K].tolist() == P.psi_p


def test_diagonal_packing_matches_oracle_and_reference_layout():
    from image_matching_b200 import diagonal_pack
    rng = np.random.default_rng(0)
    db = rng.integers(-99, 100, size=(1300, 512)).astype(np.float64)   # ragged: 3 square blocks, last one padded
    a, b = diagonal_pack(db, 2048), O.diagonal_pack(db, 2048)
    assert a.shape == b.shape == (512, 2048) and np.array_equal(a, b)
    dbn = O.plaintext_normalize(db)
    # cipher d, slot j*512 + r = db[(j*512 + r)][(r + d) % 512]
    for d, j, r in ((0, 0, 0), (5, 1, 7), (511, 2, 100), (17, 2, 300)):
        row = j * 512 + r
        want = dbn[row][(r + d) % 512] if row < 1300 else 0.0
        assert a[d, j * 512 + r] == want


def test_rotation_key_set_matches_main_cpp():
    from image_matching_b200 import hydia_rotation_set
    rot = hydia_rotation_set(16384)
    assert len(rot) == 530 and rot[:511] == list(range(1, 512))        # 530 keys at slots = 16384
    assert rot == O.hydia_rotation_set(16384)


def test_binary_rotate_decomposition():
    """OpenFHEWrapper::binaryRotate (openFHE_wrapper.cpp:103-128): 511 -> [512, -1]."""
    from image_matching_b200 import OpenFHEWrapper

    class FakeCC:
        def __init__(self): self.calls = []
        def GetBatchSize(self): return 16384
        def EvalRotate(self, ct, r): self.calls.append(r); return ct
    cc = FakeCC(); OpenFHEWrapper.binaryRotate(cc, None, 511); assert cc.calls == [512, -1]
    cc = FakeCC(); OpenFHEWrapper.binaryRotate(cc, None, -3); assert cc.calls == [-4, 1]
    cc = FakeCC(); OpenFHEWrapper.binaryRotate(cc, None, 16384); assert cc.calls == []
    assert OpenFHEWrapper.computeRequiredDepth(5) == 11


def test_chebyshev_coefficients_equal_oracle():
    from image_matching_b200 import OpenFHEWrapper
    f = lambda x: 1.0 if x >= 0.44 else -1.0
    assert OpenFHEWrapper.chebyshev_coefficients(f, -1.0, 1.0, 59) == O.cheb_coefficients(f, -1.0, 1.0, 59)
