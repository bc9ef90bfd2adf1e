"""Host-side mirror of the lbcrypto::CryptoContext<DCRTPoly> subset the reference's Sender path
uses, over the C ABI of libhydia_b200.so. Method names follow the reference's
call sites (cc->EvalAdd, cc->EvalMultNoRelin, ...); "InPlace" variants return the new ciphertext."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from ._lib import F64P, I32P, U64P, VP, check_int, check_ptr


def _u64p(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags.c_contiguous
    return a.ctypes.data_as(U64P)


class Ciphertext:
    """Ciphertext<DCRTPoly> handle (device resident)."""

    def __init__(self, cc: "CryptoContext", handle):
        self.cc = cc
        self._h = check_ptr(handle)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            _lib.load().hx_ct_free(h)

    def info(self):
        v = [C.c_int() for _ in range(4)]
        check_int(_lib.load().hx_ct_info(self._h, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)   # ncomp, nlimbs, level, noise_deg

    ncomp = property(lambda s: s.info()[0])
    nlimbs = property(lambda s: s.info()[1])
    def GetLevel(self): return self.info()[2]
    def GetNoiseScaleDeg(self): return self.info()[3]

    def download_ptr(self, host_ptr: int):
        check_int(_lib.load().hx_ct_download(self._h, C.cast(host_ptr, U64P)))

    def download_dev(self, dev_ptr: int):
        check_int(_lib.load().hx_ct_download_dev(self._h, VP(dev_ptr)))

    def download_dev_async(self, dev_ptr: int, stream: int):
        """hx_ct_download_dev_async: ordered against the caller's CUDA stream (raw cudaStream_t, 0 = default), no host sync."""
        check_int(_lib.load().hx_ct_download_dev_async(self._h, VP(dev_ptr), VP(stream)))

    def to_numpy(self) -> np.ndarray:
        ncomp, nl, _, _ = self.info()
        out = np.empty((ncomp, nl, self.cc.N), dtype=np.uint64)
        check_int(_lib.load().hx_ct_download(self._h, _u64p(out)))
        return out


class CryptoContext:
    """CryptoContext<DCRTPoly> for CKKS / FIXEDMANUAL / HYBRID (src/main.cpp:169-179)."""

    def __init__(self, N: int, q: Sequence[int], p: Sequence[int], psi_q: Sequence[int], psi_p: Sequence[int],
                 dnum: int, scale_bits: int = 45, device: int = 0):
        self.N, self.L, self.K, self.dnum, self.scale_bits, self.device = N, len(q), len(p), dnum, scale_bits, device
        self.q, self.p = list(q), list(p)
        arr = lambda v: np.array([int(x) for x in v], dtype=np.uint64)
        self._lib = _lib.load()
        self._h = check_ptr(self._lib.hx_ctx_create(N, self.L, self.K, dnum, _u64p(arr(q)), _u64p(arr(p)), _u64p(arr(psi_q)),
                                                    _u64p(arr(psi_p)), scale_bits, device))

    @classmethod
    def generate(cls, depth: int = 11, scale_bits: int = 45, first_mod_bits: int = 60, ring_dim: int = 0,
                 num_large_digits: int = 0, device: int = 0) -> "CryptoContext":
        """GenCryptoContext(HEStd_128_classic, depth, ScalingModSize 45, FIXEDMANUAL) (src/main.cpp:169-175)."""
        lib = _lib.load()
        N, L, K, d = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        q, p, sq, sp = (np.zeros(40, dtype=np.uint64) for _ in range(4))
        check_int(lib.hx_params_generate(depth, scale_bits, first_mod_bits, ring_dim, num_large_digits, C.byref(N), C.byref(L),
                                         C.byref(K), C.byref(d), _u64p(q), _u64p(p), _u64p(sq), _u64p(sp)))
        return cls(N.value, q[:L.value].tolist(), p[:K.value].tolist(), sq[:L.value].tolist(), sp[:K.value].tolist(), d.value,
                   scale_bits, device)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.hx_ctx_destroy(h)

    # ---- getters the reference uses
    def GetRingDimension(self) -> int: return self.N
    def GetBatchSize(self) -> int: return self.N // 2
    def sync(self): check_int(self._lib.hx_ctx_sync(self._h))
    def launches(self) -> int: return int(self._lib.hx_ctx_launches(self._h))
    def kernel_stamp(self) -> str: return self._lib.hx_kernel_stamp().decode()
    def set_option(self, name: str, value: int): check_int(self._lib.hx_ctx_set_option(self._h, name.encode(), int(value)))

    # ---- launch accounting / device timer (bench.py)
    def prof_enable(self, kinds=None):
        """Time the named kernel families with CUDA events on the launching stream (None = all, [] = off)."""
        names = [self._lib.hx_prof_name(k).decode() for k in range(self._lib.hx_prof_kinds())]
        mask = -1 if kinds is None else sum(1 << names.index(k) for k in kinds)
        check_int(self._lib.hx_prof_enable(self._h, mask))
    def prof_reset(self): check_int(self._lib.hx_prof_reset(self._h))
    def prof_read(self) -> dict:
        """{kernel family: {"launches", "ms", "bytes"}}; ms is filled only while the profiler is enabled."""
        out = {}
        for k in range(self._lib.hx_prof_kinds()):
            n, ms, by = C.c_uint64(), C.c_double(), C.c_double()
            check_int(self._lib.hx_prof_read(self._h, k, C.byref(n), C.byref(ms), C.byref(by)))
            out[self._lib.hx_prof_name(k).decode()] = {"launches": int(n.value), "ms": ms.value, "bytes": by.value}
        return out
    def timer_start(self): check_int(self._lib.hx_timer_start(self._h))
    def timer_stop(self) -> float:
        ms = C.c_double()
        check_int(self._lib.hx_timer_stop(self._h, C.byref(ms)))
        return ms.value

    def galois(self, rotation: int) -> int: return int(self._lib.hx_galois_for_rotation(self._h, rotation))

    # ---- keys / ciphertext transport
    def load_mult_key(self, a: np.ndarray, b: np.ndarray):
        check_int(self._lib.hx_key_load_mult(self._h, _u64p(np.ascontiguousarray(a)), _u64p(np.ascontiguousarray(b))))

    def load_auto_key(self, galois: int, a: np.ndarray, b: np.ndarray):
        check_int(self._lib.hx_key_load_auto(self._h, galois, _u64p(np.ascontiguousarray(a)), _u64p(np.ascontiguousarray(b))))

    def upload(self, data: np.ndarray, level: int = 0, noise_deg: int = 1) -> Ciphertext:
        data = np.ascontiguousarray(data, dtype=np.uint64)
        ncomp, nl, n = data.shape
        assert n == self.N
        return Ciphertext(self, self._lib.hx_ct_upload(self._h, ncomp, nl, _u64p(data), level, noise_deg))

    def upload_ptr(self, host_ptr: int, ncomp: int, nl: int, level: int = 0, noise_deg: int = 1) -> Ciphertext:
        """hx_ct_upload from a raw HOST address (e.g. a pinned staging buffer)."""
        return Ciphertext(self, self._lib.hx_ct_upload(self._h, ncomp, nl, C.cast(host_ptr, U64P), level, noise_deg))

    def upload_dev(self, dev_ptr: int, ncomp: int, nl: int, level: int = 0, noise_deg: int = 1) -> Ciphertext:
        """hx_ct_upload_dev from a DEVICE address on this context's GPU (NCCL staging buffer)."""
        return Ciphertext(self, self._lib.hx_ct_upload_dev(self._h, ncomp, nl, VP(dev_ptr), level, noise_deg))

    def upload_dev_async(self, dev_ptr: int, ncomp: int, nl: int, level: int, noise_deg: int, stream: int) -> Ciphertext:
        """hx_ct_upload_dev_async: ordered against the caller's CUDA stream (raw cudaStream_t, 0 = default), no host sync."""
        return Ciphertext(self, self._lib.hx_ct_upload_dev_async(self._h, ncomp, nl, VP(dev_ptr), level, noise_deg, VP(stream)))

    def _ct(self, h) -> Ciphertext:
        return Ciphertext(self, h)

    # ---- evaluation (names as in the reference's cc-> calls)
    def EvalAdd(self, a, b):
        if isinstance(b, (int, float)):
            return self._ct(self._lib.hx_add_const(a._h, float(b)))
        return self._ct(self._lib.hx_add(a._h, b._h))
    EvalAddInPlace = EvalAdd
    def EvalSub(self, a, b):
        if isinstance(b, (int, float)):
            return self._ct(self._lib.hx_add_const(a._h, -float(b)))
        return self._ct(self._lib.hx_sub(a._h, b._h))
    def EvalMult(self, a, b):
        if isinstance(b, (int, float)):
            return self._ct(self._lib.hx_mult_const(a._h, float(b)))
        if isinstance(b, np.ndarray):   # Plaintext = encoded limbs [nlimbs][N] (MakeCKKSPackedPlaintext)
            pt = np.ascontiguousarray(b, dtype=np.uint64)
            return self._ct(self._lib.hx_mult_plain(a._h, _u64p(pt), pt.shape[0]))
        return self._ct(self._lib.hx_mult(a._h, b._h))
    def EvalMultNoRelin(self, a, b): return self._ct(self._lib.hx_mult_norelin(a._h, b._h))
    def Relinearize(self, a): return self._ct(self._lib.hx_relin(a._h))
    RelinearizeInPlace = Relinearize
    def EvalSquare(self, a): return self._ct(self._lib.hx_square(a._h))
    def Rescale(self, a): return self._ct(self._lib.hx_rescale(a._h))
    RescaleInPlace = ModReduce = Rescale
    def LevelReduce(self, a, levels=1): return self._ct(self._lib.hx_level_reduce(a._h, levels))
    def EvalRotate(self, a, index): return self._ct(self._lib.hx_rotate(a._h, index))
    def EvalSum(self, a, n): return self._ct(self._lib.hx_eval_sum(a._h, n))
    def EvalInnerProduct(self, a, b, n): return self._ct(self._lib.hx_inner_product(a._h, b._h, n))
    def MakeCKKSPackedPlaintext(self, values) -> np.ndarray:
        """Encoded plaintext [L][N] (evaluation form). Encoding is client-side work: `encoder` is set by Client (or by a test
        to the oracle's encoder so both sides multiply by the same polynomial)."""
        if getattr(self, "encoder", None) is None:
            raise RuntimeError("MakeCKKSPackedPlaintext: no encoder attached (create a Client or set cc.encoder)")
        return self.encoder(values)
    def EvalAddMany(self, cts: List[Ciphertext]):
        arr = (VP * len(cts))(*[c._h for c in cts])
        return self._ct(self._lib.hx_add_many(arr, len(cts)))
    EvalAddManyInPlace = EvalAddMany

    class _Digits:
        def __init__(self, lib, h, cc=None):
            self._lib, self._h, self._cc = lib, check_ptr(h), cc
        def __del__(self):
            h, self._h = getattr(self, "_h", None), None
            if h:
                self._lib.hx_digits_free(h)

    def EvalFastRotationPrecompute(self, a): return CryptoContext._Digits(self._lib, self._lib.hx_rot_precompute(a._h), self)
    def EvalFastRotation(self, a, index, m, digits):
        assert m == 2 * self.N
        return self._ct(self._lib.hx_rot_fast(a._h, index, digits._h))

    def EvalChebyshevSeries(self, a, coeffs, lo, hi):
        c = np.ascontiguousarray(coeffs, dtype=np.float64)
        return self._ct(self._lib.hx_cheby_series(a._h, c.ctypes.data_as(F64P), len(c), lo, hi))
    def EvalPoly(self, a, coeffs):
        c = np.ascontiguousarray(coeffs, dtype=np.float64)
        return self._ct(self._lib.hx_poly(a._h, c.ctypes.data_as(F64P), len(c)))

    # ---- fused HyDia entry points
    def db_register(self, block: int, diag: int, data: np.ndarray):
        check_int(self._lib.hx_db_register(self._h, block, diag, _u64p(np.ascontiguousarray(data, dtype=np.uint64))))
    def db_blocks(self) -> int: return int(self._lib.hx_db_blocks(self._h))
    def score_diag(self, query: Ciphertext, ndiag: int = 512) -> List[Ciphertext]:
        out = (VP * 4096)()
        n = check_int(self._lib.hx_score_diag(self._h, query._h, ndiag, out, 4096))
        return [self._ct(out[i]) for i in range(n)]
    def membership(self, query, ndiag=512, delta=0.44, depth=10) -> Ciphertext:
        return self._ct(self._lib.hx_membership(self._h, query._h, ndiag, delta, depth))
    def index(self, query, ndiag=512, delta=0.44, depth=10) -> List[Ciphertext]:
        out = (VP * 4096)()
        n = check_int(self._lib.hx_index(self._h, query._h, ndiag, delta, depth, out, 4096))
        return [self._ct(out[i]) for i in range(n)]
    def compare(self, a, delta=0.44, depth=10): return self._ct(self._lib.hx_compare(a._h, delta, depth))
    # the fused path in two steps (what shim/openfhe.h's deferred layer calls)
    def db_register_ct(self, block: int, diag: int, ct: Ciphertext): check_int(self._lib.hx_db_register_ct(self._h, block, diag, ct._h))
    def db_get(self, block: int, diag: int) -> Ciphertext: return self._ct(self._lib.hx_db_get(self._h, block, diag))
    def score_blocks(self, query: Ciphertext, ndiag: int, blocks: Sequence[int], rotations: Sequence[int] = ()):
        """hx_rot_fast_all + hx_score_blocks (+ hx_rotset_get for the listed rotations): ([score per block], [rotated queries])."""
        rs = check_ptr(self._lib.hx_rot_fast_all(query._h, ndiag))
        try:
            b = np.array(list(blocks), dtype=np.int32)
            out = (VP * len(b))()
            check_int(self._lib.hx_score_blocks(rs, b.ctypes.data_as(_lib.I32P), len(b), out))
            return [self._ct(out[i]) for i in range(len(b))], [self._ct(self._lib.hx_rotset_get(rs, int(r))) for r in rotations]
        finally:
            self._lib.hx_rotset_free(rs)
    # ---- multi-GPU building blocks (sharded.py): diagonal-sharded partial sums, in-place collective access to ciphertext storage
    def download_many_ptr(self, cts: Sequence[Ciphertext], host_ptr: int):
        """Results back to back into one (pinned) host buffer with a single synchronisation (hx_ct_download_many)."""
        arr = (VP * len(cts))(*[c._h for c in cts])
        check_int(self._lib.hx_ct_download_many(arr, len(cts), C.c_void_p(host_ptr)))
    def diag_mac_partial(self, query: Ciphertext, d0: int, d1: int, blocks: Sequence[int]) -> Ciphertext:
        """Rotations d0..d1-1 of the query and sum over those diagonals of rotQ[d] (x) DB[g][d] for every named block: one batched
        3-component ciphertext [len(blocks)][3][L][N], not relinearised (hx_diag_mac_partial)."""
        b = np.array(list(blocks), dtype=np.int32)
        return self._ct(self._lib.hx_diag_mac_partial(self._h, query._h, int(d0), int(d1), b.ctypes.data_as(_lib.I32P), len(b)))
    def alloc(self, ncomp: int, nl: int, level: int, deg: int, nbatch: int = 1) -> Ciphertext:
        return self._ct(self._lib.hx_ct_alloc(self._h, ncomp, nl, level, deg, nbatch))
    def fold_mod(self, ct: Ciphertext): check_int(self._lib.hx_ct_fold_mod(ct._h))
    def device_view(self, ct: Ciphertext):
        """(device pointer, 64-bit words, batch entries) of a ciphertext's storage (hx_ct_device_view)."""
        ptr, words, nb = C.c_void_p(), C.c_uint64(), C.c_int()
        check_int(self._lib.hx_ct_device_view(ct._h, C.byref(ptr), C.byref(words), C.byref(nb)))
        return int(ptr.value), int(words.value), int(nb.value)
    def trace_start(self): check_int(self._lib.hx_trace_start(self._h))
    def trace_mark(self, name: str, stream: int = 0): check_int(self._lib.hx_trace_mark(self._h, name.encode(), C.c_void_p(stream or None)))
    def trace_dump(self) -> dict:
        """{mark name: [ms since trace_start, ...]} of the marks recorded since trace_start (hx_trace_dump; synchronises)."""
        buf = C.create_string_buffer(1 << 16)
        check_int(self._lib.hx_trace_dump(self._h, buf, len(buf)))
        out = {}
        for item in buf.value.decode().split(";"):
            if item:
                k, v = item.rsplit(":", 1)
                out.setdefault(k, []).append(float(v))
        return out
    def rescale_noalloc(self, in_ptr: int, ncomp: int, nl: int, out_ptr: int):
        """hx_rescale_noalloc: Rescale on caller-owned device buffers ([ncomp][nl][N] -> [ncomp][nl-1][N]) with the scratch in a
        static device array of the library; allocates nothing, queued on the engine stream."""
        check_int(self._lib.hx_rescale_noalloc(self._h, C.c_void_p(in_ptr), int(ncomp), int(nl), C.c_void_p(out_ptr)))
    def stream_order(self, stream: int, engine_waits: bool): check_int(self._lib.hx_stream_order(self._h, C.c_void_p(stream), 1 if engine_waits else 0))
    def stack(self, cts: Sequence[Ciphertext]) -> Ciphertext:
        """n ciphertexts of one shape -> one batched handle (hx_ct_stack); every Eval* call then works on all entries at once."""
        n = len(cts)
        return self._ct(self._lib.hx_ct_stack((VP * n)(*[c._h for c in cts]), n))
    def unstack(self, ct: Ciphertext) -> List[Ciphertext]:
        out = (VP * 65536)()
        n = check_int(self._lib.hx_ct_unstack(ct._h, out, 65536))
        return [self._ct(out[i]) for i in range(n)]
    def compare_many(self, cts: Sequence[Ciphertext], delta=0.44, depth=10) -> List[Ciphertext]:
        """chebyshevCompare of every ciphertext of the list through one batched chain (hx_compare_many)."""
        n = len(cts)
        inp, out = (VP * n)(*[c._h for c in cts]), (VP * n)()
        check_int(self._lib.hx_compare_many(self._h, inp, n, delta, depth, out))
        return [self._ct(out[i]) for i in range(n)]
    def score_hers(self, query: Sequence[Ciphertext]) -> List[Ciphertext]:
        """HersSender::computeSimilarity on the registered database matrices (hx_score_hers)."""
        n = len(query)
        inp, out = (VP * n)(*[c._h for c in query]), (VP * 4096)()
        m = check_int(self._lib.hx_score_hers(self._h, inp, n, out, 4096))
        return [self._ct(out[i]) for i in range(m)]

    # ---- parity-test hook
    def ntt(self, data: np.ndarray, primes: Sequence[int], inverse: bool = False) -> np.ndarray:
        out = np.ascontiguousarray(data, dtype=np.uint64).copy()
        pr = np.array(list(primes), dtype=np.int32)
        check_int(self._lib.hx_ntt(self._h, _u64p(out), out.shape[0], pr.ctypes.data_as(I32P), len(pr), 1 if inverse else 0))
        return out
