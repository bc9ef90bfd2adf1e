"""oracle/ckks_oracle.py -- CPU oracle for the HyDia Sender hot path (scheme level).

TEST INFRASTRUCTURE ONLY (see oracle/ckks_ref.c header): imported by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs as the checker
or the timed CPU arm -- never by image_matching_b200.

PARITY STATUS: polynomial level "parity unpinned" (OpenFHE v1.2.3 is an un-vendored
dependency, /root/reference/dockerfile:7, and the reference holds no ciphertext KAT); decrypted
level pinned by tests/test_oracle_golden.py against tools/figures/signApprox.csv and
test/2_10.dat.

This file restates, on top of the plain-C limb arithmetic in ckks_ref.c, the CKKS-RNS
evaluation algorithms OpenFHE v1.2.3 runs for the reference's calls:
  parameter generation            <- /root/reference/src/main.cpp:169-175
  key generation (hybrid KS keys) <- main.cpp:184-206
  encode / encrypt / decrypt      <- src/openFHE_wrapper.cpp:74-85
  EvalFastRotation(Precompute)    <- src/sender/sender_diag.cpp:22-26
  EvalMultNoRelin / EvalAddInPlace / RelinearizeInPlace / RescaleInPlace <- sender_diag.cpp:75-93
  EvalChebyshevFunction / EvalPoly / EvalAddInPlace(ct, 1.0) <- openFHE_wrapper.cpp:143-185
  EvalAddManyInPlace / EvalSum    <- sender_diag.cpp:46-47
Upstream file names are given per function (they are not under /root/reference).
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libckks_ref.so")


def build(force: bool = False) -> str:
    """Compile ckks_ref.c (gcc) if the shared object is missing or stale."""
    src = os.path.join(_HERE, "ckks_ref.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


_lib = None
U64P = C.POINTER(C.c_uint64)
U32P = C.POINTER(C.c_uint32)
I32P = C.POINTER(C.c_int)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u64 = C.c_uint64
        L.ref_mulmod.restype = u64; L.ref_mulmod.argtypes = [u64, u64, u64]
        L.ref_powmod.restype = u64; L.ref_powmod.argtypes = [u64, u64, u64]
        L.ref_invmod.restype = u64; L.ref_invmod.argtypes = [u64, u64]
        L.ref_is_prime.restype = C.c_int; L.ref_is_prime.argtypes = [u64]
        L.ref_first_prime.restype = u64; L.ref_first_prime.argtypes = [C.c_int, u64]
        L.ref_next_prime.restype = u64; L.ref_next_prime.argtypes = [u64, u64]
        L.ref_prev_prime.restype = u64; L.ref_prev_prime.argtypes = [u64, u64]
        L.ref_min_root_of_unity.restype = u64; L.ref_min_root_of_unity.argtypes = [u64, u64]
        L.ref_ctx_new.restype = C.c_void_p; L.ref_ctx_new.argtypes = [u64, C.c_int, U64P, U64P]
        L.ref_ctx_free.argtypes = [C.c_void_p]
        L.ref_ntt.argtypes = [C.c_void_p, U64P, C.c_int, I32P, C.c_int]
        for nm in ("ref_add", "ref_sub", "ref_mul", "ref_mul_acc", "ref_mul_scalar", "ref_add_scalar"):
            getattr(L, nm).argtypes = [U64P, U64P, U64P, C.c_int, u64, U64P]
        L.ref_automap.argtypes = [u64, u64, U32P]
        L.ref_automorph.argtypes = [U64P, U64P, C.c_int, u64, U32P]
        L.ref_baseconv.argtypes = [U64P, U64P, C.c_int, U64P, C.c_int, U64P, u64, U64P, U64P]
        L.ref_switch_modulus.argtypes = [U64P, U64P, u64, u64, u64]
        L.ref_tensor_acc.argtypes = [U64P, U64P, U64P, C.c_int, u64, U64P]
        L.ref_set_threads.argtypes = [C.c_int]
        L.ref_max_threads.restype = C.c_int
        _lib = L
    return _lib


def _p(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags.c_contiguous
    return a.ctypes.data_as(U64P)


def _u64arr(v: Sequence[int]) -> np.ndarray:
    return np.array([int(x) for x in v], dtype=np.uint64)


# ----------------------------------------------------------------------------- parameters

# HE-standard bound on log2(QP) for 128-bit classical security, uniform ternary secret
# (upstream lattice/stdlatticeparms.cpp, HEStd_128_classic / HEStd_ternary).
_HESTD_128_CLASSIC_TERNARY = {1024: 27, 2048: 54, 4096: 109, 8192: 218, 16384: 438, 32768: 881,
                              65536: 1747, 131072: 3523}


def estimate_qbound(depth: int, scale_bits: int = 45, first_mod: int = 60, num_large_digits: int = 0, aux_bits: int = 60) -> int:
    """qBound of ParamsGenCKKSRNS for HYBRID key switching (upstream ckksrns-parametergeneration.cpp +
    CryptoParametersRNS::EstimateLogP with addOne = true), FIXEDMANUAL (no extra modulus)."""
    L = depth + 1
    dnum = num_large_digits or (3 if depth > 3 else (2 if depth > 0 else 1))
    alpha = -(-L // dnum)
    dnum = -(-L // alpha)
    bits = [scale_bits] * L
    bits[0] = first_mod
    max_bits = max(sum(bits[j * alpha:min(L, (j + 1) * alpha)]) for j in range(dnum)) + 1   # addOne
    size_p = -(-max_bits // aux_bits)
    return first_mod + (L - 1) * scale_bits + size_p * aux_bits


def ring_dim_for(depth: int, scale_bits: int = 45, first_mod: int = 60, num_large_digits: int = 0, aux_bits: int = 60) -> int:
    """StdLatticeParm::FindRingDim(HEStd_ternary, HEStd_128_classic, qBound)."""
    qb = estimate_qbound(depth, scale_bits, first_mod, num_large_digits, aux_bits)
    for N in sorted(_HESTD_128_CLASSIC_TERNARY):
        if qb <= _HESTD_128_CLASSIC_TERNARY[N]:
            return N
    return 2 * max(_HESTD_128_CLASSIC_TERNARY)


@dataclass
class Params:
    N: int
    L: int                 # number of Q limbs (= depth + 1 in FIXEDMANUAL)
    K: int                 # number of P limbs
    dnum: int              # number of digits
    alpha: int             # limbs per digit
    q: List[int]
    p: List[int]
    psi_q: List[int]
    psi_p: List[int]
    scale_bits: int = 45

    @property
    def slots(self) -> int:
        return self.N // 2


def gen_params(depth: int, scale_bits: int = 45, first_mod: int = 60, ring_dim: Optional[int] = None,
               num_large_digits: int = 0, aux_bits: int = 60) -> Params:
    """ParameterGenerationCKKSRNS::ParamsGenCKKSRNS + CryptoParametersRNS::PrecomputeCRTTables
    (upstream scheme/ckksrns/ckksrns-parametergeneration.cpp, schemerns/rns-cryptoparameters.cpp)
    for FIXEDMANUAL / HYBRID, as configured at /root/reference/src/main.cpp:169-173.
    ring_dim=None picks the smallest HE-standard ring (128-bit classic); tests pass a small ring."""
    L = depth + 1
    dnum = num_large_digits
    if dnum == 0:
        dnum = 3 if depth > 3 else (2 if depth > 0 else 1)
    alpha = -(-L // dnum)
    dnum = -(-L // alpha)

    def pick(N: int):
        m = 2 * N
        l = lib()
        q = [0] * L
        q[L - 1] = l.ref_first_prime(scale_bits, m)
        q_prev = q_next = q[L - 1]
        cnt = 0
        for i in range(L - 2, 0, -1):
            if cnt % 2 == 0:
                q_prev = l.ref_prev_prime(q_prev, m); q[i] = q_prev
            else:
                q_next = l.ref_next_prime(q_next, m); q[i] = q_next
            cnt += 1
        if first_mod == scale_bits:
            q[0] = l.ref_prev_prime(q_prev, m)
        else:
            q[0] = l.ref_prev_prime(l.ref_first_prime(first_mod, m), m)
        # digit moduli -> number of special primes
        max_bits = 0
        for j in range(dnum):
            prod = 1
            for i in range(j * alpha, min(L, (j + 1) * alpha)):
                prod *= q[i]
            max_bits = max(max_bits, prod.bit_length())
        K = -(-max_bits // aux_bits)
        p = []
        p_prev = l.ref_first_prime(aux_bits, m)
        for _ in range(K):
            while True:
                p_prev = l.ref_prev_prime(p_prev, m)
                if p_prev not in q:
                    break
            p.append(p_prev)
        return q, p

    if ring_dim is None:
        # Ring dimension: StdLatticeParm::FindRingDim on the NOMINAL bit budget qBound = firstModSize + (L-1)*scalingModSize
        # + logP, where logP comes from CryptoParametersRNS::EstimateLogP(..., addOne = true): the largest digit's nominal
        # bits PLUS ONE, rounded up to whole 60-bit auxiliary primes.  The estimate, not the generated
        # primes, picks N -- depth 12 gives 60 + 12*45 + 300 = 900 > 881, hence N = 2^16, which is what the ciphertext byte
        # counts of /root/reference/tools/figures/15IndexTotals.csv:2,5 pin for approach 3.  The actual number of P limbs
        # below (pick) still follows PrecomputeCRTTables: ceil(bits(largest digit product) / 60) on the generated primes.
        N = ring_dim_for(depth, scale_bits, first_mod, num_large_digits, aux_bits)
        q, p = pick(N)
    else:
        N = ring_dim
        q, p = pick(N)
    l = lib()
    psi_q = [l.ref_min_root_of_unity(2 * N, x) for x in q]
    psi_p = [l.ref_min_root_of_unity(2 * N, x) for x in p]
    return Params(N=N, L=L, K=len(p), dnum=dnum, alpha=alpha, q=q, p=p, psi_q=psi_q, psi_p=psi_p,
                  scale_bits=scale_bits)


# ----------------------------------------------------------------------------- ciphertext

@dataclass
class Ct:
    """Ciphertext<DCRTPoly>: `c` = list of components, each uint64 [nl][N] in EVALUATION format,
    limbs q_0..q_{nl-1}. level = GetLevel() (limbs dropped), deg = GetNoiseScaleDeg()."""
    c: List[np.ndarray]
    level: int = 0
    deg: int = 1

    @property
    def nl(self) -> int:
        return self.c[0].shape[0]

    def clone(self) -> "Ct":
        return Ct([x.copy() for x in self.c], self.level, self.deg)


def _prod(xs):
    r = 1
    for x in xs:
        r *= x
    return r


class Oracle:
    """CryptoContext<DCRTPoly> restatement for the evaluation subset listed in the module doc."""

    def __init__(self, params: Params):
        self.P = params
        P = params
        self.N, self.L, self.K = P.N, P.L, P.K
        self.mods = list(P.q) + list(P.p)
        self.psis = list(P.psi_q) + list(P.psi_p)
        self.modarr = _u64arr(self.mods)
        self._ctx = lib().ref_ctx_new(P.N, len(self.mods), _p(self.modarr), _p(_u64arr(self.psis)))
        self.scale = float(2 ** P.scale_bits)
        self._automap_cache: Dict[int, np.ndarray] = {}
        self._modup_cache = {}
        self._moddown_cache = {}
        self.rot_keys: Dict[int, tuple] = {}
        self.mult_key = None
        self.Pprod = _prod(P.p)

    def __del__(self):
        try:
            lib().ref_ctx_free(self._ctx)
        except Exception:
            pass

    # ---- limb helpers
    def qmods(self, nl: int) -> np.ndarray:
        return self.modarr[:nl]

    def ntt(self, a: np.ndarray, idx: Sequence[int], inverse: bool = False) -> None:
        ia = np.array(list(idx), dtype=np.int32)
        lib().ref_ntt(self._ctx, _p(a), a.shape[0], ia.ctypes.data_as(I32P), 1 if inverse else 0)

    def _bin(self, fn, a, b, mods):
        out = np.empty_like(a)
        getattr(lib(), fn)(_p(out), _p(a), _p(b), a.shape[0], self.N, _p(np.ascontiguousarray(mods)))
        return out

    def padd(self, a, b, mods): return self._bin("ref_add", a, b, mods)
    def psub(self, a, b, mods): return self._bin("ref_sub", a, b, mods)
    def pmul(self, a, b, mods): return self._bin("ref_mul", a, b, mods)

    def pmul_scalar(self, a, scal: Sequence[int], mods):
        return self._bin("ref_mul_scalar", a, _u64arr(scal), mods)

    def padd_scalar(self, a, scal: Sequence[int], mods):
        return self._bin("ref_add_scalar", a, _u64arr(scal), mods)

    def automap(self, k: int) -> np.ndarray:
        m = self._automap_cache.get(k)
        if m is None:
            m = np.empty(self.N, dtype=np.uint32)
            lib().ref_automap(self.N, k, m.ctypes.data_as(U32P))
            self._automap_cache[k] = m
        return m

    def pautomorph(self, a: np.ndarray, k: int) -> np.ndarray:
        out = np.empty_like(a)
        lib().ref_automorph(_p(out), _p(a), a.shape[0], self.N, self.automap(k).ctypes.data_as(U32P))
        return out

    def baseconv(self, a: np.ndarray, src: Sequence[int], dst: Sequence[int]) -> np.ndarray:
        """ApproxSwitchCRTBasis from prime set `src` to `dst` (indices into self.mods); `a` is in
        COEFFICIENT format [len(src)][N]."""
        key = (tuple(src), tuple(dst))
        tab = self._modup_cache.get(key)
        if tab is None:
            qs = [self.mods[i] for i in src]
            ps = [self.mods[j] for j in dst]
            Q = _prod(qs)
            hatinv = [pow((Q // qi) % qi, -1, qi) for qi in qs]
            hatmod = [((Q // qi) % pj) for qi in qs for pj in ps]
            tab = (_u64arr(qs), _u64arr(ps), _u64arr(hatinv), _u64arr(hatmod))
            self._modup_cache[key] = tab
        qs, ps, hatinv, hatmod = tab
        out = np.empty((len(dst), self.N), dtype=np.uint64)
        lib().ref_baseconv(_p(out), _p(a), len(src), _p(qs), len(dst), _p(ps), self.N, _p(hatinv), _p(hatmod))
        return out

    # ---- Galois index (upstream utils: FindAutomorphismIndex2nComplex)
    def galois(self, r: int) -> int:
        m = 2 * self.N
        if r == 0:
            return 1
        g = 5 if r > 0 else pow(5, -1, m)
        return pow(g, abs(r), m)

    # ------------------------------------------------------------------ client side (boundary)
    def sample_ternary(self, rng) -> np.ndarray:
        return rng.integers(-1, 2, size=self.N).astype(np.int64)

    def sample_gauss(self, rng) -> np.ndarray:
        return np.rint(rng.normal(0.0, 3.19, size=self.N)).astype(np.int64)

    def small_to_rns(self, v: np.ndarray, idx: Sequence[int], ntt: bool = True) -> np.ndarray:
        out = np.empty((len(idx), self.N), dtype=np.uint64)
        for r, i in enumerate(idx):
            q = self.mods[i]
            out[r] = np.where(v < 0, v + q, v).astype(np.uint64)
        if ntt:
            self.ntt(out, idx)
        return out

    def sample_uniform(self, rng, idx: Sequence[int]) -> np.ndarray:
        out = np.empty((len(idx), self.N), dtype=np.uint64)
        for r, i in enumerate(idx):
            out[r] = rng.integers(0, self.mods[i], size=self.N, dtype=np.uint64)
        return out

    def keygen(self, seed: int = 1):
        """KeyGen (main.cpp:184): sk ternary; pk = (-a*s + e, a) over Q."""
        rng = np.random.default_rng(seed)
        self.rng = rng
        allidx = list(range(self.L + self.K))
        qidx = list(range(self.L))
        self.sk_small = self.sample_ternary(rng)
        self.sk = self.small_to_rns(self.sk_small, allidx)          # over QP, eval form
        a = self.sample_uniform(rng, qidx)
        e = self.small_to_rns(self.sample_gauss(rng), qidx)
        mods = self.qmods(self.L)
        b = self.padd(self.psub(np.zeros_like(a), self.pmul(a, self.sk[:self.L], mods), mods), e, mods)
        self.pk = (b, a)

    def _ks_keygen(self, s_old: np.ndarray, s_new: np.ndarray):
        """KeySwitchHYBRID::KeySwitchGenInternal (upstream keyswitch/keyswitch-hybrid.cpp):
        for digit j: b_j = -a_j*s_new + e_j + [limb in digit j]*(P mod q_i)*s_old ; over all L+K limbs.
        Returns (a[dnum][L+K][N], b[dnum][L+K][N])."""
        P = self.P
        allidx = list(range(self.L + self.K))
        A = np.empty((P.dnum, self.L + self.K, self.N), dtype=np.uint64)
        B = np.empty_like(A)
        for j in range(P.dnum):
            a = self.sample_uniform(self.rng, allidx)
            e = self.small_to_rns(self.sample_gauss(self.rng), allidx)
            b = self.padd(self.psub(np.zeros_like(a), self.pmul(a, s_new, self.modarr), self.modarr), e, self.modarr)
            lo, hi = j * P.alpha, min(self.L, (j + 1) * P.alpha)
            fac = [0] * (self.L + self.K)
            for i in range(lo, hi):
                fac[i] = self.Pprod % self.mods[i]
            b = self.padd(b, self.pmul_scalar(s_old, fac, self.modarr), self.modarr)
            A[j], B[j] = a, b
        return A, B

    def mult_keygen(self):
        """EvalMultKeyGen (main.cpp:188): switch s^2 -> s."""
        s2 = self.pmul(self.sk, self.sk, self.modarr)
        self.mult_key = self._ks_keygen(s2, self.sk)

    def rot_keygen(self, rotations: Sequence[int]):
        """EvalRotateKeyGen / EvalSumKeyGen (main.cpp:192-206): key for Galois element k switches
        s -> sigma_{k^-1}(s) (upstream base-leveledshe.cpp EvalAutomorphismKeyGen)."""
        m = 2 * self.N
        for r in rotations:
            k = self.galois(r)
            if k in self.rot_keys:
                continue
            kinv = pow(k, -1, m)
            s_perm = self.pautomorph(self.sk, kinv)
            self.rot_keys[k] = self._ks_keygen(self.sk, s_perm)

    # ---- encode / decode (upstream encoding/ckkspackedencoding.cpp; full packing, slots = N/2)
    def _slot_index(self):
        if not hasattr(self, "_slot_idx"):
            m = 2 * self.N
            e = 1
            idx = np.empty(self.N // 2, dtype=np.int64)
            for j in range(self.N // 2):
                idx[j] = (e - 1) // 2
                e = (e * 5) % m
            self._slot_idx = idx
            self._slot_idx_conj = self.N - 1 - idx   # (2N - e - 1)/2
            n = np.arange(self.N)
            self._zeta = np.exp(1j * np.pi * n / self.N)
        return self._slot_idx, self._slot_idx_conj, self._zeta

    def encode_coeffs(self, values: Sequence[float], deg: int = 1) -> np.ndarray:
        """slots -> integer coefficient vector round(m * scale^deg)."""
        idx, idxc, zeta = self._slot_index()
        z = np.zeros(self.N // 2, dtype=np.complex128)
        vals = np.asarray(values, dtype=np.float64)
        z[:len(vals)] = vals
        v = np.zeros(self.N, dtype=np.complex128)
        v[idx] = z
        v[idxc] = np.conj(z)
        c = np.fft.fft(v) / self.N
        mcoef = np.real(c * np.conj(zeta))
        return np.rint(mcoef * (self.scale ** deg)).astype(np.int64) if deg == 1 else \
            np.array([int(round(x * self.scale ** deg)) for x in mcoef], dtype=object)

    def encode(self, values: Sequence[float], nl: Optional[int] = None) -> np.ndarray:
        """MakeCKKSPackedPlaintext (openFHE_wrapper.cpp:75): EVALUATION format over nl Q limbs."""
        nl = self.L if nl is None else nl
        coeffs = self.encode_coeffs(values)
        return self.small_to_rns(coeffs, list(range(nl)))

    def encrypt(self, values: Sequence[float], rng=None) -> Ct:
        """Encrypt(pk, ptxt) (openFHE_wrapper.cpp:76): (v*b + e0 + m, v*a + e1)."""
        rng = self.rng if rng is None else rng
        qidx = list(range(self.L))
        mods = self.qmods(self.L)
        m = self.encode(values)
        v = self.small_to_rns(self.sample_ternary(rng), qidx)
        e0 = self.small_to_rns(self.sample_gauss(rng), qidx)
        e1 = self.small_to_rns(self.sample_gauss(rng), qidx)
        c0 = self.padd(self.padd(self.pmul(v, self.pk[0], mods), e0, mods), m, mods)
        c1 = self.padd(self.pmul(v, self.pk[1], mods), e1, mods)
        return Ct([c0, c1], 0, 1)

    def decrypt_poly(self, ct: Ct) -> List[int]:
        """Decrypt core: c0 + c1*s (+ c2*s^2), CRT-interpolated to centred big integers."""
        nl = ct.nl
        mods = self.qmods(nl)
        s = np.ascontiguousarray(self.sk[:nl])
        acc = ct.c[0].copy()
        spow = s
        for comp in ct.c[1:]:
            acc = self.padd(acc, self.pmul(comp, spow, mods), mods)
            spow = self.pmul(spow, s, mods)
        self.ntt(acc, list(range(nl)), inverse=True)
        qs = [self.mods[i] for i in range(nl)]
        if nl == 1:
            q = qs[0]
            v = acc[0].astype(object)
            return [int(x) - q if int(x) > q // 2 else int(x) for x in v]
        Q = _prod(qs)
        res = [0] * self.N
        for i, qi in enumerate(qs):
            Qi = Q // qi
            f = (Qi * pow(Qi % qi, -1, qi)) % Q
            col = acc[i]
            for j in range(self.N):
                res[j] += int(col[j]) * f
        return [(x % Q) - Q if (x % Q) > Q // 2 else (x % Q) for x in res]

    def decrypt(self, ct: Ct) -> np.ndarray:
        """Decrypt + decode to N/2 real slot values (no decryption-noise flooding; OpenFHE adds
        fresh Gaussian noise here, so decrypted doubles are compared with a tolerance)."""
        idx, _, zeta = self._slot_index()
        coeffs = self.decrypt_poly(ct)
        sc = self.scale ** ct.deg
        m = np.array([float(x) / sc for x in coeffs], dtype=np.float64)
        v = np.fft.ifft(m * zeta) * self.N
        return np.real(v[idx])

    # ------------------------------------------------------------------ evaluation: key switching
    def ks_precompute(self, c: np.ndarray) -> List[np.ndarray]:
        """KeySwitchHYBRID::EvalKeySwitchPrecomputeCore (upstream keyswitch-hybrid.cpp) --
        reached from EvalFastRotationPrecompute (sender_diag.cpp:22) and inside every
        relinearisation/rotation.  c: [nl][N] EVALUATION.  Returns the digits, each extended to
        the nl Q-limbs followed by the K P-limbs: list of [nl+K][N]."""
        P = self.P
        nl = c.shape[0]
        ndig = min(-(-nl // P.alpha), P.dnum)
        pidx = list(range(self.L, self.L + self.K))
        digits = []
        for j in range(ndig):
            lo = j * P.alpha
            hi = nl if j == ndig - 1 else (j + 1) * P.alpha
            src = list(range(lo, hi))
            dst = [i for i in range(nl) if i < lo or i >= hi] + pidx
            part = np.ascontiguousarray(c[lo:hi]).copy()
            self.ntt(part, src, inverse=True)
            compl = self.baseconv(part, src, dst)
            self.ntt(compl, dst)
            ext = np.empty((nl + self.K, self.N), dtype=np.uint64)
            ext[:lo] = compl[:lo]
            ext[lo:hi] = c[lo:hi]
            ext[hi:] = compl[lo:]
            digits.append(ext)
        return digits

    def ks_core_ext(self, digits: List[np.ndarray], key, nl: int):
        """EvalFastKeySwitchCoreExt: (c~0, c~1) = sum_j digit_j * (b_j, a_j) over nl Q-limbs + K
        P-limbs; P-limbs of the key sit at index L..L+K-1."""
        A, B = key
        sel = list(range(nl)) + list(range(self.L, self.L + self.K))
        mods = np.ascontiguousarray(self.modarr[sel])
        c0 = np.zeros((nl + self.K, self.N), dtype=np.uint64)
        c1 = np.zeros_like(c0)
        l = lib()
        for j, d in enumerate(digits):
            bj = np.ascontiguousarray(B[j][sel])
            aj = np.ascontiguousarray(A[j][sel])
            l.ref_mul_acc(_p(c0), _p(d), _p(bj), nl + self.K, self.N, _p(mods))
            l.ref_mul_acc(_p(c1), _p(d), _p(aj), nl + self.K, self.N, _p(mods))
        return c0, c1

    def mod_down(self, ct_ext: np.ndarray, nl: int) -> np.ndarray:
        """DCRTPoly::ApproxModDown (upstream dcrtpoly-impl.h): [nl+K][N] over Q_l*P -> [nl][N]:
        out_i = (c_i - NTT(conv_{P->q_i}(iNTT(c_P)))) * P^-1 mod q_i."""
        pidx = list(range(self.L, self.L + self.K))
        qidx = list(range(nl))
        partP = np.ascontiguousarray(ct_ext[nl:]).copy()
        self.ntt(partP, pidx, inverse=True)
        conv = self.baseconv(partP, pidx, qidx)
        self.ntt(conv, qidx)
        mods = self.qmods(nl)
        diff = self.psub(np.ascontiguousarray(ct_ext[:nl]), conv, mods)
        pinv = [pow(self.Pprod % self.mods[i], -1, self.mods[i]) for i in qidx]
        return self.pmul_scalar(diff, pinv, mods)

    def key_switch_core(self, c: np.ndarray, key):
        """KeySwitchCore = precompute + EvalFastKeySwitchCore (core-ext then ModDown of both)."""
        nl = c.shape[0]
        digits = self.ks_precompute(c)
        c0, c1 = self.ks_core_ext(digits, key, nl)
        return self.mod_down(c0, nl), self.mod_down(c1, nl)

    # ------------------------------------------------------------------ evaluation: rotations
    def fast_rotation_precompute(self, ct: Ct) -> List[np.ndarray]:
        """cc->EvalFastRotationPrecompute(ct) (sender_diag.cpp:22): digits of c1."""
        return self.ks_precompute(ct.c[1])

    def fast_rotation(self, ct: Ct, r: int, digits: List[np.ndarray]) -> Ct:
        """cc->EvalFastRotation(ct, r, 2N, digits) (sender_diag.cpp:25; upstream
        base-leveledshe.cpp): key-switch with the hoisted digits, add c0, then permute both."""
        if r == 0:
            return ct.clone()
        k = self.galois(r)
        nl = ct.nl
        c0e, c1e = self.ks_core_ext(digits, self.rot_keys[k], nl)
        b = self.mod_down(c0e, nl)
        a = self.mod_down(c1e, nl)
        b = self.padd(b, ct.c[0], self.qmods(nl))
        return Ct([self.pautomorph(b, k), self.pautomorph(a, k)], ct.level, ct.deg)

    def rotate(self, ct: Ct, r: int) -> Ct:
        """cc->EvalRotate(ct, r) (openFHE_wrapper.cpp:124) = EvalAutomorphism: KeySwitchInPlace
        then AutomorphismTransform on both components."""
        if r == 0:
            return ct.clone()
        return self.automorphism(ct, self.galois(r))

    def automorphism(self, ct: Ct, k: int) -> Ct:
        nl = ct.nl
        b, a = self.key_switch_core(ct.c[1], self.rot_keys[k])
        b = self.padd(b, ct.c[0], self.qmods(nl))
        return Ct([self.pautomorph(b, k), self.pautomorph(a, k)], ct.level, ct.deg)

    def eval_sum(self, ct: Ct, n: int) -> Ct:
        """cc->EvalSum(ct, batchSize) (sender_diag.cpp:47; upstream base-advancedshe.cpp
        EvalSum2nComplex): g = 5; ceil(log2 n) times { ct += Automorph_g(ct); g = g^2 mod 2N }."""
        m = 2 * self.N
        g = 5
        out = ct.clone()
        for _ in range(int(math.ceil(math.log2(n)))):
            out = self.add(out, self.automorphism(out, g))
            g = (g * g) % m
        return out

    # ------------------------------------------------------------------ evaluation: leveled ops
    def level_reduce(self, ct: Ct, levels: int) -> Ct:
        """LevelReduceInternalInPlace: drop the last `levels` limbs (no scaling)."""
        if levels == 0:
            return ct.clone()
        nl = ct.nl - levels
        return Ct([np.ascontiguousarray(x[:nl]).copy() for x in ct.c], ct.level + levels, ct.deg)

    def _align(self, a: Ct, b: Ct):
        """FIXEDMANUAL AdjustLevelsInPlace: drop limbs of the fresher operand."""
        if a.nl > b.nl:
            a = self.level_reduce(a, a.nl - b.nl)
        elif b.nl > a.nl:
            b = self.level_reduce(b, b.nl - a.nl)
        return a, b

    def add(self, a: Ct, b: Ct) -> Ct:
        """EvalAdd / EvalAddInPlace(ct, ct) (sender_diag.cpp:76)."""
        a, b = self._align(a, b)
        if a.deg != b.deg:
            raise ValueError("EvalAdd: noiseScaleDeg mismatch")
        mods = self.qmods(a.nl)
        n = max(len(a.c), len(b.c))
        out = []
        for i in range(n):
            if i < len(a.c) and i < len(b.c):
                out.append(self.padd(a.c[i], b.c[i], mods))
            else:
                out.append((a.c[i] if i < len(a.c) else b.c[i]).copy())
        return Ct(out, a.level, a.deg)

    def sub(self, a: Ct, b: Ct) -> Ct:
        a, b = self._align(a, b)
        if a.deg != b.deg:
            raise ValueError("EvalSub: noiseScaleDeg mismatch")
        mods = self.qmods(a.nl)
        return Ct([self.psub(x, y, mods) for x, y in zip(a.c, b.c)], a.level, a.deg)

    def add_many(self, cts: List[Ct]) -> Ct:
        """EvalAddManyInPlace (sender_diag.cpp:46; upstream base-advancedshe.cpp): pairwise tree
        for j = 1,2,4..: for i = 0; i+j < n; i += 2j: ct[i] += ct[i+j]."""
        v = [c.clone() for c in cts]
        j = 1
        while j < len(v):
            i = 0
            while i + j < len(v):
                v[i] = self.add(v[i], v[i + j])
                i += 2 * j
            j *= 2
        return v[0]

    def _const_for_mult(self, c: float, nl: int) -> List[int]:
        """LeveledSHECKKSRNS::GetElementForEvalMult: (int128)(c*2^45 + 0.5), per-limb residues."""
        large = int(c * self.scale + 0.5)          # C cast truncates toward zero; so does int()
        return [large % self.mods[i] for i in range(nl)]   # python % is already non-negative

    def _const_for_add(self, c: float, nl: int, deg: int) -> List[int]:
        """GetElementForEvalAddOrSub: (uint64)(c*2^45 + 0.5) * ((uint64)(2^45+0.5))^(deg-1), c >= 0."""
        assert c >= 0
        sc = int(c * self.scale + 0.5)
        isf = int(self.scale + 0.5)
        out = []
        for i in range(nl):
            q = self.mods[i]
            v = sc % q
            for _ in range(1, deg):
                v = (v * (isf % q)) % q
            out.append(v)
        return out

    def mult_const(self, ct: Ct, c: float) -> Ct:
        """cc->EvalMult(ct, double): every component times the per-limb integer; deg += 1."""
        f = self._const_for_mult(c, ct.nl)
        mods = self.qmods(ct.nl)
        return Ct([self.pmul_scalar(x, f, mods) for x in ct.c], ct.level, ct.deg + 1)

    def add_const(self, ct: Ct, c: float) -> Ct:
        """cc->EvalAdd(ct, double) / EvalAddInPlace(ct, double) (openFHE_wrapper.cpp:182): the
        CryptoContext front end routes negative constants to EvalSub(ct, -c)."""
        mods = self.qmods(ct.nl)
        out = ct.clone()
        if c >= 0:
            out.c[0] = self.padd_scalar(ct.c[0], self._const_for_add(c, ct.nl, ct.deg), mods)
        else:
            f = self._const_for_add(-c, ct.nl, ct.deg)
            neg = [(self.mods[i] - f[i]) % self.mods[i] for i in range(ct.nl)]
            out.c[0] = self.padd_scalar(ct.c[0], neg, mods)
        return out

    def mult_norelin(self, a: Ct, b: Ct) -> Ct:
        """cc->EvalMultNoRelin (sender_diag.cpp:93): EvalMultCore on 2-component inputs."""
        a, b = self._align(a, b)
        mods = self.qmods(a.nl)
        assert len(a.c) == 2 and len(b.c) == 2
        c0 = self.pmul(a.c[0], b.c[0], mods)
        c1 = self.padd(self.pmul(a.c[0], b.c[1], mods), self.pmul(a.c[1], b.c[0], mods), mods)
        c2 = self.pmul(a.c[1], b.c[1], mods)
        return Ct([c0, c1, c2], a.level, a.deg + b.deg)

    def relinearize(self, ct: Ct) -> Ct:
        """cc->RelinearizeInPlace (sender_diag.cpp:79): key-switch c2 with the s^2->s key; no-op
        on a 2-component ciphertext."""
        if len(ct.c) == 2:
            return ct.clone()
        assert len(ct.c) == 3
        mods = self.qmods(ct.nl)
        b, a = self.key_switch_core(ct.c[2], self.mult_key)
        return Ct([self.padd(ct.c[0], b, mods), self.padd(ct.c[1], a, mods)], ct.level, ct.deg)

    def mult(self, a: Ct, b: Ct) -> Ct:
        """cc->EvalMult(ct, ct): EvalMultCore + relinearisation, no rescale."""
        return self.relinearize(self.mult_norelin(a, b))

    def mult_plain(self, ct: Ct, pt: np.ndarray) -> Ct:
        """EvalMult(ct, Plaintext) (openFHE_wrapper.cpp:235,244,289): pt = encode() output at level 0; a ciphertext at a
        lower level uses the plaintext's first nl limbs (LevelReduce of the plaintext); noiseScaleDeg adds up."""
        nl = ct.nl
        mods = self.qmods(nl)
        return Ct([self.pmul(c, np.ascontiguousarray(pt[:nl]), mods) for c in ct.c], ct.level, ct.deg + 1)

    def inner_product(self, a: Ct, b: Ct, n: int) -> Ct:
        """EvalInnerProduct (sender_base.cpp:93): EvalSum(EvalMult(a, b), n) (upstream base-advancedshe.cpp)."""
        return self.eval_sum(self.mult(a, b), n)

    def square(self, a: Ct) -> Ct:
        """cc->EvalSquare: same residues as EvalMult(a, a)."""
        return self.mult(a, a)

    def rescale(self, ct: Ct) -> Ct:
        """cc->RescaleInPlace / ModReduceInPlace (sender_diag.cpp:80) ->
        DCRTPoly::DropLastElementAndScale (upstream dcrtpoly-impl.h)."""
        nl = ct.nl
        ql = self.mods[nl - 1]
        mods = self.qmods(nl - 1)
        qlinv = [pow(ql % self.mods[i], -1, self.mods[i]) for i in range(nl - 1)]
        neg_qlinv = [(self.mods[i] - qlinv[i]) % self.mods[i] for i in range(nl - 1)]
        out = []
        l = lib()
        for comp in ct.c:
            last = np.ascontiguousarray(comp[nl - 1:nl]).copy()
            self.ntt(last, [nl - 1], inverse=True)
            tmp = np.empty((nl - 1, self.N), dtype=np.uint64)
            for i in range(nl - 1):
                l.ref_switch_modulus(_p(tmp[i]), _p(last[0]), self.N, ql, self.mods[i])
            tmp = self.pmul_scalar(tmp, neg_qlinv, mods)
            self.ntt(tmp, list(range(nl - 1)))
            head = self.pmul_scalar(np.ascontiguousarray(comp[:nl - 1]), qlinv, mods)
            out.append(self.padd(head, tmp, mods))
        return Ct(out, ct.level + 1, ct.deg - 1)

    # ------------------------------------------------------------------ HyDia (approach 5) flow
    def chebyshev_compare(self, ct: Ct, delta: float = 0.44, sign_depth: int = 10) -> Ct:
        return chebyshev_compare(self, ct, delta, sign_depth)


# ----------------------------------------------------------------------------- PS evaluators
# Written against an abstract evaluator `ev` (Oracle above, or PlainEval below for the
# plaintext-math golden check) exposing: add, sub, mult, square, mult_const, add_const, rescale,
# level_reduce, and ciphertext objects with .level.

_PREC = 2.0 ** -20


def _not_one(v: float) -> bool:
    """IsNotEqualOne (upstream math/chebyshev / polynomials): |v - 1| > 2^-20."""
    return (1 - _PREC > v) or (1 + _PREC < v)


def degree(c: Sequence[float]) -> int:
    """Degree(): index of the last non-zero coefficient."""
    d = len(c) - 1
    while d > 0 and c[d] == 0:
        d -= 1
    return d


def _populate_ps(upper: int = 2204) -> List[int]:
    """PopulateParameterPS (upstream ckksrns-utils.cpp): m for each degree n (index n-1)."""
    m = [0] * upper
    def fill(a, b, v):
        for i in range(a, min(b, upper)):
            m[i] = v
    fill(0, 2, 1); fill(2, 11, 2); fill(11, 13, 3); fill(13, 17, 2); fill(17, 55, 3)
    fill(55, 59, 4); fill(59, 76, 3); fill(76, 239, 4); fill(239, 247, 5); fill(247, 284, 4)
    fill(284, 991, 5); fill(991, 1021, 6); fill(1021, 1078, 5); fill(1078, 2204, 6)
    return m


_MLIST = _populate_ps()


def compute_degrees_ps(n: int):
    """ComputeDegreesPS: (k, m) with k = floor(n / (2^m - 1)) + 1."""
    m = _MLIST[n - 1]
    return n // ((1 << m) - 1) + 1, m


def cheb_coefficients(func: Callable[[float], float], a: float, b: float, deg: int) -> List[float]:
    """EvalChebyshevCoefficients (upstream core/lib/math/chebyshev.cpp), same association order."""
    n = deg + 1
    b_minus_a = 0.5 * (b - a)
    b_plus_a = 0.5 * (b + a)
    pi_by_deg = math.pi / float(n)
    fp = [func(math.cos(pi_by_deg * (i + 0.5)) * b_minus_a + b_plus_a) for i in range(n)]
    mult = 2.0 / float(n)
    out = []
    for i in range(n):
        s = 0.0
        for j in range(n):
            s += fp[j] * math.cos(pi_by_deg * i * (j + 0.5))
        out.append(s * mult)
    return out


def long_div_cheb(f: List[float], g: List[float]):
    """LongDivisionChebyshev (upstream ckksrns-utils.cpp): f = q*g + r in the Chebyshev basis,
    q returned with q[0] doubled ("[c0] in the last spot, not [c0/2]")."""
    n, k = degree(f), degree(g)
    assert n == len(f) - 1 and k == len(g) - 1
    r = list(f)
    if n - k >= 0:
        q = [0.0] * (n - k + 1)
        while n - k > 0:
            q[n - k] = 2 * r[-1]
            if _not_one(g[k]):
                q[n - k] /= g[-1]
            d = [0.0] * (n + 1)
            if k == n - k:
                d[0] = 2 * g[n - k]
                for i in range(1, 2 * k + 1):
                    d[i] = g[abs(n - k - i)]
            elif k > n - k:
                d[0] = 2 * g[n - k]
                for i in range(1, k - (n - k) + 1):
                    d[i] = g[abs(n - k - i)] + g[n - k + i]
                for i in range(k - (n - k) + 1, n + 1):
                    d[i] = g[abs(i - n + k)]
            else:
                d[n - k] = g[0]
                for i in range(n - 2 * k, n + 1):
                    if i != n - k:
                        d[i] = g[abs(i - n + k)]
            if _not_one(r[-1]):
                d = [x * r[-1] for x in d]
            if _not_one(g[-1]):
                d = [x / g[-1] for x in d]
            r = [x - y for x, y in zip(r, d)]
            if len(r) > 1:
                n = degree(r)
                r = r[:n + 1]
        if n == k:
            q[0] = r[-1]
            if _not_one(g[-1]):
                q[0] /= g[-1]
            d = list(g)
            if _not_one(r[-1]):
                d = [x * r[-1] for x in d]
            if _not_one(g[-1]):
                d = [x / g[-1] for x in d]
            r = [x - y for x, y in zip(r, d)]
            if len(r) > 1:
                n = degree(r)
                r = r[:n + 1]
        q[0] *= 2
    else:
        q = [0.0]
        r = list(f)
    return q, r


def long_div_poly(f: List[float], g: List[float]):
    """LongDivisionPoly (upstream ckksrns-utils.cpp): power-basis f = q*g + r."""
    n, k = degree(f), degree(g)
    assert n == len(f) - 1 and k == len(g) - 1
    r = list(f)
    if n - k >= 0:
        q = [0.0] * (n - k + 1)
        while n - k >= 0:
            d = [0.0] * (n - k) + list(g)
            q[n - k] = r[-1]
            if _not_one(g[-1]):
                q[n - k] /= g[-1]
            d = [x * q[n - k] for x in d]
            r = [x - y for x, y in zip(r, d)]
            if len(r) > 1:
                n = degree(r)
                r = r[:n + 1]
            else:
                break
    else:
        q = [0.0]
    return q, r


def _resize(v: List[float], n: int) -> List[float]:
    return list(v[:n]) + [0.0] * max(0, n - len(v))


def linear_wsum(ev, cts, weights):
    """EvalLinearWSumMutable (upstream ckksrns-advancedshe.cpp): sum_i w_i*ct_i, then ModReduce."""
    acc = ev.mult_const(cts[0], weights[0])
    for c, w in zip(cts[1:], weights[1:]):
        acc = ev.add(acc, ev.mult_const(c, w))
    return ev.rescale(acc)


def _ps_split(coeffs: List[float], k: int, m: int, longdiv):
    """Shared by the (Inner)Eval{Chebyshev,Poly}PS routines: divide by the degree k*2^(m-1) basis
    element, then r - basis_{k(2^(m-1)-1)} by q.  Returns (q, c, s2)."""
    k2m2k = k * (1 << (m - 1)) - k
    tkm = [0.0] * (k2m2k + k) + [1.0]
    q, r = longdiv(coeffs, tkm)
    r2 = list(r)
    if k2m2k - degree(r) <= 0:
        r2[k2m2k] -= 1
        r2 = r2[:degree(r2) + 1]
    else:
        r2 = _resize(r2, k2m2k + 1)
        r2[-1] = -1.0
    c, s = longdiv(r2, q)
    s2 = _resize(s, k2m2k + 1)
    s2[-1] = 1.0
    return q, c, s2


def _eval_c(ev, basis, c, half: bool):
    """cu = c(u) on the baby-step basis; `half`: Chebyshev free term is c0/2."""
    dc = degree(c)
    if dc < 1:
        return None
    if dc == 1:
        if c[1] != 1:
            cu = ev.rescale(ev.mult_const(basis[0], c[1]))
        else:
            cu = basis[0]
    else:
        cu = linear_wsum(ev, [basis[i] for i in range(dc)], [c[i + 1] for i in range(dc)])
    return ev.add_const(cu, c[0] / 2 if half else c[0])


def _inner_cheb_ps(ev, coeffs, k, m, T, T2):
    """AdvancedSHECKKSRNS::InnerEvalChebyshevPS (upstream ckksrns-advancedshe.cpp)."""
    q, c, s2 = _ps_split(coeffs, k, m, long_div_cheb)
    cu = _eval_c(ev, T, c, True)
    if degree(q) > k:
        qu = _inner_cheb_ps(ev, q, k, m - 1, T, T2)
    else:
        qcopy = _resize(q, k)
        dq = degree(qcopy)
        ssum = T[k - 1]
        i = 0
        while i < math.log2(q[-1]):
            ssum = ev.add(ssum, ssum)
            i += 1
        if dq > 0:
            qu = linear_wsum(ev, [T[i] for i in range(dq)], [q[i + 1] for i in range(dq)])
            qu = ev.add(qu, ssum)
        else:
            qu = ssum
        qu = ev.add_const(qu, q[0] / 2)
    if degree(s2) > k:
        su = _inner_cheb_ps(ev, s2, k, m - 1, T, T2)
    else:
        scopy = _resize(s2, k)
        ds = degree(scopy)
        if ds > 0:
            su = linear_wsum(ev, [T[i] for i in range(ds)], [s2[i + 1] for i in range(ds)])
            su = ev.add(su, T[k - 1])
        else:
            su = T[k - 1]
        su = ev.add_const(su, s2[0] / 2)
        su = ev.level_reduce(su, 1)
    if cu is not None:
        res = ev.add(T2[m - 1], cu)
    else:
        res = ev.add_const(T2[m - 1], c[0] / 2)
    res = ev.rescale(ev.mult(res, qu))
    return ev.add(res, su)


def eval_chebyshev_series_ps(ev, x, coefficients: List[float], a: float, b: float):
    """AdvancedSHECKKSRNS::EvalChebyshevSeriesPS (upstream ckksrns-advancedshe.cpp), reached via
    cc->EvalChebyshevFunction (openFHE_wrapper.cpp:174).  FIXEDMANUAL branch."""
    n = degree(coefficients)
    f2 = list(coefficients)
    if coefficients[-1] == 0:
        f2 = f2[:n + 1]
    k, m = compute_degrees_ps(n)
    T = [None] * k
    if (a - round(a) < 1e-10) and (b - round(b) < 1e-10) and round(a) == -1 and round(b) == 1:
        T[0] = x
    else:
        alpha = 2 / (b - a)
        beta = 2 * a / (b - a)
        T[0] = ev.add_const(ev.rescale(ev.mult_const(x, alpha)), -1.0 - beta)
    y = T[0]
    for i in range(2, k + 1):
        if not (i & (i - 1)) or i % 2 == 0:
            sq = ev.square(T[i // 2 - 1])
            T[i - 1] = ev.add_const(ev.rescale(ev.add(sq, sq)), -1.0)
        else:
            prod = ev.mult(T[i // 2 - 1], T[i // 2])
            T[i - 1] = ev.sub(ev.rescale(ev.add(prod, prod)), y)
    for i in range(1, k):
        T[i - 1] = ev.level_reduce(T[i - 1], T[k - 1].level - T[i - 1].level)
    T2 = [None] * m
    T2[0] = T[-1]
    for i in range(1, m):
        sq = ev.square(T2[i - 1])
        T2[i] = ev.add_const(ev.rescale(ev.add(sq, sq)), -1.0)
    T2km1 = T2[0]
    for i in range(1, m):
        prod = ev.mult(T2km1, T2[i])
        T2km1 = ev.sub(ev.rescale(ev.add(prod, prod)), T2[0])
    k2m2k = k * (1 << (m - 1)) - k
    f2 = _resize(f2, 2 * k2m2k + k + 1)
    f2[-1] = 1.0
    q, c, s2 = _ps_split(f2, k, m, long_div_cheb)
    cu = _eval_c(ev, T, c, True)
    if cu is not None:
        cu = ev.level_reduce(cu, T2[m - 1].level - cu.level)
    if degree(q) > k:
        qu = _inner_cheb_ps(ev, q, k, m - 1, T, T2)
    else:
        qcopy = _resize(q, k)
        dq = degree(qcopy)
        if dq > 0:
            qu = linear_wsum(ev, [T[i] for i in range(dq)], [q[i + 1] for i in range(dq)])
            qu = ev.add(qu, ev.add(T[k - 1], T[k - 1]))
        else:
            qu = T[k - 1]
            i = 1
            while i < q[-1]:
                qu = ev.add(qu, T[k - 1])
                i += 1
        qu = ev.add_const(qu, q[0] / 2)
    if degree(s2) > k:
        su = _inner_cheb_ps(ev, s2, k, m - 1, T, T2)
    else:
        scopy = _resize(s2, k)
        ds = degree(scopy)
        if ds > 0:
            su = linear_wsum(ev, [T[i] for i in range(ds)], [s2[i + 1] for i in range(ds)])
            su = ev.add(su, T[k - 1])
        else:
            su = T[k - 1]
        su = ev.add_const(su, s2[0] / 2)
    if cu is not None:
        res = ev.add(T2[m - 1], cu)
    else:
        res = ev.add_const(T2[m - 1], c[0] / 2)
    res = ev.rescale(ev.mult(res, qu))
    res = ev.add(res, su)
    return ev.sub(res, T2km1)


def _inner_poly_ps(ev, coeffs, k, m, powers, powers2):
    """AdvancedSHECKKSRNS::InnerEvalPolyPS."""
    q, c, s2 = _ps_split(coeffs, k, m, long_div_poly)
    cu = _eval_c(ev, powers, c, False)
    qu = _poly_leaf_or_recurse(ev, q, q, k, m, powers, powers2)
    su = _poly_leaf_or_recurse(ev, s2, s2, k, m, powers, powers2)
    if cu is not None:
        res = ev.add(powers2[m - 1], cu)
    else:
        res = ev.add_const(powers2[m - 1], c[0])
    res = ev.rescale(ev.mult(res, qu))
    return ev.add(res, su)


def _poly_leaf_or_recurse(ev, v, _unused, k, m, powers, powers2):
    if degree(v) > k:
        return _inner_poly_ps(ev, v, k, m - 1, powers, powers2)
    vc = _resize(v, k)
    dv = degree(vc)
    if dv > 0:
        out = linear_wsum(ev, [powers[i] for i in range(dv)], [v[i + 1] for i in range(dv)])
        out = ev.add(out, powers[k - 1])
    else:
        out = powers[k - 1]
    return ev.add_const(out, v[0])


def eval_poly_ps(ev, x, coefficients: List[float]):
    """AdvancedSHECKKSRNS::EvalPolyPS (upstream ckksrns-advancedshe.cpp), reached via cc->EvalPoly
    (openFHE_wrapper.cpp:179) for degree >= 5.  FIXEDMANUAL branch."""
    n = degree(coefficients)
    f2 = list(coefficients)
    if coefficients[-1] == 0:
        f2 = f2[:n + 1]
    k, m = compute_degrees_ps(n)
    indices = [0] * k
    for i in range(k, 0, -1):
        if not (i & (i - 1)):
            indices[i - 1] = 1
        else:
            indices[i - 1] = 1
            p2 = 1 << int(math.floor(math.log2(i)))
            rem = i % p2
            indices[rem - 1] = 1
            while rem & (rem - 1):
                p2 = 1 << int(math.floor(math.log2(rem)))
                rem = rem % p2
                indices[rem - 1] = 1
    powers = [None] * k
    powers[0] = x
    for i in range(2, k + 1):
        if not (i & (i - 1)):
            powers[i - 1] = ev.rescale(ev.square(powers[i // 2 - 1]))
        elif indices[i - 1] == 1:
            p2 = 1 << int(math.floor(math.log2(i)))
            rem = i % p2
            powers[rem - 1] = ev.level_reduce(powers[rem - 1], powers[p2 - 1].level - powers[rem - 1].level)
            powers[i - 1] = ev.rescale(ev.mult(powers[p2 - 1], powers[rem - 1]))
    for i in range(1, k):
        if indices[i - 1] == 1:
            powers[i - 1] = ev.level_reduce(powers[i - 1], powers[k - 1].level - powers[i - 1].level)
    powers2 = [None] * m
    powers2[0] = powers[-1]
    for i in range(1, m):
        powers2[i] = ev.rescale(ev.square(powers2[i - 1]))
    power2km1 = powers2[0]
    for i in range(1, m):
        power2km1 = ev.rescale(ev.mult(power2km1, powers2[i]))
    k2m2k = k * (1 << (m - 1)) - k
    f2 = _resize(f2, 2 * k2m2k + k + 1)
    f2[-1] = 1.0
    q, c, s2 = _ps_split(f2, k, m, long_div_poly)
    cu = _eval_c(ev, powers, c, False)
    qu = _poly_leaf_or_recurse(ev, q, q, k, m, powers, powers2)
    su = _poly_leaf_or_recurse(ev, s2, s2, k, m, powers, powers2)
    if cu is not None:
        res = ev.add(powers2[m - 1], cu)
    else:
        res = ev.add_const(powers2[m - 1], c[0])
    res = ev.rescale(ev.mult(res, qu))
    res = ev.add(res, su)
    return ev.sub(res, power2km1)


DEPTH_TO_DEGREE = [-1, -1, -1, 5, 13, 27, 59, 119, 247, 495, 1007, 2031]
F4_COEFS = [0.0, 315.0 / 128.0, 0.0, -420.0 / 128.0, 0.0, 378.0 / 128.0, 0.0, -180.0 / 128.0, 0.0, 35.0 / 128.0]


def chebyshev_compare(ev, ct, delta: float = 0.44, sign_depth: int = 10):
    """OpenFHEWrapper::chebyshevCompare (/root/reference/src/openFHE_wrapper.cpp:143-185)."""
    if sign_depth < 7 or sign_depth > 15:
        return ct
    deg = DEPTH_TO_DEGREE[sign_depth - 4]
    coeffs = cheb_coefficients(lambda x: 1.0 if x >= delta else -1.0, -1.0, 1.0, deg)
    ct = eval_chebyshev_series_ps(ev, ct, coeffs, -1.0, 1.0)
    ct = eval_poly_ps(ev, ct, F4_COEFS)
    return ev.add_const(ct, 1.0)


class PlainCt:
    def __init__(self, v, level=0, deg=1):
        self.v, self.level, self.deg = v, level, deg


class PlainEval:
    """Plain-double stand-in for the evaluator: same op/level discipline, values are the slot
    values themselves.  Used to pin the PS restatement against signApprox.csv without keys."""
    def __init__(self, scale_bits=45):
        self.scale = float(2 ** scale_bits)
    def add(self, a, b):
        assert a.deg == b.deg
        return PlainCt(a.v + b.v, max(a.level, b.level), a.deg)
    def sub(self, a, b):
        assert a.deg == b.deg
        return PlainCt(a.v - b.v, max(a.level, b.level), a.deg)
    def mult(self, a, b): return PlainCt(a.v * b.v, max(a.level, b.level), a.deg + b.deg)
    def square(self, a): return self.mult(a, a)
    def mult_const(self, a, c): return PlainCt(a.v * (int(c * self.scale + 0.5) / self.scale), a.level, a.deg + 1)
    def add_const(self, a, c): return PlainCt(a.v + c, a.level, a.deg)
    def rescale(self, a): return PlainCt(a.v, a.level + 1, a.deg - 1)
    def level_reduce(self, a, n): return PlainCt(a.v, a.level + n, a.deg)


# ----------------------------------------------------------------------------- HyDia approach 5
VECTOR_DIM = 512
MATCH_THRESHOLD = 0.44
COMP_DEPTH = 10


def plaintext_normalize(v: np.ndarray) -> np.ndarray:
    n = np.linalg.norm(v, axis=-1, keepdims=True)
    return v / np.where(n == 0, 1.0, n)


def diagonal_pack(db: np.ndarray, batch: int) -> np.ndarray:
    """DiagonalEnroller::serializeDB packing (/root/reference/src/enroller/enroller_diag.cpp:12-53 via
    splitIntoSquareMatrices :57-86, preprocessToDiagonalForm :99-115, concatenateRows :118-156),
    written as the explicit loops of the reference."""
    dim = VECTOR_DIM
    dbn = plaintext_normalize(db.astype(np.float64))
    nrows = int(math.ceil(db.shape[0] / dim)) * dim
    mats = []
    for i in range(0, nrows, dim):
        sq = np.zeros((dim, dim))
        hi = min(i + dim, db.shape[0])
        if hi > i:
            sq[:hi - i] = dbn[i:hi]
        mats.append(sq)
    diag_mats = []
    for mtx in mats:
        d = np.zeros((dim, dim))
        for i in range(dim):
            cols = (np.arange(dim) + i) % dim
            d[i] = mtx[np.arange(dim), cols]
        diag_mats.append(d)
    per = batch // dim
    nout = int(math.ceil(len(diag_mats) / per)) * dim
    out = np.zeros((nout, batch))
    for i in range(nout):
        for j in range(per):
            m = (i // dim) * per + j
            if m < len(diag_mats):
                out[i, j * dim:(j + 1) * dim] = diag_mats[m][i % dim]
    return out


def encrypt_query_diag(orc: Oracle, query: np.ndarray) -> Ct:
    """DiagonalReceiver::encryptQuery (/root/reference/src/receiver/receiver_diag.cpp:13-26)."""
    q = plaintext_normalize(np.asarray(query, dtype=np.float64))
    return orc.encrypt(np.tile(q, orc.P.slots // VECTOR_DIM))


def diag_compute_similarity(orc: Oracle, query: Ct, db_cts: List[Ct], ndiag: int = VECTOR_DIM) -> List[Ct]:
    """DiagonalSender::computeSimilarity (/root/reference/src/sender/sender_diag.cpp:12-33) with
    computeSimilarityMatrix (:66-83); db_cts[g*ndiag + d] stands for serial/db_diagonal/index*.bin."""
    digits = orc.fast_rotation_precompute(query)
    rot = [query] + [orc.fast_rotation(query, i, digits) for i in range(1, ndiag)]
    scores = []
    for g in range(len(db_cts) // ndiag):
        acc = orc.mult_norelin(rot[0], db_cts[g * ndiag])
        for i in range(1, ndiag):
            acc = orc.add(acc, orc.mult_norelin(rot[i], db_cts[g * ndiag + i]))
        scores.append(orc.rescale(orc.relinearize(acc)))
    return scores


def diag_index_scenario(orc: Oracle, query: Ct, db_cts: List[Ct], ndiag: int = VECTOR_DIM) -> List[Ct]:
    """DiagonalSender::indexScenario (sender_diag.cpp:52-63)."""
    return [chebyshev_compare(orc, s, MATCH_THRESHOLD, COMP_DEPTH) for s in diag_compute_similarity(orc, query, db_cts, ndiag)]


def diag_membership_scenario(orc: Oracle, query: Ct, db_cts: List[Ct], ndiag: int = VECTOR_DIM) -> Ct:
    """DiagonalSender::membershipScenario (sender_diag.cpp:35-50)."""
    s = diag_index_scenario(orc, query, db_cts, ndiag)
    return orc.eval_sum(orc.add_many(s), orc.P.slots)


# ---- approach 4 (HERS), SURVEY.md section 8 row R7 ---------------------------------------------------
def hers_pack(db: np.ndarray, batch: int) -> np.ndarray:
    """HersEnroller::serializeDB packing (/root/reference/src/enroller/enroller_hers.cpp:14-36 normalisation,
    :106-121 serializeDBThread), written as the reference's loops: out[matrix][index][k % batch] =
    normalize(db)[k][index] for k in [matrix*batch, (matrix+1)*batch)."""
    n, dim = db.shape
    dbn = plaintext_normalize(db.astype(np.float64))
    nmat = int(math.ceil(n / batch))
    out = np.zeros((nmat, dim, batch))
    for m in range(nmat):
        for index in range(dim):
            for k in range(m * batch, min((m + 1) * batch, n)):
                out[m, index, k % batch] = dbn[k][index]
    return out


def encrypt_query_hers(orc: Oracle, query: np.ndarray) -> List[Ct]:
    """HersReceiver::encryptQuery (/root/reference/src/receiver/receiver_hers.cpp:13-24) with encryptQueryThread
    (:56-61): one ciphertext per coordinate, every slot holding that coordinate of the normalised query."""
    q = plaintext_normalize(np.asarray(query, dtype=np.float64))
    return [orc.encrypt(np.full(orc.P.slots, v)) for v in q]


def hers_compute_similarity(orc: Oracle, query: List[Ct], db_cts: List[List[Ct]]) -> List[Ct]:
    """HersSender::computeSimilarity (/root/reference/src/sender/sender_hers.cpp:13-26) with computeSimilarityHelper
    (:61-84: EvalMultNoRelin, RelinearizeInPlace, RescaleInPlace per column, then the EvalAddInPlace chain);
    db_cts[m][i] stands for serial/db_hers/matrix{m}/index{i}.bin (:88-98)."""
    out = []
    for mat in db_cts:
        score = [orc.rescale(orc.relinearize(orc.mult_norelin(query[i], mat[i]))) for i in range(len(query))]
        for i in range(1, len(score)):
            score[0] = orc.add(score[0], score[i])
        out.append(score[0])
    return out


def hers_index_scenario(orc: Oracle, query: List[Ct], db_cts: List[List[Ct]]) -> List[Ct]:
    """HersSender::indexScenario (sender_hers.cpp:29-40)."""
    return [chebyshev_compare(orc, s, MATCH_THRESHOLD, COMP_DEPTH) for s in hers_compute_similarity(orc, query, db_cts)]


def hers_membership_scenario(orc: Oracle, query: List[Ct], db_cts: List[List[Ct]]) -> Ct:
    """HersSender::membershipScenario (sender_hers.cpp:43-57)."""
    return orc.eval_sum(orc.add_many(hers_index_scenario(orc, query, db_cts)), orc.P.slots)


# ---- OpenFHEWrapper helpers used by approaches 1 and 3 ---------------------
def binary_rotate(orc: Oracle, ct: Ct, factor: int) -> Ct:
    """OpenFHEWrapper::binaryRotate (/root/reference/src/openFHE_wrapper.cpp:103-128)."""
    batch = orc.P.slots
    needed = []
    while factor != 0:
        sign = factor // abs(factor)
        counter = int(pow(2, round(math.log2(abs(factor)))))
        cur = int(math.fmod(counter * sign, batch))
        if cur != 0:
            needed.append(counter * sign)
        factor -= counter * sign
    for r in needed:
        ct = orc.rotate(ct, r)
    return ct


def generate_merge_mask(orc: Oracle, dimension: int, segment: int) -> np.ndarray:
    """OpenFHEWrapper::generateMergeMask (openFHE_wrapper.cpp:253-271)."""
    batch = orc.P.slots
    mask = np.zeros(batch)
    if segment > batch // dimension:
        return orc.encode(mask)
    i = 0
    while i < batch:
        mask[i:i + segment] = 1.0
        i += dimension * segment
    return orc.encode(mask)


def merge_single_cipher(orc: Oracle, ct: Ct, dimension: int) -> Ct:
    """OpenFHEWrapper::mergeSingleCipher (openFHE_wrapper.cpp:221-249)."""
    output_size = orc.P.slots // dimension
    padding, rot = 1, dimension - 1
    i = 1
    while i < output_size:
        if i >= padding:
            ct = orc.rescale(orc.relinearize(orc.mult_plain(ct, generate_merge_mask(orc, dimension, i))))
            padding = i * dimension
        ct = orc.add(ct, binary_rotate(orc, ct, rot * i))
        i *= 2
    return orc.rescale(orc.relinearize(orc.mult_plain(ct, generate_merge_mask(orc, dimension, output_size))))


def merge_ciphers(orc: Oracle, cts: List[Ct], dimension: int) -> List[Ct]:
    """OpenFHEWrapper::mergeCiphers (openFHE_wrapper.cpp:189-217)."""
    batch = orc.P.slots
    per = batch // dimension
    needed = int(math.ceil(per * len(cts) / batch))
    cts = [merge_single_cipher(orc, c, dimension) for c in cts]
    merged: List[Optional[Ct]] = [None] * needed
    for i, c in enumerate(cts):
        oc, slot = (per * i) // batch, (per * i) % batch
        merged[oc] = c if slot == 0 else orc.add(merged[oc], binary_rotate(orc, c, -slot))
    return merged


def compress_ciphers(orc: Oracle, cts: List[Ct], dimension: int) -> List[Ct]:
    """OpenFHEWrapper::compressCiphers (openFHE_wrapper.cpp:273-312)."""
    batch = orc.P.slots
    needed = int(math.ceil(len(cts) / dimension))
    mask = np.zeros(batch)
    mask[::dimension] = 1.0
    pt = orc.encode(mask)
    cts = [orc.rescale(orc.relinearize(orc.mult_plain(c, pt))) for c in cts]
    out: List[Optional[Ct]] = [None] * needed
    for i, c in enumerate(cts):
        rot, slot = -(i % dimension), i // dimension
        out[slot] = c if rot == 0 else orc.add(out[slot], binary_rotate(orc, c, rot))
    return out


# ---- approach 1 (literature baseline), row R9 ----------------------------------------------------------
def base_pack(db: np.ndarray, batch: int) -> np.ndarray:
    """BaseEnroller::serializeDB packing (/root/reference/src/enroller/enroller_base.cpp:36-56), as the reference's loops."""
    n, dim = db.shape
    per = batch // dim
    dbn = plaintext_normalize(db.astype(np.float64))
    nb = int(math.ceil(n / per))
    out = np.zeros((nb, batch))
    for i in range(nb):
        for j in range(min(per, n - i * per)):
            out[i, j * dim:(j + 1) * dim] = dbn[j + i * per]
    return out


def encrypt_query_base(orc: Oracle, query: np.ndarray) -> List[Ct]:
    """BaseReceiver::encryptQuery (/root/reference/src/receiver/receiver_base.cpp:13-26)."""
    q = plaintext_normalize(np.asarray(query, dtype=np.float64))
    batch = np.zeros(orc.P.slots)
    for i in range(0, orc.P.slots, len(q)):
        batch[i:i + len(q)] = q
    return [orc.encrypt(batch)]


def base_compute_similarity(orc: Oracle, query: List[Ct], db_cts: List[Ct], dim: int = VECTOR_DIM) -> List[Ct]:
    """BaseSender::computeSimilarity (/root/reference/src/sender/sender_base.cpp:13-27) with computeSimilarityThread
    (:84-98: EvalInnerProduct, RelinearizeInPlace, RescaleInPlace) and OpenFHEWrapper::mergeCiphers."""
    sim = [orc.rescale(orc.relinearize(orc.inner_product(query[0], d, dim))) for d in db_cts]
    return merge_ciphers(orc, sim, dim)


# ---- approach 3 (Blind-Match), row R8 --------------------------------------------------------------------
def blind_pack(db: np.ndarray, batch: int, chunk: int) -> np.ndarray:
    """BlindEnroller::serializeDB packing (/root/reference/src/enroller/enroller_blind.cpp:40-90), as the reference's loops:
    out[matrix][index/chunk][i*chunk : (i+1)*chunk] = normalize(db)[i + matrix*chunksPerBatch][index : index+chunk]."""
    n, dim = db.shape
    per = batch // chunk
    dbn = plaintext_normalize(db.astype(np.float64))
    nmat = int(math.ceil(n / per))
    out = np.zeros((nmat, dim // chunk, batch))
    for m in range(nmat):
        remaining = n - m * per
        for index in range(0, dim, chunk):
            for i in range(min(per, remaining)):
                out[m, index // chunk, i * chunk:(i + 1) * chunk] = dbn[i + m * per][index:index + chunk]
    return out


def encrypt_query_blind(orc: Oracle, query: np.ndarray, chunk: int) -> List[Ct]:
    """BlindReceiver::encryptQuery (/root/reference/src/receiver/receiver_blind.cpp:13-27) with encryptQueryThread (:58-70)."""
    q = plaintext_normalize(np.asarray(query, dtype=np.float64))
    out = []
    for index in range(0, len(q), chunk):
        v = np.zeros(orc.P.slots)
        for i in range(0, orc.P.slots, chunk):
            v[i:i + chunk] = q[index:index + chunk]
        out.append(orc.encrypt(v))
    return out


def blind_compute_similarity(orc: Oracle, query: List[Ct], db_cts: List[List[Ct]], chunk: int) -> List[Ct]:
    """BlindSender::computeSimilarity (/root/reference/src/sender/sender_blind.cpp:43-57) with computeSimilarityMatrix
    (:59-83) and computeSimilaritySerial (:85-94); db_cts[m][c] stands for serial/db_blind/matrix{m}/batch{c*chunk}.bin."""
    scores = []
    for mat in db_cts:
        m = [orc.mult_norelin(query[i], mat[i]) for i in range(len(query))]
        for i in range(1, len(m)):
            m[0] = orc.add(m[0], m[i])
        acc = orc.rescale(orc.relinearize(m[0]))
        r = 1
        while r < chunk:
            acc = orc.add(acc, binary_rotate(orc, acc, r))
            r *= 2
        scores.append(acc)
    return compress_ciphers(orc, scores, chunk)


def blind_decrypt_index(orc: Oracle, cts: List[Ct], chunk: int) -> List[int]:
    """BlindReceiver::decryptIndex (receiver_blind.cpp:29-54)."""
    batch = orc.P.slots
    per = batch // chunk
    out = []
    for i, ct in enumerate(cts):
        v = orc.decrypt(ct)
        out.extend(int(i * batch + j // chunk + (j % chunk) * per) for j in np.nonzero(v >= 1.0)[0])
    return out


def decrypt_membership(orc: Oracle, ct: Ct) -> bool:
    """HersReceiver::decryptMembership (/root/reference/src/receiver/receiver_hers.cpp:26-35)."""
    return bool(orc.decrypt(ct)[0] >= 1.0)


def decrypt_index(orc: Oracle, cts: List[Ct]) -> List[int]:
    """HersReceiver::decryptIndex (receiver_hers.cpp:37-54)."""
    out = []
    for i, ct in enumerate(cts):
        v = orc.decrypt(ct)
        out.extend(int(j + i * orc.P.slots) for j in np.nonzero(v >= 1.0)[0])
    return out


def hydia_rotation_set(batch: int) -> List[int]:
    """Rotation keys generated at /root/reference/src/main.cpp:195-206."""
    rot = list(range(1, VECTOR_DIM))
    i = VECTOR_DIM
    while i < batch:
        rot.append(i); i *= 2
    i = 1
    while i < batch:
        rot.append(-i); i *= 2
    return rot


def read_dat(path: str):
    """Dataset format of /root/reference/src/main.cpp:56-57, 216-230: n, query[512], db[n][512]."""
    with open(path) as f:
        tok = f.read().split()
    n = int(tok[0])
    vals = np.array(tok[1:], dtype=np.float64)
    return n, vals[:VECTOR_DIM], vals[VECTOR_DIM:VECTOR_DIM + n * VECTOR_DIM].reshape(n, VECTOR_DIM)
