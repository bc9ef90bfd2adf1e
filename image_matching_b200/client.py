"""Synthetic client side over the C ABI (hx_client_*): KeyGen / EvalMultKeyGen / EvalRotateKeyGen
(src/main.cpp:184-206), DiagonalEnroller packing + encryption (src/enroller/enroller_diag.cpp),
DiagonalReceiver::encryptQuery (src/receiver/receiver_diag.cpp:13-26) and HersReceiver decisions
(src/receiver/receiver_hers.cpp:26-54). Used by smoke / bench / tests to produce inputs; the
Sender hot path never needs it."""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Sequence

import numpy as np

from . import _lib
from ._lib import F64P, I32P, check_int, check_ptr
from .config import VECTOR_DIM
from .engine import Ciphertext, CryptoContext


def hydia_rotation_set(batch_size: int) -> List[int]:
    """Rotation indices of src/main.cpp:195-206."""
    rot = list(range(1, VECTOR_DIM))
    i = VECTOR_DIM
    while i < batch_size:
        rot.append(i); i *= 2
    i = 1
    while i < batch_size:
        rot.append(-i); i *= 2
    return rot


def plaintext_normalize(v: np.ndarray) -> np.ndarray:
    """VectorUtils::plaintextNormalize (src/vector_utils.cpp)."""
    n = np.linalg.norm(v, axis=-1, keepdims=True)
    return v / np.where(n == 0, 1.0, n)


def diagonal_pack(db: np.ndarray, batch: int) -> np.ndarray:
    """DiagonalEnroller packing (enroller_diag.cpp:57-156): returns [G*512][batch] rows; row
    g*512+d, slot j*512+r = normalize(db)[(g*(batch/512)+j)*512 + r][(r+d) mod 512]."""
    dim = VECTOR_DIM
    n = db.shape[0]
    dbn = plaintext_normalize(db.astype(np.float64))
    nmat = int(math.ceil(n / dim))
    padded = np.zeros((nmat * dim, dim))
    padded[:n] = dbn
    mats = padded.reshape(nmat, dim, dim)
    r = np.arange(dim)
    # diagonals[m][d][r] = mats[m][r][(r+d) % dim]
    diags = np.stack([mats[:, r, (r + d) % dim] for d in range(dim)], axis=1)   # [nmat][dim(d)][dim(r)]
    per = batch // dim
    G = int(math.ceil(nmat / per))
    out = np.zeros((G * dim, batch))
    for g in range(G):
        for j in range(per):
            m = g * per + j
            if m < nmat:
                out[g * dim:(g + 1) * dim, j * dim:(j + 1) * dim] = diags[m]
    return out


def hers_pack(db: np.ndarray, batch: int) -> np.ndarray:
    """HersEnroller packing (enroller_hers.cpp:14-36, 106-121): [matrices][dim][batch]; matrix m, column j, slot k =
    normalize(db)[m*batch + k][j]; rows beyond the database stay zero."""
    n, dim = db.shape
    dbn = plaintext_normalize(db.astype(np.float64))
    nmat = int(math.ceil(n / batch))
    padded = np.zeros((nmat * batch, dim))
    padded[:n] = dbn
    return np.ascontiguousarray(padded.reshape(nmat, batch, dim).transpose(0, 2, 1))


def base_pack(db: np.ndarray, batch: int) -> np.ndarray:
    """BaseEnroller::serializeDB packing (enroller_base.cpp:36-56): [numBatches][batch]; row i = normalised vectors
    i*(batch/dim) .. (i+1)*(batch/dim)-1, one after the other."""
    n, dim = db.shape
    per = batch // dim
    dbn = plaintext_normalize(db.astype(np.float64))
    nb = int(math.ceil(n / per))
    padded = np.zeros((nb * per, dim))
    padded[:n] = dbn
    return np.ascontiguousarray(padded.reshape(nb, per * dim))


def blind_pack(db: np.ndarray, batch: int, chunk: int) -> np.ndarray:
    """BlindEnroller::serializeDB packing (enroller_blind.cpp:40-90): [matrices][dim/chunk][batch]; matrix m, chunk c,
    slots i*chunk..(i+1)*chunk-1 = normalize(db)[m*(batch/chunk) + i][c*chunk:(c+1)*chunk]."""
    n, dim = db.shape
    per = batch // chunk
    dbn = plaintext_normalize(db.astype(np.float64))
    nmat = int(math.ceil(n / per))
    padded = np.zeros((nmat * per, dim))
    padded[:n] = dbn
    x = padded.reshape(nmat, per, dim // chunk, chunk)           # [m][i][c][k]
    return np.ascontiguousarray(x.transpose(0, 2, 1, 3).reshape(nmat, dim // chunk, per * chunk))


class Client:
    def __init__(self, cc: CryptoContext, seed: int = 1):
        self.cc = cc
        self._lib = _lib.load()
        self._h = check_ptr(self._lib.hx_client_create(cc._h, seed))
        if getattr(cc, "encoder", None) is None:
            cc.encoder = self.MakeCKKSPackedPlaintext

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._lib.hx_client_destroy(h)

    def EvalMultKeyGen(self): check_int(self._lib.hx_client_gen_mult_key(self._h))
    def EvalSumKeyGen(self): check_int(self._lib.hx_client_gen_sum_keys(self._h))
    def EvalRotateKeyGen(self, rotations: Sequence[int]):
        r = np.array(list(rotations), dtype=np.int32)
        check_int(self._lib.hx_client_gen_rot_keys(self._h, r.ctypes.data_as(I32P), len(r)))

    def MakeCKKSPackedPlaintext(self, values) -> np.ndarray:
        v = np.ascontiguousarray(values, dtype=np.float64)
        out = np.empty((self.cc.L, self.cc.N), dtype=np.uint64)
        check_int(self._lib.hx_client_encode(self._h, v.ctypes.data_as(F64P), len(v), out.ctypes.data_as(_lib.U64P)))
        return out

    def encrypt(self, values) -> Ciphertext:
        v = np.ascontiguousarray(values, dtype=np.float64)
        return Ciphertext(self.cc, self._lib.hx_client_encrypt(self._h, v.ctypes.data_as(F64P), len(v)))

    def decrypt(self, ct: Ciphertext) -> np.ndarray:
        out = np.empty(self.cc.N // 2, dtype=np.float64)
        check_int(self._lib.hx_client_decrypt(self._h, ct._h, out.ctypes.data_as(F64P)))
        return out

    def secret(self) -> np.ndarray:
        out = np.empty(self.cc.N, dtype=np.int8)
        check_int(self._lib.hx_client_secret(self._h, out.ctypes.data_as(C.POINTER(C.c_int8))))
        return out

    # ---- Enroller / Receiver roles for approach 5
    def enroll_diagonal(self, db: np.ndarray, first_block: int = 0) -> int:
        """DiagonalEnroller::serializeDB: encrypt every packed row and keep it HBM resident."""
        rows = diagonal_pack(db, self.cc.GetBatchSize())
        for i, row in enumerate(rows):
            v = np.ascontiguousarray(row)
            check_int(self._lib.hx_client_enroll(self._h, first_block + i // VECTOR_DIM, i % VECTOR_DIM,
                                                 v.ctypes.data_as(F64P), len(v)))
        return len(rows) // VECTOR_DIM

    def enroll_row(self, block: int, diag: int, row: np.ndarray):
        """Encrypt one packed row as database ciphertext index{block*512+diag} (enroller_diag.cpp:158-166)."""
        v = np.ascontiguousarray(row, dtype=np.float64)
        check_int(self._lib.hx_client_enroll(self._h, block, diag, v.ctypes.data_as(F64P), len(v)))

    def encryptQuery(self, query: np.ndarray) -> List[Ciphertext]:
        """DiagonalReceiver::encryptQuery: normalise, tile batch/512 times, encrypt."""
        q = plaintext_normalize(np.asarray(query, dtype=np.float64))
        return [self.encrypt(np.tile(q, self.cc.GetBatchSize() // VECTOR_DIM))]

    # ---- Enroller / Receiver roles for approach 4 (HERS)
    def enroll_hers(self, db: np.ndarray) -> int:
        """HersEnroller::serializeDB: column j of matrix m -> database ciphertext matrix{m}/index{j}, HBM resident."""
        packed = hers_pack(db, self.cc.GetBatchSize())
        for m in range(packed.shape[0]):
            for j in range(packed.shape[1]):
                self.enroll_row(m, j, packed[m, j])
        return packed.shape[0]

    def encryptQueryHers(self, query: np.ndarray) -> List[Ciphertext]:
        """HersReceiver::encryptQuery (receiver_hers.cpp:13-24, 56-61): one ciphertext per coordinate, all slots equal."""
        q = plaintext_normalize(np.asarray(query, dtype=np.float64))
        return [self.encrypt(np.full(self.cc.GetBatchSize(), v)) for v in q]

    # ---- Receiver roles for approaches 1 and 3
    def encryptQueryBase(self, query: np.ndarray) -> List[Ciphertext]:
        """BaseReceiver::encryptQuery (receiver_base.cpp:13-26): normalise, tile over the slots."""
        q = plaintext_normalize(np.asarray(query, dtype=np.float64))
        return [self.encrypt(np.tile(q, self.cc.GetBatchSize() // len(q)))]

    def encryptQueryBlind(self, query: np.ndarray, chunk: int) -> List[Ciphertext]:
        """BlindReceiver::encryptQuery (receiver_blind.cpp:13-27, 58-70): chunk c of the query tiled over the slots."""
        q = plaintext_normalize(np.asarray(query, dtype=np.float64))
        return [self.encrypt(np.tile(q[c:c + chunk], self.cc.GetBatchSize() // chunk)) for c in range(0, len(q), chunk)]

    def decryptIndexBlind(self, cts: List[Ciphertext], chunk: int) -> List[int]:
        """BlindReceiver::decryptIndex (receiver_blind.cpp:29-54): undo the slot order of compressCiphers."""
        batch = self.cc.GetBatchSize()
        per = batch // chunk
        out = []
        for i, ct in enumerate(cts):
            v = self.decrypt(ct)
            out.extend(int(i * batch + j // chunk + (j % chunk) * per) for j in np.nonzero(v >= 1.0)[0])
        return out

    def decryptMembership(self, ct: Ciphertext) -> bool:
        return bool(self.decrypt(ct)[0] >= 1.0)

    def decryptIndex(self, cts: List[Ciphertext]) -> List[int]:
        batch = self.cc.GetBatchSize()
        out = []
        for i, ct in enumerate(cts):
            v = self.decrypt(ct)
            out.extend(int(j + i * batch) for j in np.nonzero(v >= 1.0)[0])
        return out
