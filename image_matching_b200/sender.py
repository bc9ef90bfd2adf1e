"""Mirror of the Sender class surface (/root/reference/include/sender.h:19-43): DiagonalSender (approach 5,
/root/reference/src/sender/sender_diag.cpp) and HersSender (approach 4, src/sender/sender_hers.cpp).
DiagonalSender: Two code paths with identical
results: `fused=True` (default) uses the fused C-ABI entry points; `fused=False` issues the same
per-op call sequence as the reference (511 EvalFastRotation, 512 EvalMultNoRelin, 511 EvalAdd...)."""
from __future__ import annotations

import math
from typing import Callable, List, Optional

from . import openfhe_wrapper as OpenFHEWrapper
from .config import COMP_DEPTH, MATCH_THRESHOLD, VECTOR_DIM
from .engine import Ciphertext, CryptoContext


class Sender:
    """include/sender.h: cc, pk (unused on the evaluation path), numVectors."""

    def __init__(self, cc: CryptoContext, pk, numVectors: int):
        self.cc, self.pk, self.numVectors = cc, pk, numVectors

    def computeSimilarity(self, queryCipher: List[Ciphertext]) -> List[Ciphertext]:
        raise NotImplementedError

    def membershipScenario(self, queryCipher: List[Ciphertext]) -> Ciphertext:
        raise NotImplementedError

    def indexScenario(self, queryCipher: List[Ciphertext]) -> List[Ciphertext]:
        raise NotImplementedError


class DiagonalSender(Sender):
    """sender_diag.cpp. `db_loader(index)` stands for Serial::DeserializeFromFile of
    serial/db_diagonal/index{i}.bin (sender_diag.cpp:87-91) on the per-op path; the fused path
    reads the blocks registered with cc.db_register / client.enroll."""

    def __init__(self, cc: CryptoContext, pk, numVectors: int, fused: bool = True,
                 db_loader: Optional[Callable[[int], Ciphertext]] = None):
        super().__init__(cc, pk, numVectors)
        self.fused, self.db_loader = fused, db_loader

    def _num_matrices(self) -> int:
        return int(math.ceil(float(self.numVectors) / float(self.cc.GetBatchSize())))

    def computeSimilarity(self, queryCipher):
        cc = self.cc
        if self.fused:
            return cc.score_diag(queryCipher[0], VECTOR_DIM)
        cyclotomicOrder = 2 * cc.GetRingDimension()
        rotated = [None] * VECTOR_DIM
        rotated[0] = queryCipher[0]
        precomp = cc.EvalFastRotationPrecompute(queryCipher[0])
        for i in range(1, VECTOR_DIM):
            rotated[i] = cc.EvalFastRotation(queryCipher[0], i, cyclotomicOrder, precomp)
        return [self.computeSimilarityMatrix(rotated, m) for m in range(self._num_matrices())]

    def computeSimilarityMatrix(self, queryCipher, matrix: int) -> Ciphertext:
        cc = self.cc
        score = [self.computeSimilarityThread(queryCipher[i], matrix, i) for i in range(VECTOR_DIM)]
        for i in range(1, VECTOR_DIM):
            score[0] = cc.EvalAddInPlace(score[0], score[i])
        score[0] = cc.RelinearizeInPlace(score[0])
        score[0] = cc.RescaleInPlace(score[0])
        return score[0]

    def computeSimilarityThread(self, queryCipher, matrix: int, index: int) -> Ciphertext:
        databaseCipher = self.db_loader(matrix * VECTOR_DIM + index)
        return self.cc.EvalMultNoRelin(queryCipher, databaseCipher)

    def membershipScenario(self, queryCipher):
        cc = self.cc
        if self.fused:
            return cc.membership(queryCipher[0], VECTOR_DIM, MATCH_THRESHOLD, COMP_DEPTH)
        score = self.computeSimilarity(queryCipher)
        score = [OpenFHEWrapper.chebyshevCompare(cc, s, MATCH_THRESHOLD, COMP_DEPTH) for s in score]
        m = cc.EvalAddManyInPlace(score)
        return cc.EvalSum(m, cc.GetBatchSize())

    def indexScenario(self, queryCipher):
        cc = self.cc
        if self.fused:
            return cc.index(queryCipher[0], VECTOR_DIM, MATCH_THRESHOLD, COMP_DEPTH)
        score = self.computeSimilarity(queryCipher)
        return [OpenFHEWrapper.chebyshevCompare(cc, s, MATCH_THRESHOLD, COMP_DEPTH) for s in score]


class HersSender(Sender):
    """sender_hers.cpp (approach 4). queryCipher = VECTOR_DIM ciphertexts, one per
    coordinate (HersReceiver::encryptQuery); database matrix m, column j = serial/db_hers/matrix{m}/index{j}.bin,
    registered with cc.db_register(m, j, ...) / client.enroll_hers. `fused=False` issues the reference's per-op
    call sequence (EvalMultNoRelin, RelinearizeInPlace, RescaleInPlace per column, then the EvalAddInPlace chain)
    through `db_loader(matrix, index)`."""

    def __init__(self, cc: CryptoContext, pk, numVectors: int, fused: bool = True,
                 db_loader: Optional[Callable[[int, int], Ciphertext]] = None):
        super().__init__(cc, pk, numVectors)
        self.fused, self.db_loader = fused, db_loader

    def computeSimilarity(self, queryCipher):
        cc = self.cc
        if self.fused:
            return cc.score_hers(queryCipher)
        ciphersNeeded = int(math.ceil(float(self.numVectors) / float(cc.GetBatchSize())))
        return [self.computeSimilarityHelper(i, queryCipher) for i in range(ciphersNeeded)]

    def computeSimilarityHelper(self, matrixIndex: int, queryCipher) -> Ciphertext:
        cc = self.cc
        score = []
        for i in range(len(queryCipher)):
            s = self.computeSimilaritySerial(matrixIndex, i, queryCipher[i])
            s = cc.RelinearizeInPlace(s)
            score.append(cc.RescaleInPlace(s))
        for i in range(1, len(score)):
            score[0] = cc.EvalAddInPlace(score[0], score[i])
        return score[0]

    def computeSimilaritySerial(self, matrix: int, index: int, queryCipher: Ciphertext) -> Ciphertext:
        return self.cc.EvalMultNoRelin(queryCipher, self.db_loader(matrix, index))

    def _compare(self, score):
        if self.fused:
            return self.cc.compare_many(score, MATCH_THRESHOLD, COMP_DEPTH)
        return [OpenFHEWrapper.chebyshevCompare(self.cc, s, MATCH_THRESHOLD, COMP_DEPTH) for s in score]

    def indexScenario(self, queryCipher):
        return self._compare(self.computeSimilarity(queryCipher))

    def membershipScenario(self, queryCipher):
        cc = self.cc
        m = cc.EvalAddManyInPlace(self._compare(self.computeSimilarity(queryCipher)))
        return cc.EvalSum(m, cc.GetBatchSize())


class BaseSender(HersSender):
    """sender_base.cpp (approach 1, literature baseline). queryCipher[0] = the normalised
    query tiled over the slots (BaseReceiver::encryptQuery); database ciphertext i = serial/db_baseline/batch{i}.bin =
    batch/dim consecutive vectors (`db_loader(i)`). Per database ciphertext: EvalInnerProduct(q, db, dim), Relinearize,
    Rescale; then OpenFHEWrapper::mergeCiphers packs the scores. `dim` is VECTOR_DIM in the reference."""

    def __init__(self, cc: CryptoContext, pk, numVectors: int, db_loader: Callable[[int], Ciphertext], dim: int = VECTOR_DIM,
                 batched: bool = True):
        super().__init__(cc, pk, numVectors, fused=True)
        self.db_loader, self.dim, self.batched, self._db = db_loader, dim, batched, None

    def computeSimilarity(self, queryCipher):
        """batched=True: the "embarrassingly parallel" loop over database ciphertexts (sender_base.cpp:20-24) and the per-cipher
        merge run as one batched ciphertext (the database batch is stacked once and stays HBM resident)."""
        cc = self.cc
        vectorsPerBatch = cc.GetBatchSize() // self.dim
        numBatches = int(math.ceil(float(self.numVectors) / float(vectorsPerBatch)))
        if self.batched and numBatches > 1:
            if self._db is None:
                self._db = cc.stack([self.db_loader(i) for i in range(numBatches)])
            s = cc.EvalInnerProduct(queryCipher[0], self._db, self.dim)
            s = cc.RescaleInPlace(cc.RelinearizeInPlace(s))
            return OpenFHEWrapper.mergeCiphers(cc, cc.unstack(s), self.dim, batched=True)
        similarity = [self.computeSimilarityThread(queryCipher[0], i) for i in range(numBatches)]
        return OpenFHEWrapper.mergeCiphers(cc, similarity, self.dim)

    def computeSimilarityThread(self, queryCipher: Ciphertext, databaseIndex: int) -> Ciphertext:
        cc = self.cc
        s = cc.EvalInnerProduct(queryCipher, self.db_loader(databaseIndex), self.dim)
        s = cc.RelinearizeInPlace(s)
        return cc.RescaleInPlace(s)


class BlindSender(HersSender):
    """sender_blind.cpp (approach 3, Blind-Match). queryCipher[i] = chunk i of the query tiled
    over the slots (BlindReceiver::encryptQuery); database ciphertext (matrix m, chunk offset j) =
    serial/db_blind/matrix{m}/batch{j}.bin (`db_loader(m, j)`). `dim` / `chunk` are VECTOR_DIM / CHUNK_LEN in the reference."""

    def __init__(self, cc: CryptoContext, pk, numVectors: int, db_loader: Callable[[int, int], Ciphertext],
                 dim: int = VECTOR_DIM, chunk: int = 0, batched: bool = True):
        super().__init__(cc, pk, numVectors, fused=True)
        from .config import CHUNK_LEN
        self.db_loader, self.dim, self.chunk, self.batched, self._db = db_loader, dim, chunk or CHUNK_LEN, batched, None

    def computeSimilarity(self, queryCipher):
        """batched=True: the loop over matrices (sender_blind.cpp:52-54) runs as one batched ciphertext per query chunk."""
        cc = self.cc
        chunksPerBatch = cc.GetBatchSize() // self.chunk
        numMatrices = int(math.ceil(float(self.numVectors) / float(chunksPerBatch)))
        if self.batched and numMatrices > 1:
            chunksPerVector = self.dim // self.chunk
            if self._db is None:
                self._db = [cc.stack([self.db_loader(m, c * self.chunk) for m in range(numMatrices)]) for c in range(chunksPerVector)]
            m0 = cc.EvalMultNoRelin(queryCipher[0], self._db[0])
            for c in range(1, chunksPerVector):
                m0 = cc.EvalAddInPlace(m0, cc.EvalMultNoRelin(queryCipher[c], self._db[c]))
            m0 = cc.RescaleInPlace(cc.RelinearizeInPlace(m0))
            r = 1
            while r < self.chunk:
                m0 = cc.EvalAddInPlace(m0, OpenFHEWrapper.binaryRotate(cc, m0, r))
                r *= 2
            return OpenFHEWrapper.compressCiphers(cc, cc.unstack(m0), self.chunk, batched=True)
        score = [self.computeSimilarityMatrix(queryCipher, self.chunk, i) for i in range(numMatrices)]
        return OpenFHEWrapper.compressCiphers(cc, score, self.chunk)

    def computeSimilarityMatrix(self, queryCipher, chunkLength: int, matrix: int) -> Ciphertext:
        cc = self.cc
        chunksPerVector = self.dim // chunkLength
        m = [self.computeSimilaritySerial(queryCipher[i], matrix, i * chunkLength) for i in range(chunksPerVector)]
        for i in range(1, chunksPerVector):
            m[0] = cc.EvalAddInPlace(m[0], m[i])
        m[0] = cc.RelinearizeInPlace(m[0])
        m[0] = cc.RescaleInPlace(m[0])
        r = 1
        while r < chunkLength:
            m[0] = cc.EvalAddInPlace(m[0], OpenFHEWrapper.binaryRotate(cc, m[0], r))
            r *= 2
        return m[0]

    def computeSimilaritySerial(self, queryCipher: Ciphertext, matrix: int, index: int) -> Ciphertext:
        return self.cc.EvalMultNoRelin(queryCipher, self.db_loader(matrix, index))
