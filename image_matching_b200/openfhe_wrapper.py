"""Mirror of the OpenFHEWrapper namespace (/root/reference/src/openFHE_wrapper.cpp,
include/openFHE_wrapper.h) for the functions on the HyDia Sender path."""
from __future__ import annotations

import math
import sys
from typing import List

from .engine import Ciphertext, CryptoContext

# openFHE_wrapper.cpp:153-155 and :158-169
DEPTH_TO_DEGREE = [-1, -1, -1, 5, 13, 27, 59, 119, 247, 495, 1007, 2031]
F4_COEFS = [0.0, 315.0 / 128.0, 0.0, -420.0 / 128.0, 0.0, 378.0 / 128.0, 0.0, -180.0 / 128.0, 0.0, 35.0 / 128.0]


def computeRequiredDepth(approach: int) -> int:
    """openFHE_wrapper.cpp:6-44."""
    from .config import ALPHA_DEPTH, COMP_DEPTH
    return {1: 3 + COMP_DEPTH, 2: 6 + COMP_DEPTH + ALPHA_DEPTH, 3: 2 + COMP_DEPTH, 4: 1 + COMP_DEPTH, 5: 1 + COMP_DEPTH}[approach]


def chebyshev_coefficients(func, a: float, b: float, degree: int) -> List[float]:
    """What cc->EvalChebyshevFunction computes before the series evaluation
    (OpenFHE EvalChebyshevCoefficients; call site openFHE_wrapper.cpp:174)."""
    n = degree + 1
    b_minus_a, b_plus_a = 0.5 * (b - a), 0.5 * (b + a)
    pi_by_deg = math.pi / float(n)
    pts = [func(math.cos(pi_by_deg * (i + 0.5)) * b_minus_a + b_plus_a) for i in range(n)]
    mult = 2.0 / float(n)
    out = []
    for i in range(n):
        s = 0.0
        for j in range(n):
            s += pts[j] * math.cos(pi_by_deg * i * (j + 0.5))
        out.append(s * mult)
    return out


def chebyshevCompare(cc: CryptoContext, ctxt: Ciphertext, delta: float, signDepth: int) -> Ciphertext:
    """OpenFHEWrapper::chebyshevCompare (openFHE_wrapper.cpp:143-185): x -> ~2 if x >= delta else ~0.
    Issued as the reference does, one cc-> call per step (the C ABI also has the fused hx_compare)."""
    if signDepth < 7 or signDepth > 15:
        print("Error: chebshevCompare requires a depth parameter between 7 and 15", file=sys.stderr)
        return ctxt
    degree = DEPTH_TO_DEGREE[signDepth - 4]
    coeffs = chebyshev_coefficients(lambda x: 1.0 if x >= delta else -1.0, -1.0, 1.0, degree)
    ctxt = cc.EvalChebyshevSeries(ctxt, coeffs, -1.0, 1.0)
    ctxt = cc.EvalPoly(ctxt, F4_COEFS)
    return cc.EvalAdd(ctxt, 1.0)


def binaryRotate(cc: CryptoContext, ctxt: Ciphertext, factor: int) -> Ciphertext:
    """OpenFHEWrapper::binaryRotate (openFHE_wrapper.cpp:103-128): decompose into +-2^k rotations."""
    batch = cc.GetBatchSize()
    needed = []
    while factor != 0:
        sign = 1 if factor > 0 else -1
        step = int(pow(2, round(math.log2(abs(factor)))))
        cur = int(math.fmod(step * sign, batch))
        if cur != 0:
            needed.append(step * sign)
        factor -= step * sign
    for r in needed:
        ctxt = cc.EvalRotate(ctxt, r)
    return ctxt


def generateMergeMask(cc: CryptoContext, dimension: int, segmentLength: int):
    """OpenFHEWrapper::generateMergeMask (openFHE_wrapper.cpp:253-271): segments of ones every dimension*segmentLength slots."""
    batch = cc.GetBatchSize()
    cache = cc.__dict__.setdefault("_merge_masks", {})      # the same (dimension, segment) always encodes to the same plaintext
    if (dimension, segmentLength) in cache:
        return cache[(dimension, segmentLength)]
    mask = [0.0] * batch
    if segmentLength > batch // dimension:
        print("Mask generation index error", file=sys.stderr)
        return cc.MakeCKKSPackedPlaintext(mask)
    i = 0
    while i < batch:
        for k in range(i, min(batch, i + segmentLength)):
            mask[k] = 1.0
        i += dimension * segmentLength
    cache[(dimension, segmentLength)] = cc.MakeCKKSPackedPlaintext(mask)
    return cache[(dimension, segmentLength)]


def mergeSingleCipher(cc: CryptoContext, ctxt: Ciphertext, dimension: int) -> Ciphertext:
    """OpenFHEWrapper::mergeSingleCipher (openFHE_wrapper.cpp:221-249): packs every dimension-th slot to the front."""
    batch = cc.GetBatchSize()
    outputSize = batch // dimension
    paddingSize = 1
    rotationFactor = dimension - 1
    i = 1
    while i < outputSize:
        if i >= paddingSize:
            ctxt = cc.EvalMult(ctxt, generateMergeMask(cc, dimension, i))
            ctxt = cc.RelinearizeInPlace(ctxt)
            ctxt = cc.RescaleInPlace(ctxt)
            paddingSize = i * dimension
        ctxt = cc.EvalAddInPlace(ctxt, binaryRotate(cc, ctxt, rotationFactor * i))
        i *= 2
    ctxt = cc.EvalMult(ctxt, generateMergeMask(cc, dimension, outputSize))
    ctxt = cc.RelinearizeInPlace(ctxt)
    return cc.RescaleInPlace(ctxt)


def mergeCiphers(cc: CryptoContext, ctxts: List[Ciphertext], dimension: int, batched: bool = False) -> List[Ciphertext]:
    """OpenFHEWrapper::mergeCiphers (openFHE_wrapper.cpp:189-217). batched=True runs the per-cipher mergeSingleCipher
    loop (the reference's "#pragma omp parallel for", :199-202) as ONE batched ciphertext: same operations per entry."""
    batch = cc.GetBatchSize()
    elementsPerCipher = batch // dimension
    outputSize = elementsPerCipher * len(ctxts)
    neededCiphers = int(math.ceil(float(outputSize) / float(batch)))
    if batched and len(ctxts) > 1:
        ctxts = cc.unstack(mergeSingleCipher(cc, cc.stack(ctxts), dimension))
    else:
        ctxts = [mergeSingleCipher(cc, c, dimension) for c in ctxts]
    merged: List[Ciphertext] = [None] * neededCiphers
    for i, c in enumerate(ctxts):
        outputCipher = (elementsPerCipher * i) // batch
        outputSlot = (elementsPerCipher * i) % batch
        if outputSlot == 0:
            merged[outputCipher] = c
        else:
            merged[outputCipher] = cc.EvalAddInPlace(merged[outputCipher], binaryRotate(cc, c, -outputSlot))
    return merged


def compressCiphers(cc: CryptoContext, ctxts: List[Ciphertext], dimension: int, batched: bool = False) -> List[Ciphertext]:
    """OpenFHEWrapper::compressCiphers (openFHE_wrapper.cpp:273-312): keeps the dimension-th slots, not in order.
    batched=True runs the masking loop (:287-292) on one batched ciphertext."""
    batch = cc.GetBatchSize()
    ciphersNeeded = int(math.ceil(float(len(ctxts)) / float(dimension)))
    maskVec = [0.0] * batch
    for i in range(0, batch, dimension):
        maskVec[i] = 1.0
    maskPtxt = cc.MakeCKKSPackedPlaintext(maskVec)
    out = []
    if batched and len(ctxts) > 1:
        c = cc.EvalMult(cc.stack(ctxts), maskPtxt)
        c = cc.RelinearizeInPlace(c)
        out = cc.unstack(cc.RescaleInPlace(c))
        ctxts = []
    for c in ctxts:
        c = cc.EvalMult(c, maskPtxt)
        c = cc.RelinearizeInPlace(c)
        out.append(cc.RescaleInPlace(c))
    compressed: List[Ciphertext] = [None] * ciphersNeeded
    for i, c in enumerate(out):
        rotFactor = -(i % dimension)
        outputSlot = i // dimension
        if rotFactor == 0:
            compressed[outputSlot] = c
        else:
            compressed[outputSlot] = cc.EvalAddInPlace(compressed[outputSlot], binaryRotate(cc, c, rotFactor))
    return compressed
