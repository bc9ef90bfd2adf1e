"""image_matching_b200: B200-native CKKS evaluation engine for the HyDia Sender hot path
(n7koirala/image_matching approach 5). The compute lives in libhydia_b200.so (hand-written
sm_100a CUDA behind the C ABI in include/hydia_b200.h); this package is the thin host mirror of
the reference's Sender / OpenFHEWrapper / cc->Eval* surface. There is no CPU fallback."""
from . import _lib
from .client import Client, base_pack, blind_pack, diagonal_pack, hers_pack, hydia_rotation_set, plaintext_normalize
from .engine import Ciphertext, CryptoContext
from .sender import BaseSender, BlindSender, DiagonalSender, HersSender, Sender
from .sharded import ShardedDiagonalSender, block_range
from . import openfhe_wrapper as OpenFHEWrapper

__all__ = ["CryptoContext", "Ciphertext", "Client", "DiagonalSender", "HersSender", "BaseSender", "BlindSender", "Sender", "ShardedDiagonalSender", "block_range", "OpenFHEWrapper", "diagonal_pack", "hers_pack", "base_pack", "blind_pack",
           "hydia_rotation_set", "plaintext_normalize"]
