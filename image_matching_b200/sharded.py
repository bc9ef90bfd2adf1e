"""Multi-GPU HyDia Sender.  One process per GPU over torch.distributed (NCCL over NVLink on the GPU box;
gloo in the CPU tests).  Evaluation keys are generated from one seed on every rank.

Two shardings of /root/reference/src/sender/sender_diag.cpp's work, same results bit for bit:

shard = "diagonals" (default on GPUs, round 2)
  The 512 diagonals of EVERY block are split over the ranks.  Rank r holds DB[g][d] for its diagonals d only, computes only ITS
  hoisted rotations (sender_diag.cpp:20-26 restricted to d0 <= i < d1: 64 instead of 511 at 8 ranks) and the partial sums
  sum_{d0 <= d < d1} rotQ[d] (x) DB[g][d] for all G blocks (:70-77).  The ONE data-path exchange is a reduce-scatter of those
  un-relinearised accumulators as raw 64-bit words (world * q < 2^63), 9.4 MB per block; one fold brings the words back to
  residues.  Modular addition is associative and commutative, so the block's accumulator equals the single-GPU sum over 512
  diagonals exactly.  Each rank then relinearises, rescales and compares its contiguous share of the blocks (:79-80, :40-43).
  Round 1 replicated the 511 rotations on every rank (18.6 of 55 ms at 8 GPUs); all-gathering sharded rotations would move
  3.2 GB into every GPU per scenario, the partial sums are 0.6 GB in total.

shard = "blocks" (round 1; kept for comparison and for databases enrolled block-wise)
  Rank r owns whole blocks [r*B, (r+1)*B) and every rank computes all rotations.

Result exchange (both shardings):
  indexScenario      gather of the G result ciphertexts (2 x 1 limb each) to rank 0 in global block order, so a decoded index
                     stays j + g*slots (receiver_hers.cpp:46-49)
  membershipScenario local EvalAddMany, SUM-reduce of the residues as int64, one mod-q0 fold, EvalSum on rank 0 -- bit-identical
                     to the reference's EvalAddManyInPlace tree over all G blocks.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np

from .config import COMP_DEPTH, MATCH_THRESHOLD, VECTOR_DIM
from .sender import Sender


def diag_range(rank: int, world: int, ndiag: int = VECTOR_DIM):
    """Contiguous diagonal partition (same rule as block_range)."""
    return block_range(rank, world, ndiag)


def _mark(cc, name, stream=0):
    """Timeline mark on the context if it has the facility (the CPU contexts of the gloo tests do not)."""
    f = getattr(cc, "trace_mark", None)
    if f is not None:
        f(name, stream)


def block_range(rank: int, world: int, total_blocks: int):
    """Contiguous block partition: the first (total % world) ranks own one block more."""
    base, extra = divmod(total_blocks, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class TorchTransport:
    """Moves ciphertext residues between the engine and torch tensors on the same device, stream-ordered (hx_ct_*_dev_async): the
    host never waits, so the next scenario is issued while this one's compare chain still runs (the engine's rotation lane).
    The exchange (copies, stack / pad / fold kernels, the NCCL collective's entry and exit) runs on a transport stream of its own at
    the highest priority: the engine's rotation lanes run at the lowest, which is torch's default too."""

    def __init__(self, cc, device, async_exchange: bool = True):
        import torch
        self.cc, self.device, self.torch = cc, device, torch
        self.async_exchange = async_exchange
        self.stream = torch.cuda.Stream(device=device, priority=-1) if async_exchange else None

    def exchange(self):
        """Context for a scenario's exchange step: the transport stream is torch's current stream inside; it first waits for the
        caller's stream and the caller's stream waits for it on exit."""
        import contextlib
        if not self.async_exchange:
            return contextlib.nullcontext()
        torch, hs = self.torch, self.stream

        @contextlib.contextmanager
        def ctx():
            outer = torch.cuda.current_stream(self.device)
            hs.wait_stream(outer)
            with torch.cuda.stream(hs):
                yield
            outer.wait_stream(hs)
        return ctx()

    class _View:
        """CUDA array interface over a ciphertext's device storage: torch wraps it without a copy."""

        def __init__(self, ptr, words):
            self.__cuda_array_interface__ = {"shape": (words,), "typestr": "<i8", "data": (ptr, False), "version": 3, "strides": None}

    def view(self, ct):
        """int64 tensor [batch entries][words per entry] aliasing the ciphertext's storage (no copy).  The caller orders the
        streams (engine_to_exchange before reading, exchange_to_engine after writing)."""
        ptr, words, nb = self.cc.device_view(ct)
        return self.torch.as_tensor(self._View(ptr, words), device=self.device).view(nb, words // nb)

    def engine_to_exchange(self):
        """torch's current stream waits for everything queued on the engine stream."""
        self.cc.stream_order(int(self.torch.cuda.current_stream(self.device).cuda_stream), engine_waits=False)

    def exchange_to_engine(self):
        """the engine stream waits for everything queued on torch's current stream."""
        self.cc.stream_order(int(self.torch.cuda.current_stream(self.device).cuda_stream), engine_waits=True)

    def reduce_scatter_blocks(self, dist, group, acc, lo, hi, total, world):
        """acc: this rank's partial sums [total][3][L][N].  Returns this rank's blocks lo..hi-1 summed over the ranks as raw 64-bit
        words (a new batched ciphertext; hx_ct_fold_mod still to be applied).  reduce_scatter when the blocks split evenly,
        all_reduce + slice otherwise."""
        ncomp, nl, level, deg = acc.info()
        out = self.cc.alloc(ncomp, nl, level, deg, hi - lo)
        self.engine_to_exchange()
        cur = int(self.torch.cuda.current_stream(self.device).cuda_stream)
        _mark(self.cc, "exchange_begin", cur)   # timeline marks: no-ops unless a trace is on (hx_trace_start)
        src, dst = self.view(acc), self.view(out)
        if total % world == 0:
            dist.reduce_scatter_tensor(dst, src, op=dist.ReduceOp.SUM, group=group)
        else:
            dist.all_reduce(src, op=dist.ReduceOp.SUM, group=group)
            dst.copy_(src[lo:hi])
        _mark(self.cc, "exchange_end", cur)
        self.exchange_to_engine()
        return out, acc   # acc must outlive the collective: the caller drops it after the engine stream has joined

    def reduce_inplace(self, dist, group, ct):
        """int64 SUM of the ciphertext's raw words over the ranks into rank 0, on the ciphertext's own storage (no staging tensor,
        no torch arithmetic: hx_ct_fold_mod brings the sums back below q_i afterwards).  Contents on other ranks are unspecified."""
        self.engine_to_exchange()
        dist.reduce(self.view(ct), dst=0, op=dist.ReduceOp.SUM, group=group)
        self.exchange_to_engine()

    def gather_batched(self, dist, group, ct, rank, world):
        """Every rank's batched ciphertext [B][...] (same B on all ranks) into one batched ciphertext [world * B][...] on rank 0, in
        rank order, received straight into the result's storage.  Returns (result or None, ct): ct must outlive the collective."""
        ncomp, nl, level, deg = ct.info()
        nb = self.cc.device_view(ct)[2]
        out = self.cc.alloc(ncomp, nl, level, deg, nb * world) if rank == 0 else None
        self.engine_to_exchange()
        src = self.view(ct)
        bufs = list(self.view(out).chunk(world, dim=0)) if rank == 0 else None
        dist.gather(src, bufs, dst=0, group=group)
        self.exchange_to_engine()
        return out, ct

    def to_tensor(self, ct):
        ncomp, nl, _, _ = ct.info()
        t = self.torch.empty((ncomp, nl, self.cc.N), dtype=self.torch.int64, device=self.device)
        if self.async_exchange:
            ct.download_dev_async(t.data_ptr(), int(self.torch.cuda.current_stream(self.device).cuda_stream))
        else:
            ct.download_dev(t.data_ptr())
        return t

    def from_tensor(self, t, level: int, deg: int):
        t = t.contiguous()
        if self.async_exchange:
            return self.cc.upload_dev_async(t.data_ptr(), t.shape[0], t.shape[1], level, deg,
                                            int(self.torch.cuda.current_stream(self.device).cuda_stream))
        return self.cc.upload_dev(t.data_ptr(), t.shape[0], t.shape[1], level, deg)


class ShardedDiagonalSender(Sender):
    """DiagonalSender over `world` ranks. `cc` is this rank's context holding its local blocks
    (registered as local block ids 0..B-1) and the full (replicated) key set."""

    def __init__(self, cc, pk, numVectors: int, group=None, transport=None, device=None, shard: str = None):
        super().__init__(cc, pk, numVectors)
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.transport = transport or TorchTransport(cc, device)
        # "diagonals" needs a transport that can run the reduce-scatter on ciphertext storage and a context whose database holds
        # this rank's diagonal range of every block (enroll with diag_range); "blocks" is the round-1 sharding
        self.shard = shard or ("diagonals" if hasattr(self.transport, "reduce_scatter_blocks") else "blocks")
        if self.shard not in ("diagonals", "blocks"):
            raise ValueError("shard must be 'diagonals' or 'blocks'")
        # membershipScenario SUM-reduces residues as int64: the un-reduced sum of `world` residues must stay below 2^63
        if self.world * max(int(m) for m in cc.q) >= 1 << 63:
            raise ValueError(f"ShardedDiagonalSender: world size {self.world} x max modulus exceeds the int64 range of the membership reduce")

    # -- per-rank compute
    def computeSimilarity(self, queryCipher):
        """Scores of THIS rank's blocks (global blocks block_range(rank)), as one batched ciphertext in "diagonals" mode."""
        cc = self.cc
        if self.shard == "blocks":
            return cc.score_diag(queryCipher[0], VECTOR_DIM)
        total = self._total_blocks()
        d0, d1 = diag_range(self.rank, self.world)
        lo, hi = block_range(self.rank, self.world, total)
        acc = cc.diag_mac_partial(queryCipher[0], d0, d1, range(total))       # [total][3][L][N]: partial sums over my diagonals
        with _exchange(self.transport):
            mine, keep = self.transport.reduce_scatter_blocks(self.dist, self.group, acc, lo, hi, total, self.world)
        cc.fold_mod(mine)                                                     # raw sums of `world` residues -> residues
        del keep, acc                                                         # freed in engine-stream order, behind the join
        score = cc.Rescale(cc.Relinearize(mine))                              # sender_diag.cpp:79-80 on my blocks, batched
        _mark(cc, "score_end")
        return score

    def _local_index(self, queryCipher):
        if self.shard == "blocks":
            return self.cc.index(queryCipher[0], VECTOR_DIM, MATCH_THRESHOLD, COMP_DEPTH)
        cc = self.cc
        out = cc.unstack(cc.compare(self.computeSimilarity(queryCipher), MATCH_THRESHOLD, COMP_DEPTH))
        _mark(cc, "compare_end")
        return out

    def indexScenario(self, queryCipher) -> Optional[List]:
        """Rank 0 returns the G result ciphertexts in global block order; other ranks return None."""
        total = self._total_blocks()
        if self.world > 1 and self.shard == "diagonals" and hasattr(self.transport, "gather_batched") and total % self.world == 0:
            # the compare results stay one batched ciphertext and are gathered straight into rank 0's result storage
            cc = self.cc
            cmp = cc.compare(self.computeSimilarity(queryCipher), MATCH_THRESHOLD, COMP_DEPTH)
            _mark(cc, "compare_end")
            with _exchange(self.transport):
                out, keep = self.transport.gather_batched(self.dist, self.group, cmp, self.rank, self.world)
            del keep, cmp                                     # freed in engine-stream order, behind the join
            return cc.unstack(out) if self.rank == 0 else None
        local = self._local_index(queryCipher)
        if self.world == 1:
            return local
        with _exchange(self.transport):
            return self._gather_index(local)

    def _gather_index(self, local):
        import torch
        meta = local[0].info() if local else None
        stacked = torch.stack([self.transport.to_tensor(c) for c in local])          # [B][2][1][N]
        counts = [block_range(r, self.world, self._total_blocks())[1] - block_range(r, self.world, self._total_blocks())[0]
                  for r in range(self.world)]
        bmax = max(counts)
        if stacked.shape[0] < bmax:                                                   # pad ragged shards for the collective
            pad = torch.zeros((bmax - stacked.shape[0],) + tuple(stacked.shape[1:]), dtype=stacked.dtype, device=stacked.device)
            stacked = torch.cat([stacked, pad])
        bufs = [torch.empty_like(stacked) for _ in range(self.world)] if self.rank == 0 else None
        self.dist.gather(stacked, bufs, dst=0, group=self.group)
        if self.rank != 0:
            return None
        out = []
        for r in range(self.world):
            for b in range(counts[r]):
                out.append(self.transport.from_tensor(bufs[r][b], meta[2], meta[3]))
        return out

    def membershipScenario(self, queryCipher):
        """Rank 0 returns the membership ciphertext; other ranks return None."""
        cc = self.cc
        local = self._local_index(queryCipher)
        part = cc.EvalAddMany(local)
        if self.world == 1:
            return cc.EvalSum(part, cc.GetBatchSize())
        _, _, level, deg = part.info()
        if hasattr(self.transport, "reduce_inplace"):
            with _exchange(self.transport):
                self.transport.reduce_inplace(self.dist, self.group, part)
            if self.rank != 0:
                return None
            cc.fold_mod(part)                                 # sums of `world` residues -> residues (hx_ct_fold_mod)
            return cc.EvalSum(part, cc.GetBatchSize())
        with _exchange(self.transport):
            t = self.transport.to_tensor(part)
            self.dist.reduce(t, dst=0, op=self.dist.ReduceOp.SUM, group=self.group)
            if self.rank != 0:
                return None
            t = fold_mod(t, cc.q[: t.shape[1]])
            total = self.transport.from_tensor(t, level, deg)
        return cc.EvalSum(total, cc.GetBatchSize())

    def _total_blocks(self) -> int:
        import math
        return int(math.ceil(float(self.numVectors) / float(self.cc.GetBatchSize())))


def _exchange(transport):
    """The transport's exchange context if it has one (TorchTransport), else nothing (CPU transports of the gloo tests)."""
    import contextlib
    return transport.exchange() if hasattr(transport, "exchange") else contextlib.nullcontext()


_FOLD_Q = {}


def fold_mod(t, moduli):
    """Residue-wise t mod q_i over the limb axis (axis 1) of an int64 [ncomp][nl][N] tensor of
    un-reduced sums (each < world * q_i < 2^63).  The modulus tensor is built once per (device, moduli):
    torch.tensor(list, device=cuda) is a synchronous copy from pageable memory and would stall the host
    behind the collective it follows."""
    import torch
    key = (str(t.device), tuple(int(m) for m in moduli))
    q = _FOLD_Q.get(key)
    if q is None:
        q = _FOLD_Q[key] = torch.tensor(list(key[1]), dtype=torch.int64, device=t.device).view(1, -1, 1)
    return torch.remainder(t, q)
