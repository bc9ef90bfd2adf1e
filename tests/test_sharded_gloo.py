"""World-size-2 (gloo, CPU) test of the multi-GPU host logic of ShardedDiagonalSender:
contiguous block partition, gather order of the index scenario over ragged shards, SUM-reduce + mod-q0
fold of the membership scenario.  The engine under the sender is replaced by a stand-in whose
"compare outputs" are seeded residues (the device arithmetic is covered by the -m gpu parity tests);
what is checked is that the combined results equal the single-process EvalAddMany / block order."""
import os
import socket

import numpy as np
import pytest

N, Q0 = 256, (1 << 60) - 93          # a 60-bit modulus: world * q0 < 2^63 is the property the int64 reduce relies on
TOTAL_BLOCKS = 5                     # ragged over 2 ranks: 3 + 2


def block_residues(g):
    return np.random.default_rng(1000 + g).integers(0, Q0, size=(2, 1, N), dtype=np.uint64)


class FakeCt:
    def __init__(self, data, level=10, deg=1):
        self.data, self.level, self.deg = data, level, deg

    def info(self):
        return self.data.shape[0], self.data.shape[1], self.level, self.deg


class FakeContext:
    """Stands for this rank's CryptoContext: local blocks lo..hi registered as local ids 0..B-1."""

    def __init__(self, lo, hi):
        self.N, self.q, self.blocks = N, [Q0], list(range(lo, hi))

    def GetBatchSize(self):
        return N // 2

    def index(self, q, ndiag, delta, depth):
        return [FakeCt(block_residues(g)) for g in self.blocks]

    def EvalAddMany(self, cts):
        acc = np.zeros_like(cts[0].data)
        for c in cts:
            acc = (acc + c.data) % np.uint64(Q0)
        return FakeCt(acc, cts[0].level, cts[0].deg)

    def EvalSum(self, ct, n):
        return ct                     # rotations are engine work; identity keeps the reduce observable


class CpuTransport:
    def to_tensor(self, ct):
        import torch
        return torch.from_numpy(ct.data.astype(np.int64))

    def from_tensor(self, t, level, deg):
        return FakeCt(t.numpy().astype(np.uint64), level, deg)


def _worker(rank, world, port, out):
    import torch.distributed as dist
    from image_matching_b200.sharded import ShardedDiagonalSender, block_range
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = block_range(rank, world, TOTAL_BLOCKS)
        cc = FakeContext(lo, hi)
        snd = ShardedDiagonalSender(cc, None, TOTAL_BLOCKS * cc.GetBatchSize(), transport=CpuTransport())
        idx = snd.indexScenario([None])
        mem = snd.membershipScenario([None])
        if rank == 0:
            out.put(([c.data for c in idx], mem.data, [(c.level, c.deg) for c in idx]))
        else:
            assert idx is None and mem is None
            out.put("ok")
    finally:
        dist.destroy_process_group()


def test_block_range_partitions_contiguously():
    from image_matching_b200.sharded import block_range
    for world in (1, 2, 3, 4, 8):
        for total in (1, 5, 8, 64):
            ranges = [block_range(r, world, total) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == total
            assert all(ranges[i][1] == ranges[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in ranges]
            assert max(sizes) - min(sizes) <= 1


def test_fold_mod_is_exact_for_sums_of_world_residues():
    import torch
    from image_matching_b200.sharded import fold_mod
    rng = np.random.default_rng(0)
    parts = [rng.integers(0, Q0, size=(2, 1, 64), dtype=np.uint64) for _ in range(8)]
    s = sum(torch.from_numpy(p.astype(np.int64)) for p in parts)        # 8 * q0 < 2^63: no wrap
    want = np.zeros_like(parts[0])
    for p in parts:
        want = (want + p) % np.uint64(Q0)
    assert np.array_equal(fold_mod(s, [Q0]).numpy().astype(np.uint64), want)


@pytest.mark.timeout(300)
def test_sharded_sender_world2_gloo():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    idx, mem, meta = next(r for r in results if r != "ok")
    assert len(idx) == TOTAL_BLOCKS and all(m == (10, 1) for m in meta)
    for g in range(TOTAL_BLOCKS):                                        # global block order survives the ragged gather
        assert np.array_equal(idx[g], block_residues(g))
    want = np.zeros((2, 1, N), dtype=np.uint64)
    for g in range(TOTAL_BLOCKS):
        want = (want + block_residues(g)) % np.uint64(Q0)
    assert np.array_equal(mem, want)                                     # == single-process EvalAddMany over all blocks


# ---- "diagonals" sharding: partial sums over a rank's diagonals for ALL blocks, reduce-scattered, folded --------------------------
Q_LIMBS = [Q0, (1 << 45) + 59, (1 << 45) - 229]          # stand-ins for the limb moduli (the fold is per limb)


def partial_residues(rank, g):
    """what rank `rank`'s diagonal range contributes to block g: [3][L][N] residues"""
    rng = np.random.default_rng(7000 + 100 * rank + g)
    return np.stack([np.stack([rng.integers(0, q, size=N, dtype=np.uint64) for q in Q_LIMBS]) for _ in range(3)])


class FakeBatch:
    def __init__(self, data, level=0, deg=2):
        self.data, self.level, self.deg = data, level, deg      # [nb][ncomp][L][N]

    def info(self):
        return self.data.shape[1], self.data.shape[2], self.level, self.deg


class FakeDiagContext:
    def __init__(self, rank):
        self.rank, self.N, self.q = rank, N, Q_LIMBS

    def GetBatchSize(self):
        return N // 2

    def diag_mac_partial(self, q, d0, d1, blocks):
        self.range = (d0, d1)
        return FakeBatch(np.stack([partial_residues(self.rank, g) for g in blocks]))

    def fold_mod(self, ct):
        for i, q in enumerate(Q_LIMBS):
            ct.data[:, :, i] %= np.uint64(q)

    def Relinearize(self, ct): return ct
    def Rescale(self, ct): return ct
    def compare(self, ct, delta, depth): return ct
    def unstack(self, ct): return [FakeCt(ct.data[i][:2, :1]) for i in range(ct.data.shape[0])]

    def EvalAddMany(self, cts):
        acc = np.zeros_like(cts[0].data)
        for c in cts:
            acc = (acc + c.data) % np.uint64(Q0)
        return FakeCt(acc, cts[0].level, cts[0].deg)

    def EvalSum(self, ct, n): return ct


class CpuDiagTransport(CpuTransport):
    def reduce_scatter_blocks(self, dist, group, acc, lo, hi, total, world):
        import torch
        t = torch.from_numpy(acc.data.astype(np.int64))
        if total % world == 0:
            out = torch.empty((hi - lo,) + tuple(t.shape[1:]), dtype=torch.int64)
            dist.reduce_scatter_tensor(out, t, op=dist.ReduceOp.SUM, group=group)
        else:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
            out = t[lo:hi].clone()
        return FakeBatch(out.numpy().astype(np.uint64)), acc


def _diag_worker(rank, world, port, total, out):
    import torch.distributed as dist
    from image_matching_b200.sharded import ShardedDiagonalSender, diag_range
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cc = FakeDiagContext(rank)
        snd = ShardedDiagonalSender(cc, None, total * cc.GetBatchSize(), transport=CpuDiagTransport())
        assert snd.shard == "diagonals"
        idx = snd.indexScenario([None])
        assert cc.range == diag_range(rank, world)
        out.put([c.data for c in idx] if rank == 0 else "ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("total", [4, 5])          # even split: reduce_scatter; ragged: all_reduce + slice
def test_diagonal_sharding_world2_gloo(total):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_diag_worker, args=(r, 2, port, total, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    idx = next(r for r in results if not isinstance(r, str))
    assert len(idx) == total
    for g in range(total):                          # block g = (sum over the ranks' partial sums) mod q, in global block order
        want = (partial_residues(0, g).astype(object) + partial_residues(1, g).astype(object))
        want = np.stack([np.stack([(want[c][i] % Q_LIMBS[i]).astype(np.uint64) for i in range(len(Q_LIMBS))]) for c in range(3)])
        assert np.array_equal(idx[g], want[:2, :1])
