"""oracle/cpu_path.py -- the timed CPU arm: DiagonalSender::membershipScenario + indexScenario on the host cores.

TEST / BENCH INFRASTRUCTURE ONLY (same rule as ckks_oracle.py): used by bench.py's `cpu_baseline` leg and `--impl reference` arm and
by tests/; never by image_matching_b200.  OpenFHE v1.2.3 is not installable here, so this is the oracle PORT of the
path, with the reference's own parallel structure restated:

  * /root/reference/src/sender/sender_diag.cpp:23-26   omp parallel for over the 511 EvalFastRotation calls
  * sender_diag.cpp:70-73                              omp parallel for over the 512 EvalMultNoRelin calls of a block, each of
                                                       which first deserialises its database ciphertext (:87-91)
  * sender_diag.cpp:75-77                              511 EvalAddInPlace on ONE thread
  * sender_diag.cpp:40-43 / :57-60                     omp parallel for over the blocks for chebyshevCompare
  * both scenarios recompute the similarity (src/main.cpp:343,361)

One host thread per loop iteration (a thread pool of `threads` workers; ctypes releases the GIL inside the C primitives), every
worker running the limb arithmetic single-threaded -- OpenFHE's inner per-limb OpenMP loops do not nest under the application's
loops either.  The serial sections use the limb-parallel primitives with all threads.
"""
from __future__ import annotations

import os
import time
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, List, Optional

import numpy as np

from . import ckks_oracle as O


def host_threads() -> int:
    """Hardware threads this process may use (torchrun exports OMP_NUM_THREADS=1 to its workers: ignored on purpose)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class DiskDb:
    """The reference's database: one serialized ciphertext file per (block, diagonal), re-read on every query
    (serial/db_diagonal/index{g*512+d}.bin, sender_diag.cpp:87-91).  Here: raw uint64 limbs, one .npy per ciphertext."""

    def __init__(self, root: str):
        self.root = root
        os.makedirs(root, exist_ok=True)

    def path(self, g: int, d: int) -> str:
        return os.path.join(self.root, f"index{g * O.VECTOR_DIM + d}.npy")

    def store(self, g: int, d: int, ct: O.Ct):
        np.save(self.path(g, d), np.stack(ct.c))

    def load(self, g: int, d: int) -> O.Ct:
        a = np.load(self.path(g, d))
        return O.Ct([a[0], a[1]], 0, 1)


class CpuDiagPath:
    def __init__(self, orc: O.Oracle, threads: Optional[int] = None):
        self.orc = orc
        self.threads = threads or host_threads()
        self.pool = ThreadPoolExecutor(max_workers=self.threads, initializer=lambda: O.lib().ref_set_threads(1))
        O.lib().ref_set_threads(self.threads)   # the calling thread: serial sections run limb-parallel

    def close(self):
        self.pool.shutdown()

    def compute_similarity(self, q: O.Ct, load_db: Callable[[int, int], O.Ct], blocks: List[int], ndiag: int):
        orc = self.orc
        digits = orc.fast_rotation_precompute(q)                                        # sender_diag.cpp:22
        rot = [q] + list(self.pool.map(lambda i: orc.fast_rotation(q, i, digits), range(1, ndiag)))   # :23-26
        scores = []
        for g in blocks:                                                                # :28-30, serial over blocks
            prods = list(self.pool.map(lambda i: orc.mult_norelin(rot[i], load_db(g, i)), range(ndiag)))   # :70-73 (+ :87-91)
            acc = prods[0]
            for i in range(1, ndiag):                                                   # :75-77, one thread
                acc = orc.add(acc, prods[i])
            scores.append(orc.rescale(orc.relinearize(acc)))                            # :79-80
        return scores

    def index(self, q, load_db, blocks, ndiag, delta=O.MATCH_THRESHOLD, depth=O.COMP_DEPTH):
        scores = self.compute_similarity(q, load_db, blocks, ndiag)
        # :57-60 is an omp loop over the blocks; with fewer blocks than threads the port keeps the cores busy with the limb-parallel
        # primitives instead (the reference's single-block compare runs on one thread -- part of its ~2.9 s fixed cost)
        if 2 * len(scores) >= self.threads:
            return list(self.pool.map(lambda s: O.chebyshev_compare(self.orc, s, delta, depth), scores))
        return [O.chebyshev_compare(self.orc, s, delta, depth) for s in scores]

    def membership(self, q, load_db, blocks, ndiag, delta=O.MATCH_THRESHOLD, depth=O.COMP_DEPTH):
        res = self.index(q, load_db, blocks, ndiag, delta, depth)                       # :38-43
        return self.orc.eval_sum(self.orc.add_many(res), self.orc.P.slots)              # :46-47

    def step(self, q, load_db, blocks, ndiag):
        """One query through both scenarios (main.cpp:343,361); returns (membership ct, index cts, seconds)."""
        t0 = time.perf_counter()
        m = self.membership(q, load_db, blocks, ndiag)
        idx = self.index(q, load_db, blocks, ndiag)
        return m, idx, time.perf_counter() - t0
