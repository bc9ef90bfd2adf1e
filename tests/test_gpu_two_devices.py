"""Two contexts on two devices in ONE process: every entry point makes its engine's device current, the
dynamic shared-memory opt-in of the kernels is done per device, handles are freed on their own device.  Interleaved calls from one
thread and concurrent calls from two threads, each result bit-exact against the oracle (key switch, rescale, hoisted rotations,
the diagonal scenario of /root/reference/src/sender/sender_diag.cpp:12-63 on a small ring).  Skipped on a one-GPU box."""
import threading

import numpy as np
import pytest

from oracle import ckks_oracle as O
from tests.helpers import make_gpu_context, make_oracle, same, up

pytestmark = pytest.mark.gpu


def _two_gpus():
    from image_matching_b200._lib import load
    return load().hx_device_count() >= 2


@pytest.mark.skipif("not _two_gpus()")
def test_two_contexts_two_devices_one_process():
    ndiag = 8
    orc = make_oracle(ring_dim=4096, depth=11, rotations=list(range(1, ndiag)) + [1, 2, 4, 8, 16, 32, 64, 128, 256, 512, 1024], seed=9)
    ccs = [make_gpu_context(orc, device=d) for d in (0, 1)]
    rng = np.random.default_rng(3)
    a, b = rng.uniform(-1, 1, orc.P.slots), rng.uniform(-1, 1, orc.P.slots)
    ca, cb = orc.encrypt(a), orc.encrypt(b)
    want_mul = orc.rescale(orc.mult(ca, cb))
    want_rot = orc.rotate(ca, 3)
    dw = orc.fast_rotation_precompute(ca)
    want_fast = orc.fast_rotation(ca, 5, dw)
    # interleaved from one thread: device 0, device 1, device 0, ... (the thread's current device flips on every call)
    ups = [(up(cc, ca), up(cc, cb)) for cc in ccs]
    prods = [cc.EvalMult(x, y) for cc, (x, y) in zip(ccs, ups)]
    res = [cc.Rescale(p) for cc, p in zip(ccs, prods)]
    rots = [cc.EvalRotate(x, 3) for cc, (x, _) in zip(ccs, ups)]
    digs = [cc.EvalFastRotationPrecompute(x) for cc, (x, _) in zip(ccs, ups)]
    fast = [cc.EvalFastRotation(x, 5, 2 * orc.N, d) for cc, (x, _), d in zip(ccs, ups, digs)]
    for i in (0, 1):
        assert same(res[i], want_mul) and same(rots[i], want_rot) and same(fast[i], want_fast), f"device {i}"
    # the diagonal scenario on both devices from two threads at once
    dbc = [orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots)) for _ in range(ndiag)]
    qc = orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots))
    want = O.diag_compute_similarity(orc, qc, dbc, ndiag)
    widx = [O.chebyshev_compare(orc, s) for s in want]
    for cc in ccs:
        for d, c in enumerate(dbc):
            cc.db_register(0, d, np.stack(c.c))
    out, errs = [None, None], []

    def work(i):
        try:
            q = up(ccs[i], qc)
            r = []
            for _ in range(3):
                r.append((ccs[i].score_diag(q, ndiag), ccs[i].index(q, ndiag)))
            out[i] = r
        except Exception as e:   # noqa: BLE001 -- reported below
            errs.append((i, repr(e)))

    th = [threading.Thread(target=work, args=(i,)) for i in (0, 1)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    for i in (0, 1):
        for sc, ix in out[i]:
            assert len(sc) == 1 and same(sc[0], want[0]) and same(ix[0], widx[0]), f"device {i}"
    # handles of both devices are released from a thread whose current device is neither's last one
    del ups, prods, res, rots, digs, fast, out
    for cc in ccs:
        cc.sync()
