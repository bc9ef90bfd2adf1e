"""Timeline marks (hx_trace_start / hx_trace_mark / hx_trace_dump): the diagnostic used to name the limiter of a multi-GPU step
(DESIGN.md section 7).  No counterpart in the reference.  The engine's own marks around the rotation lane and the diagonal MAC of
/root/reference/src/sender/sender_diag.cpp:12-33 must come out in stream order, and tracing must leave the residues untouched."""
import numpy as np
import pytest

from oracle import ckks_oracle as O
from tests.helpers import make_gpu_context, make_oracle, same, up

pytestmark = pytest.mark.gpu


def test_trace_marks_are_ordered_and_do_not_change_results():
    ndiag = 8
    orc = make_oracle(ring_dim=4096, depth=11, rotations=list(range(1, ndiag)), seed=4)
    cc = make_gpu_context(orc)
    rng = np.random.default_rng(8)
    dbc = [orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots)) for _ in range(ndiag)]
    for d, c in enumerate(dbc):
        cc.db_register(0, d, np.stack(c.c))
    qc = orc.encrypt(rng.uniform(-0.05, 0.05, orc.P.slots))
    want = O.diag_compute_similarity(orc, qc, dbc, ndiag)
    q = up(cc, qc)
    assert cc.trace_dump() == {}                       # nothing recorded while the trace is off
    cc.trace_mark("ignored")                           # a mark outside a trace is a no-op
    cc.trace_start()
    got = cc.score_diag(q, ndiag)
    cc.trace_mark("user_mark")
    marks = cc.trace_dump()
    assert same(got[0], want[0])
    for name in ("rot_begin", "rot_end", "mac_begin", "mac_end", "user_mark"):
        assert name in marks and len(marks[name]) == 1, marks
    assert "ignored" not in marks
    t = {k: v[0] for k, v in marks.items()}
    assert 0.0 <= t["rot_begin"] <= t["rot_end"] <= t["mac_begin"] <= t["mac_end"] <= t["user_mark"], t
    assert cc.trace_dump() == {}                       # the dump ended the trace
    assert same(cc.score_diag(q, ndiag)[0], want[0])
