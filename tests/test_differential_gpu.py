"""Randomised differential test: arbitrary small task lists (hand-picked adversarial due dates,
kinds, flags, duplicate creators and creation times) through tt_load_soa, every query compared
with the oracle in both reference modes.  Complements the generated-table tests: here the rows
are NOT in TaskId order, so the id index, the padding and the ragged tail are exercised."""
import uuid

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from helpers import load_from_oracle, tt

pytestmark = pytest.mark.gpu
TPD = tt.TICKS_PER_DAY
TODAY = tt.DEFAULT_TODAY
YDAY = TODAY - TPD

due_ticks = st.sampled_from([YDAY, YDAY + 1, YDAY - 1, YDAY + TPD - 1, TODAY, YDAY - TPD, YDAY + 36_000_000_000, 0, TPD])
kinds = st.sampled_from([0, 0, 0, 1, 2])
creators = st.sampled_from([b"a@x.io", b"A@x.io", b"a@x.io ", b"b@x.io", "zoë@x.io".encode(), b""])
task = st.tuples(due_ticks, kinds, st.booleans(), st.booleans(), st.integers(0, 5), creators)


@settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(tasks=st.lists(task, min_size=0, max_size=70), skew=st.sampled_from([0, 0, -3, 2]), seed=st.integers(0, 2**31))
def test_random_small_tables(oracle, tasks, skew, seed):
    rng = np.random.default_rng(seed)
    ot = oracle.OracleTable()
    for due, kind, comp, over, created, who in tasks:
        tid = bytes(rng.integers(0, 256, size=16, dtype=np.uint8))
        ot.add(tid, "n", who, "a", (TODAY - 1000 + created) | (1 << 62), due | (kind << 62), comp, over)
    run_at = TODAY + 5 * 60 * 10_000_000 - skew * 3600 * 10_000_000
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        table = load_from_oracle(ot) if len(ot) else tt.TaskTable(0)
        want = ot.get_yesterdays_due(mode, TODAY, 60)
        got = table.get_yesterdays_due(mode, TODAY).rows()
        assert got.tolist() == want.tolist()
        for who in (b"a@x.io", b"A@x.io", b"", b"nobody", "zoë@x.io".encode()):
            assert table.get_by_creator(who).rows().tolist() == ot.get_by_creator(mode, who).tolist()
        table.close()
    # the tick itself, STORE then FAKE on the same evolving state
    table = load_from_oracle(ot) if len(ot) else tt.TaskTable(0)
    for mode in (tt.MODE_STORE, tt.MODE_FAKE, tt.MODE_STORE):
        threw, listed, n_over = ot.cron_job(mode, TODAY, run_at, 60)
        res, n_marked, gthrew = table.sweep(mode, TODAY, run_at)
        assert gthrew == threw and res.rows().tolist() == sorted(listed.tolist())
        assert n_marked == (n_over if mode == tt.MODE_STORE else 0)
        if len(ot):
            c, o = ot.flags()
            gc, go = table.flags_bytes()
            assert (gc == c).all() and (go == o).all()
    table.close()
