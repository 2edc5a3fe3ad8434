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


# ---- the text either side of the path: K6 (results as JSON) and K7 (documents into a table) -------------
# strings over an alphabet that holds every class of character the two encoders and the reader treat
# differently: plain ASCII, the escapes, HTML-sensitive ASCII, controls, DEL, C1, 2/3/4-byte UTF-8,
# separators, private use, non-characters
ALPHABET = ['a', 'Z', '0', ' ', '"', '\\', '/', "'", '+', '<', '>', '&', '`', '\b', '\t', '\n', '\f', '\r', '\x00', '\x1f',
            '\x7f', '\x80', '\x9f', '\xa0', '\xe9', '\u0394', '\u1680', '\u2003', '\u2028', '\u2029', '\u4e2d', '\ue000',
            '\ufdd0', '\ufffd', '\uffff', '\U0001F600', '\U0010FFFF']
text = st.text(alphabet=st.sampled_from(ALPHABET), min_size=0, max_size=40)
stamp = st.tuples(st.integers(0, 3_652_058), st.sampled_from([0, 0, 1, 10_000_000 - 1, 5_000_000, 123_000, 863_999_999_999]),
                  st.sampled_from([0, 1, 2]))
record = st.tuples(text, text, text, stamp, stamp, st.booleans(), st.booleans())


@settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(records=st.lists(record, min_size=1, max_size=40), offset=st.sampled_from([0, 60, -480, 345]), seed=st.integers(0, 2**31))
def test_random_text_round_trips(oracle, records, offset, seed):
    """random TaskModels -> oracle documents -> K7 -> K6 (both encoders) == oracle's text; and K6's own
    documents load back through K7 into a table that prints the same text again"""
    rng = np.random.default_rng(seed)
    src = oracle.OracleTable()
    for name, who, assignee, (cd, ct, ck), (dd, dtod, dk), comp, over in records:
        tid = bytes(rng.integers(0, 256, size=16, dtype=np.uint8))
        created = min(cd * TPD + ct % TPD, 3_155_378_975_999_999_999) | (ck << 62)
        due = min(dd * TPD + dtod % TPD, 3_155_378_975_999_999_999) | (dk << 62)
        src.add(tid, name.encode(), b"same@creator", assignee.encode(), created, due, comp, over)
        src.add(tid, name.encode(), who.encode(), assignee.encode(), created, due, comp, over)
    rows = np.arange(len(src))
    for flavor in (tt.JSON_RESPONSE, tt.JSON_DOCUMENT):
        docs, off = src.to_documents(rows, flavor, offset)
        ot = oracle.OracleTable()
        assert ot.add_documents(docs, off, offset) == -1
        table = tt.TaskTable(0)
        st_ = table.load_documents(docs, off, offset)
        assert st_.first_bad == -1 and st_.n_docs == len(src)
        res = table.get_by_creator(b"same@creator")
        listed = ot.get_by_creator(oracle.MODE_FAKE, b"same@creator")
        assert res.rows().tolist() == listed.tolist() and res.count >= len(records)
        for out_flavor in (tt.JSON_RESPONSE, tt.JSON_DOCUMENT):
            assert res.to_json(out_flavor, offset) == ot.to_json(listed, out_flavor, offset)
        # device to device: what K6 wrote is what K7 reads
        j = res.json(flavor, offset)
        again = tt.TaskTable(0)
        again.load_documents(j.device_ptr(), j.offsets(res.count), offset)
        back = again.get_by_creator(b"same@creator")
        assert back.count == res.count and back.to_json(flavor, offset) == j.bytes()
        j.free()
        again.close()
        table.close()
