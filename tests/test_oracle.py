"""Known-answer tests for the CPU restatement (oracle/oracle.cpp).

The reference has no tests, so these KATs are derived by hand from the cited
reference lines; the only external pins are the JSON samples in the workshop docs.
"""
import datetime as dt
import os

import numpy as np
import pytest

TPD = 864000000000


def ticks(y, m, d, hh=0, mm=0, ss=0, frac=0):
    days = dt.date(y, m, d).toordinal() - 1
    return days * TPD + ((hh * 60 + mm) * 60 + ss) * 10_000_000 + frac


def raw(t, kind=0):
    return t | (kind << 62)


UNSPEC, UTC, LOCAL = 0, 1, 2
TODAY = ticks(2026, 11, 6)
YDAY = TODAY - TPD


def gid(i):
    return bytes([i] * 6 + [0x40, i] + [0x80] + [i] * 7)


def test_doc_samples_pin_the_json_date_format(oracle):
    # docs/aca/05-aca-dapr-pubsubapi/index.md:65-66
    assert oracle.format_stj(raw(ticks(2023, 2, 12, 0, 24, 37, 7361348), UTC)) == "2023-02-12T00:24:37.7361348Z"
    assert oracle.format_stj(raw(ticks(2023, 2, 20), UNSPEC)) == "2023-02-20T00:00:00"
    # trailing zeros of the fraction are trimmed; Local carries the offset
    assert oracle.format_stj(raw(ticks(2023, 2, 20, 1, 2, 3, 5000000), UNSPEC)) == "2023-02-20T01:02:03.5"
    assert oracle.format_stj(raw(ticks(2023, 2, 20), LOCAL), 120) == "2023-02-20T00:00:00+02:00"
    assert oracle.format_stj(raw(ticks(2023, 2, 20), LOCAL), -330) == "2023-02-20T00:00:00-05:30"
    # DateTimeConverter("yyyy-MM-ddTHH:mm:ss") drops fraction and kind (Api/Utilities/DateTimeConverter.cs:28)
    assert oracle.format_converter(raw(ticks(2026, 11, 5, 0, 0, 0, 1234567), UTC)) == "2026-11-05T00:00:00"
    assert oracle.format_converter(raw(ticks(2024, 2, 29, 23, 59, 59), LOCAL)) == "2024-02-29T23:59:59"


def test_default_today_constant(oracle):
    assert oracle.DEFAULT_TODAY == TODAY
    assert TODAY // TPD == 739925


def test_guid_string_is_canonical_big_endian(oracle):
    b = bytes(range(0x10, 0x20))
    assert oracle.guid_string(b) == "10111213-1415-1617-1819-1a1b1c1d1e1f"


def test_fake_seed_rows_config1(oracle):
    # Api/Services/FakeTasksManager.cs:12-24 — every due date >= now, with a time of day.
    now = ticks(2026, 11, 6, 9, 30, 15, 1234567)
    t = oracle.OracleTable()
    t.seed_fake(now)
    assert len(t) == 10
    for mode in (oracle.MODE_FAKE, oracle.MODE_STORE):
        assert t.get_yesterdays_due(mode, TODAY).size == 0
        threw, listed, n_over = t.cron_job(mode, TODAY, now)
        assert (threw, listed.size, n_over) == (False, 0, 0)  # Mark never invoked => no NotImplementedException
        rows = t.get_by_creator(mode, "tjoudeh@bitoftech.net")
        assert rows.tolist() == list(range(9, -1, -1))  # OrderByDescending(TaskCreatedOn)
        assert t.get_by_creator(mode, "Tjoudeh@bitoftech.net").size == 0  # ordinal, case-sensitive
    # even the day after, tick equality with midnight fails for the seed rows (time of day)
    assert t.get_yesterdays_due(oracle.MODE_FAKE, TODAY + TPD).size == 0


def build_kat(oracle):
    t = oracle.OracleTable()
    rows = [
        # due,                       kind,   completed, overdue
        (YDAY, UNSPEC, False, False),            # 0 exact hit
        (YDAY + 1, UNSPEC, False, False),        # 1 one tick past midnight: never matches
        (YDAY, UTC, False, False),               # 2 Kind Utc: FAKE hit (ticks only), STORE miss ("...Z")
        (YDAY, LOCAL, False, False),             # 3 Kind Local: FAKE hit, STORE miss ("+hh:mm")
        (YDAY, UNSPEC, True, False),             # 4 completed: FAKE hit, STORE filtered (:135)
        (YDAY, UNSPEC, False, True),             # 5 already overdue: FAKE hit, STORE filtered
        (YDAY - TPD, UNSPEC, False, False),      # 6 two days ago: miss
        (TODAY, UNSPEC, False, False),           # 7 today: miss
        (YDAY, UNSPEC, False, False),            # 8 exact hit, created earlier than row 0
        (YDAY, UNSPEC, False, False),            # 9 exact hit, same TaskCreatedOn as row 0 (tie)
        (YDAY + 10_000_000 * 3600, UTC, False, False),  # 10 queue-ingested shape: never matches
    ]
    created = [500, 100, 200, 300, 400, 450, 50, 60, 10, 500, 70]
    for i, (due, kind, comp, over) in enumerate(rows):
        t.add(gid(i + 1), f"Task {i}", "a@b.c", "x@y.z", raw(TODAY - TPD * 30 + created[i], UTC), raw(due, kind), comp, over)
    return t


def test_sweep_fake_mode(oracle):
    t = build_kat(oracle)
    # Api/Services/FakeTasksManager.cs:109 — ticks only, no flag filter, source order
    assert t.get_yesterdays_due(oracle.MODE_FAKE, TODAY).tolist() == [0, 2, 3, 4, 5, 8, 9]
    c0, o0 = t.flags()
    threw, listed, n_over = t.cron_job(oracle.MODE_FAKE, TODAY, TODAY + 3000000000)
    assert threw and listed.tolist() == [0, 2, 3, 4, 5, 8, 9] and n_over == 7
    c1, o1 = t.flags()
    assert (c0 == c1).all() and (o0 == o1).all()  # the throw leaves the list untouched


def test_sweep_store_mode(oracle):
    for string_eq in (False, True):
        t = build_kat(oracle)
        # Api/Services/TasksStoreManager.cs:135-136 — pending only, OrderBy(TaskCreatedOn) stable
        assert t.get_yesterdays_due(oracle.MODE_STORE, TODAY, 0, string_eq).tolist() == [8, 0, 9]
        threw, listed, n_over = t.cron_job(oracle.MODE_STORE, TODAY, TODAY + 3000000000, 0, string_eq)
        assert not threw and listed.tolist() == [8, 0, 9] and n_over == 3
        c, o = t.flags()
        assert o.tolist() == [1, 0, 0, 0, 0, 1, 0, 0, 1, 1, 0]
        assert c.tolist() == [0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0]
        # second tick the same day: nothing left (idempotent)
        threw, listed, n_over = t.cron_job(oracle.MODE_STORE, TODAY, TODAY + 3000000000, 0, string_eq)
        assert listed.size == 0 and n_over == 0


def test_cron_middle_predicate_utc_vs_local_skew(oracle):
    # Svc/Controllers/ScheduledTasksManagerController.cs:34 — runAt is UTC, Today is local.
    # Local zone ahead of UTC: at local 00:05 the UTC date is still yesterday => nothing survives.
    t = build_kat(oracle)
    run_at_utc = TODAY + 5 * 60 * 10_000_000 - 2 * 3600 * 10_000_000  # UTC+2 host
    threw, listed, n_over = t.cron_job(oracle.MODE_STORE, TODAY, run_at_utc, 120)
    assert listed.tolist() == [8, 0, 9] and n_over == 0 and not threw
    assert t.flags()[1].tolist() == [0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0]
    # FAKE mode with no survivors does not reach MarkOverdueTasks => no throw
    threw, listed, n_over = t.cron_job(oracle.MODE_FAKE, TODAY, run_at_utc, 120)
    assert not threw and n_over == 0 and listed.size == 7


def test_mark_overdue_by_id(oracle):
    t = build_kat(oracle)
    ids = np.stack([np.frombuffer(gid(i), dtype=np.uint8) for i in (2, 7, 7, 99)])
    threw, missing = t.mark_overdue(oracle.MODE_STORE, ids)
    assert not threw and missing == 1
    assert t.flags()[1].tolist() == [0, 1, 0, 0, 0, 1, 1, 0, 0, 0, 0]
    threw, _ = t.mark_overdue(oracle.MODE_FAKE, ids)
    assert threw  # Api/Services/FakeTasksManager.cs:104


def test_get_by_creator_order_and_ties(oracle):
    t = oracle.OracleTable()
    who = ["ann@x.io", "bob@x.io", "ann@x.io", "Ann@x.io", "ann@x.io", "ann@x.io ", "ann@x.io"]
    created = [5, 9, 7, 8, 7, 1, 2]
    for i, (w, c) in enumerate(zip(who, created)):
        t.add(gid(i + 1), "n", w, "z", raw(TODAY + c, UTC), raw(YDAY), False, False)
    for mode in (oracle.MODE_FAKE, oracle.MODE_STORE):
        # descending, stable: rows 2 and 4 tie and keep source order
        assert t.get_by_creator(mode, "ann@x.io").tolist() == [2, 4, 0, 6]
        assert t.get_by_creator(mode, "bob@x.io").tolist() == [1]
        assert t.get_by_creator(mode, "nobody").size == 0
        assert t.get_by_creator(mode, "").size == 0


def test_string_eq_equals_arithmetic_form_on_synthetic_rows(oracle):
    cfg = oracle.default_cfg(0x5EED0001, 20000, n_creators=50)
    t = oracle.OracleTable()
    t.add_synth(cfg, 0, 20000)
    a = t.get_yesterdays_due(oracle.MODE_STORE, TODAY, 60, False)
    b = t.get_yesterdays_due(oracle.MODE_STORE, TODAY, 60, True)
    assert a.size > 300 and (a == b).all()
    f = t.get_yesterdays_due(oracle.MODE_FAKE, TODAY)
    assert set(a.tolist()) < set(f.tolist())  # FAKE ignores flags and Kind


def test_stream_checker_matches_table_run(oracle):
    n = 50000
    cfg = oracle.default_cfg(0x5EED0002, 4 * n)
    for first in (0, 3 * n):
        for mode in (oracle.MODE_FAKE, oracle.MODE_STORE):
            t = oracle.OracleTable()
            t.add_synth(cfg, first, n)
            ids, _, _ = t.export()
            threw, listed, n_over = t.cron_job(mode, TODAY, oracle.DEFAULT_RUN_AT)
            s = oracle.stream_sweep(cfg, first, n, mode, threads=3)
            assert s.n_listed == listed.size and s.n_marked == n_over and bool(s.threw) == threw
            hi = np.bitwise_xor.reduce(ids[listed][:, :8].copy().view(">u8").ravel()) if listed.size else 0
            lo = np.bitwise_xor.reduce(ids[listed][:, 8:].copy().view(">u8").ravel()) if listed.size else 0
            assert int(hi) == s.xor_id_hi and int(lo) == s.xor_id_lo
            assert int(listed.astype(np.uint64).sum()) == s.sum_rows
            c, o = t.flags()
            assert int(o.sum()) == s.overdue_after and int(c.sum()) == s.completed


def test_synthetic_rows_shape(oracle):
    n = 100000
    cfg = oracle.default_cfg(7, n)
    t = oracle.OracleTable()
    t.add_synth(cfg, 0, n)
    ids, created, due = t.export()
    # TaskIds: v4 shape, strictly increasing in canonical byte order (sorted table)
    assert ((ids[:, 6] >> 4) == 4).all() and ((ids[:, 8] >> 6) == 2).all()
    hi = ids[:, :8].copy().view(">u8").ravel().astype(np.uint64)
    assert (np.diff(hi.astype(np.float64)) > 0).all()
    day = (due & np.uint64((1 << 62) - 1)) // np.uint64(TPD)
    yd = TODAY // TPD - 1
    frac = (day == yd).mean()
    assert 0.055 < frac < 0.065  # 5 % exact + 1 % adversarial
    assert int(day.min()) >= yd - 364 and int(day.max()) <= yd + 366
    c, o = t.flags()
    assert 0.49 < c.mean() < 0.51
    assert (o[day >= yd] == 0).all() and 0.03 < o.mean() < 0.07
    cr = created & np.uint64((1 << 62) - 1)
    assert (cr < TODAY).all() and (cr >= TODAY - 730 * TPD).all()
    dup = (cr[1:] == cr[:-1]).mean()
    assert 0.0005 < dup < 0.002
    assert t.created_by(0).startswith(b"user") and t.created_by(0).endswith(b".example")


def test_soa_multicore_sweep_agrees(oracle):
    n = 200000
    cfg = oracle.default_cfg(11, n)
    s = oracle.SoaTable(cfg, 0, n, threads=4)
    cnt, rows = s.sweep()
    ref = oracle.stream_sweep(cfg, 0, n, oracle.MODE_STORE)
    assert cnt == ref.n_listed and int(np.sort(rows).astype(np.uint64).sum()) == ref.sum_rows
    cnt2, _ = s.sweep()
    assert cnt2 == 0


def test_point_operations_follow_the_two_implementations(oracle):
    # FakeTasksManager.cs:32-100 / TasksStoreManager.cs:22-102
    for mode in (oracle.MODE_FAKE, oracle.MODE_STORE):
        t = build_kat(oracle)
        missing = gid(200)
        assert t.get_by_id(mode, gid(3)) == 2 and t.get_by_id(mode, missing) == -1
        assert t.mark_completed(mode, gid(1)) == 1 and t.mark_completed(mode, missing) == 0
        assert t.flags()[0][0] == 1
        # completed row 0 no longer listed by the store implementation; the fake one ignores flags
        assert t.get_yesterdays_due(mode, TODAY).tolist() == ([0, 2, 3, 4, 5, 8, 9] if mode == oracle.MODE_FAKE else [8, 9])
        assert t.update_task(mode, gid(8), "n", "a", raw(YDAY)) == 1   # row 7 (due today) moves to yesterday
        assert 7 in t.get_yesterdays_due(mode, TODAY).tolist()
        # a missing id: false from the fake list, NullReferenceException from the store code (:88)
        assert t.update_task(mode, missing, "n", "a", raw(YDAY)) == (0 if mode == oracle.MODE_FAKE else 2)
        # DeleteTask: found/not found from the list; always true from the store
        assert t.delete_task(mode, gid(9)) == 1 and t.get_by_id(mode, gid(9)) == -1
        assert t.delete_task(mode, gid(9)) == (0 if mode == oracle.MODE_FAKE else 1)
        assert len(t) == 11 and t.live_count() == 10
        assert 8 not in t.get_yesterdays_due(mode, TODAY).tolist()
        # CreateNewTask appends at the end of the list
        row = t.add(gid(50), "new", "a@b.c", "x", raw(TODAY - 5, UTC), raw(YDAY))
        assert row == 11 and t.get_yesterdays_due(oracle.MODE_FAKE, TODAY).tolist()[-1] == 11
        threw, listed, n_over = t.cron_job(mode, TODAY, TODAY + 3000000000)
        assert threw == (mode == oracle.MODE_FAKE) and n_over == listed.size


# ---- the JSON form of a task list (System.Text.Json restatement) ------------------------------------
def _doc_sample_table(oracle):
    """the body documented at docs/aca/05-aca-dapr-pubsubapi/index.md:60-69"""
    import uuid

    def ticks(y, mo, d, h=0, mi=0, s=0, frac=0):
        import datetime

        days = (datetime.date(y, mo, d) - datetime.date(1, 1, 1)).days
        return (days * 86400 + h * 3600 + mi * 60 + s) * 10_000_000 + frac

    ot = oracle.OracleTable()
    ot.add(np.frombuffer(uuid.UUID("fbc55b2c-d9fa-405e-aec8-22e53f4306dd").bytes, dtype=np.uint8).copy(),
           b"Testing Pub Sub Publisher", b"user@mail.net", b"user2@mail.com",
           ticks(2023, 2, 12, 0, 24, 37, 7361348) | (1 << 62), ticks(2023, 2, 20))
    return ot


def test_json_matches_the_documented_sample_body(oracle):
    import json

    ot = _doc_sample_table(oracle)
    for flavor in (oracle.JSON_RESPONSE, oracle.JSON_DOCUMENT):
        text = ot.to_json([0], flavor)
        assert text == (b'[{"taskId":"fbc55b2c-d9fa-405e-aec8-22e53f4306dd","taskName":"Testing Pub Sub Publisher",'
                        b'"taskCreatedBy":"user@mail.net","taskCreatedOn":"2023-02-12T00:24:37.7361348Z",'
                        b'"taskDueDate":"2023-02-20T00:00:00","taskAssignedTo":"user2@mail.com",'
                        b'"isCompleted":false,"isOverDue":false}]')
        doc = json.loads(text)[0]
        sample = {"taskId": "fbc55b2c-d9fa-405e-aec8-22e53f4306dd", "taskName": "Testing Pub Sub Publisher",
                  "taskCreatedBy": "user@mail.net", "taskCreatedOn": "2023-02-12T00:24:37.7361348Z",
                  "taskDueDate": "2023-02-20T00:00:00", "taskAssignedTo": "user2@mail.com"}
        assert {k: doc[k] for k in sample} == sample


def test_json_encoders_differ_where_dotnet_says_they_do(oracle):
    import json
    import uuid

    ot = oracle.OracleTable()
    name = "a\"b\\c<d>&'+`\t\x01\x7f caf\u00e9 \u2028 \U0001F600"
    ot.add(np.frombuffer(uuid.UUID(int=5).bytes, dtype=np.uint8).copy(), name.encode(), b"c", b"", 0, 0)
    relaxed = ot.to_json([0], oracle.JSON_RESPONSE).decode()
    default = ot.to_json([0], oracle.JSON_DOCUMENT).decode()
    assert '"taskName":"a\\"b\\\\c<d>&\'+`\\t\\u0001\\u007F caf\u00e9 \\u2028 \\uD83D\\uDE00"' in relaxed
    assert ('"taskName":"a\\u0022b\\\\c\\u003Cd\\u003E\\u0026\\u0027\\u002B\\u0060\\t\\u0001\\u007F caf\\u00E9 \\u2028 '
            '\\uD83D\\uDE00"') in default
    assert json.loads(relaxed)[0]["taskName"] == json.loads(default)[0]["taskName"] == name
    assert ot.to_json([], 0) == b"[]"
    # 0001-01-01T00:00:00 with Kind Unspecified: no fraction, no suffix
    assert '"taskCreatedOn":"0001-01-01T00:00:00","taskDueDate":"0001-01-01T00:00:00"' in relaxed


def _doc_samples():
    import json

    here = os.path.dirname(os.path.abspath(__file__))
    return json.load(open(os.path.join(here, "golden", "reference_doc_samples.json")))


def test_documents_from_the_reference_tutorial_deserialise_as_documented(oracle):
    """tests/golden/reference_doc_samples.json: the TaskModel bodies the workshop text shows (pub/sub body,
    Service Bus message, POST body, queue message).  Read by the restated deserializer and written again by
    the restated serializer, every member the tutorial shows must come back unchanged (dates included: they
    are already in System.Text.Json's canonical form), every member it omits must have its default."""
    import json

    samples = _doc_samples()
    texts = [s["text"].encode() for s in samples]
    off = np.zeros(len(texts) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(t) for t in texts])
    ot = oracle.OracleTable()
    assert ot.add_documents(b"".join(texts), off) == -1
    back = json.loads(ot.to_json(np.arange(len(texts)), oracle.JSON_RESPONSE))
    defaults = {"taskId": "00000000-0000-0000-0000-000000000000", "taskName": "", "taskCreatedBy": "",
                "taskCreatedOn": "0001-01-01T00:00:00", "taskDueDate": "0001-01-01T00:00:00", "taskAssignedTo": "",
                "isCompleted": False, "isOverDue": False}
    for s, doc in zip(samples, back):
        assert doc == {**defaults, **s["members"]}, s["source"]


def test_dates_agree_with_an_independent_calendar(oracle):
    """the restated DateTime -> text conversion against Python's proleptic Gregorian calendar (the same
    calendar .NET's DateTime uses): every 13th day of 0001-01-01 .. 9999-12-31 and both ends"""
    first = dt.date(1, 1, 1)
    day = 864_000_000_000
    for d in list(range(0, 3_652_059, 13)) + [3_652_058]:
        want = (first + dt.timedelta(days=d)).isoformat()
        assert oracle.format_stj(d * day + 1) == f"{want}T00:00:00.0000001"
        assert oracle.format_converter(d * day + 86_399 * 10_000_000) == f"{want}T23:59:59"


def test_every_scalar_survives_an_independent_json_reader(oracle):
    """all Unicode scalar values, 16 per TaskName, written by both restated encoders and read by Python's
    json module (an independent reader): whatever is escaped, the text is JSON and means the same string;
    the restated reader takes its own documents back unchanged"""
    import json

    scalars = [c for c in range(0x110000) if not 0xD800 <= c <= 0xDFFF]
    names = ["".join(map(chr, scalars[i:i + 16])) for i in range(0, len(scalars), 16)]
    ot = oracle.OracleTable()
    zero = np.zeros(16, dtype=np.uint8)
    for nm in names:
        ot.add(zero, nm.encode(), b"u@x", b"", 0, 0)
    rows = np.arange(len(names))
    for flavor in (oracle.JSON_RESPONSE, oracle.JSON_DOCUMENT):
        text = ot.to_json(rows, flavor)
        assert [d["taskName"] for d in json.loads(text)] == names
        if flavor == oracle.JSON_DOCUMENT:
            assert max(text) < 0x7F and min(text) >= 0x20  # Default encoder: plain printable ASCII only
        docs, off = ot.to_documents(rows, flavor)
        again = oracle.OracleTable()
        assert again.add_documents(docs, off) == -1 and again.to_json(rows, flavor) == text


def test_restated_reader_on_handwritten_and_malformed_documents(oracle):
    """the same hand-written and malformed documents the GPU loader is tested with (tests/test_ingest_gpu.py),
    through the oracle alone: what it accepts, what it rejects, and what it makes of the accepted ones"""
    import json

    from test_ingest_gpu import BAD, HAND, pack

    ot = oracle.OracleTable()
    assert ot.add_documents(*pack(HAND), 120) == -1 and len(ot) == len(HAND)
    docs = json.loads(ot.to_json(np.arange(len(HAND)), oracle.JSON_RESPONSE, 120))
    assert docs[0]["taskId"] == "fbc55b2c-d9fa-405e-aec8-22e53f4306dd" and docs[0]["isCompleted"] is True
    assert docs[1]["taskName"] == "q\"b\\s/\b\f\n\r\t\u00e9\u4e2d\U0001F600 end"
    assert docs[2]["taskName"] == "second" and docs[2]["isCompleted"] is False
    assert docs[2]["taskCreatedBy"] is None and docs[2]["taskAssignedTo"] is None  # null stays null; missing members are ""
    assert ot.get_by_creator(oracle.MODE_STORE, b"").tolist() == [7]               # the store's EQ "" does not match a null
    with pytest.raises(RuntimeError):                                              # FakeTasksManager.cs:70: .Equals on null
        ot.get_by_creator(oracle.MODE_FAKE, b"a@b")
    assert docs[3]["taskCreatedOn"] == "2020-01-01T10:30:00" and docs[3]["taskDueDate"] == "2026-11-05T00:00:00.5"
    assert docs[4]["taskCreatedOn"] == "2020-01-01T10:30:15.1234567"          # 16 digits truncated to 100 ns
    assert docs[4]["taskDueDate"] == "2026-11-05T02:00:00+02:00"              # +02:00 read on a +02:00 clock: same wall time
    assert docs[5]["taskCreatedOn"] == "2020-01-01T17:30:15.1+02:00"          # 10:30:15.1-05 is 15:30:15.1Z
    assert docs[5]["taskDueDate"] == "2026-11-05T02:00:00+02:00"              # 18:30-05:30 is midnight UTC
    assert docs[6]["taskDueDate"] == "9999-12-31T23:59:59.9999999"
    assert docs[7] == {"taskId": "00000000-0000-0000-0000-000000000000", "taskName": "", "taskCreatedBy": "",
                       "taskCreatedOn": "0001-01-01T00:00:00", "taskDueDate": "0001-01-01T00:00:00", "taskAssignedTo": "",
                       "isCompleted": False, "isOverDue": False}
    assert docs[8]["isOverDue"] is True and docs[9]["taskDueDate"] == "2026-11-05T00:00:00Z"
    good = b'{"taskCreatedBy":"a@b"}'
    for bad, _ in BAD:
        fresh = oracle.OracleTable()
        assert fresh.add_documents(*pack([good, bad, good])) == 1 and len(fresh) == 0, bad


def test_backlog_extension_known_answers(oracle):
    """the generalised predicate (oracle.cpp orc_backlog; an extension): twelve hand-checked rows"""
    import uuid

    day = 864_000_000_000
    today = oracle.DEFAULT_TODAY
    y = today - day
    dues = [y, y + 1, y | (1 << 62), (y + 5 * 36_000_000_000) | (2 << 62), y - 400 * day, 1 * day, 0,
            today, today + 1, today + 30 * day, y + day - 1, y - 1]
    ot = oracle.OracleTable()
    for i, d in enumerate(dues):
        ot.add(uuid.UUID(int=i + 1).bytes, "n", "c@d", "a", (700_000 * day + (len(dues) - i)) | (1 << 62), d,
               completed=(i == 4), overdue=(i == 6))
    rows, n_over, threw = ot.backlog(oracle.MODE_FAKE)
    assert rows.tolist() == [0, 1, 2, 3, 4, 5, 6, 10, 11] and n_over == 9 and not threw   # list order, no flag test
    rows, n_over, threw = ot.backlog(oracle.MODE_STORE)
    assert rows.tolist() == [11, 10, 5, 3, 2, 1, 0] and n_over == 7                       # pending only, by TaskCreatedOn
    rows, n_over, threw = ot.backlog(oracle.MODE_FAKE, mark=True)
    assert threw and ot.flags()[1].tolist() == [0] * 6 + [1] + [0] * 5                    # FAKE's Mark throws, nothing changes
    rows, n_over, threw = ot.backlog(oracle.MODE_STORE, mark=True)
    assert not threw and n_over == 7 and ot.flags()[1].tolist() == [1, 1, 1, 1, 0, 1, 1, 0, 0, 0, 1, 1]
    assert ot.backlog(oracle.MODE_STORE)[0].size == 0                                      # idempotent


def test_streaming_creator_filter_agrees_with_the_list_oracle(oracle):
    """orc_stream_creators (the full-size checker of the GPU creator filter) against GetTasksByCreator over a
    stored List<TaskModel> of the same synthetic rows."""
    n = 300_000
    cfg = oracle.default_cfg(0x5EED0003, n, n_creators=3000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    ids, _, _ = ot.export()
    queries = [ot.created_by(i) for i in (0, 5, 77, 5)] + [b"nobody@example.org", ot.created_by(9).upper()]
    for threads in (1, 3):
        out = oracle.stream_creators(cfg, 0, n, queries, threads=threads)
        for q, o in zip(queries, out):
            want = ot.get_by_creator(oracle.MODE_FAKE, q)
            assert o.count == len(want)
            if len(want):
                sel = ids[want]
                hi = int(np.bitwise_xor.reduce(np.ascontiguousarray(sel[:, :8]).view(">u8").ravel().astype(np.uint64)))
                lo = int(np.bitwise_xor.reduce(np.ascontiguousarray(sel[:, 8:]).view(">u8").ravel().astype(np.uint64)))
                assert (o.xor_id_hi, o.xor_id_lo) == (hi, lo)
                assert (o.sum_rows, o.min_row, o.max_row) == (int(want.astype(np.uint64).sum()), want.min(), want.max())
            else:
                assert (o.min_row, o.max_row) == (-1, -1)
    assert out[4].count == out[5].count == 0 and out[1].count == out[3].count > 50


def expected_relaxed(s):
    """JavaScriptEncoder.UnsafeRelaxedJsonEscaping on a string, computed from Python's own Unicode database
    (not from the table the oracle and the device share): \\" \\\\ and the five short escapes, a BMP code point as
    it is when its general category is a letter, mark, number, punctuation or symbol (or it is U+0020), anything
    else as \\uXXXX in upper-case hex, supplementary scalars as their two surrogates."""
    import unicodedata

    short = {'"': '\\"', "\\": "\\\\", "\b": "\\b", "\t": "\\t", "\n": "\\n", "\f": "\\f", "\r": "\\r"}
    out = []
    for ch in s:
        cp = ord(ch)
        if ch in short:
            out.append(short[ch])
        elif cp >= 0x10000:
            v = cp - 0x10000
            out.append("\\u%04X\\u%04X" % (0xD800 + (v >> 10), 0xDC00 + (v & 0x3FF)))
        elif cp == 0x20 or unicodedata.category(ch)[0] in "LMNPS":
            out.append(ch)
        else:
            out.append("\\u%04X" % cp)
    return "".join(out)


def test_relaxed_encoder_against_the_unicode_database(oracle):
    """every BMP scalar and a sample of the astral ones through the restated UnsafeRelaxedJsonEscaping, against
    expected_relaxed(); and the generated table (include/tt_bmp_defined.h) is up to date with this Python's UCD"""
    import subprocess
    import sys
    import unicodedata

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "tt_bmp_defined.h")).read()
    assert f'TT_BMP_UNICODE_VERSION "{unicodedata.unidata_version}"' in header, "run scripts/gen_unicode_bitmap.py"
    scalars = [c for c in range(0x10000) if not 0xD800 <= c <= 0xDFFF] + list(range(0x10000, 0x110000, 4099))
    names = ["".join(map(chr, scalars[i:i + 32])) for i in range(0, len(scalars), 32)]
    ot = oracle.OracleTable()
    zero = np.zeros(16, dtype=np.uint8)
    for nm in names:
        ot.add(zero, nm.encode(), b"", b"", 0, 0)
    text = ot.to_json(np.arange(len(names)), oracle.JSON_RESPONSE).decode()
    pieces = text.split('","taskCreatedBy":"')
    assert len(pieces) == len(names) + 1
    for nm, piece in zip(names, pieces):
        assert piece[piece.index('"taskName":"') + 12:] == expected_relaxed(nm)
    # spot checks of the classes the round-1 restatement got wrong or right
    for cp, raw_ok in ((0xAD, False), (0x200B, False), (0xFEFF, False), (0x378, False), (0xA0, False), (0x2028, False),
                       (0xE000, False), (0xFFFE, False), (0xE9, True), (0x4E2D, True), (0x20AC, True), (0x300, True),
                       (0x3C, True), (0x2B, True), (0x27, True)):
        ch = chr(cp)
        assert (expected_relaxed(ch) == ch) == raw_ok, hex(ord(ch))


def test_documented_dotnet_known_answers(oracle):
    """Facts the .NET documentation states, used as known answers for the restated BCL pieces:
    DateTime.UnixEpoch.Ticks and DateTime.MaxValue.Ticks (System.DateTime reference pages), DateTime(2000,1,1).Ticks,
    TimeSpan.TicksPerDay; Guid.ToByteArray() byte order against ToString("D") (System.Guid.ToByteArray remarks: the
    first three groups are little-endian in the array) — this library keeps ids in the order of the "D" string;
    the JavaScriptEncoder.Default escapes the System.Text.Json character-encoding article lists
    (< > & ' + as \\u003C \\u003E \\u0026 \\u0027 \\u002B, non-ASCII as \\uXXXX, the quote as \\u0022) and the relaxed
    encoder's \\" ."""
    import uuid

    assert TPD == 864_000_000_000
    assert ticks(1970, 1, 1) == 621_355_968_000_000_000            # DateTime.UnixEpoch.Ticks
    assert ticks(2000, 1, 1) == 630_822_816_000_000_000
    assert ticks(9999, 12, 31, 23, 59, 59, 9_999_999) == 3_155_378_975_999_999_999  # DateTime.MaxValue.Ticks
    assert oracle.format_stj(raw(3_155_378_975_999_999_999, UTC)) == "9999-12-31T23:59:59.9999999Z"
    assert oracle.format_stj(raw(621_355_968_000_000_000, UTC)) == "1970-01-01T00:00:00Z"
    # Guid "00112233-4455-6677-8899-aabbccddeeff": ToByteArray() gives 33 22 11 00 55 44 77 66 88 99 aa bb cc dd ee ff;
    # the canonical (string-order) bytes kept here are Python's UUID.bytes, the .NET array is UUID.bytes_le
    u = uuid.UUID("00112233-4455-6677-8899-aabbccddeeff")
    assert u.bytes_le.hex() == "33221100554477668899aabbccddeeff"
    assert oracle.guid_string(u.bytes) == "00112233-4455-6677-8899-aabbccddeeff"
    ot = oracle.OracleTable()
    ot.add(np.frombuffer(u.bytes, dtype=np.uint8).copy(), "<a href='x'>+&é\"".encode(), b"", b"", 0, 0)
    assert b'"taskName":"\\u003Ca href=\\u0027x\\u0027\\u003E\\u002B\\u0026\\u00E9\\u0022"' in ot.to_json([0], oracle.JSON_DOCUMENT)
    assert '"taskName":"<a href=\'x\'>+&é\\""'.encode() in ot.to_json([0], oracle.JSON_RESPONSE)
