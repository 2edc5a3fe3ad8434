"""K6 — a result written as the JSON the reference's controllers return (tt_result_json), against the
oracle's restatement of System.Text.Json (oracle.cpp JsonWriteTaskModel).  Byte-exact, both encoders:
JSON_RESPONSE = what OverdueTasksController.Get / TasksController.Get send (MVC output formatter),
JSON_DOCUMENT = what Dapr's client stores / publishes (Api/Services/TasksStoreManager.cs:35,155)."""
import json
import uuid

import numpy as np
import pytest

from helpers import load_from_oracle, tt

pytestmark = pytest.mark.gpu
FLAVORS = [tt.JSON_RESPONSE, tt.JSON_DOCUMENT]
KIND_UTC, KIND_LOCAL = 1 << 62, 2 << 62
DAY = tt.TICKS_PER_DAY


def with_payload(ot, table):
    n = len(ot)
    return table.load_payload([ot.string_field(i, 0) for i in range(n)], [ot.string_field(i, 2) for i in range(n)])


@pytest.mark.parametrize("mode", [tt.MODE_FAKE, tt.MODE_STORE])
def test_overdue_list_as_json_matches_the_oracle(oracle, mode):
    n = 300_000
    cfg = oracle.default_cfg(0x15A, n, n_creators=500, n_names=1000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n).generate_payload(cfg, 0)
    res = table.get_yesterdays_due(mode)
    listed = ot.get_yesterdays_due(mode)
    assert (res.rows() == listed).all() and res.count > 5_000
    for flavor in FLAVORS:
        got = res.to_json(flavor, local_offset_min=120)
        want = ot.to_json(listed, flavor, 120)
        assert len(got) == len(want)
        assert got == want
    st = table.last_json_stats()
    assert st.records == res.count and st.bytes == len(want) and st.launches == 3
    head = json.loads(got)[0]
    assert list(head) == ["taskId", "taskName", "taskCreatedBy", "taskCreatedOn", "taskDueDate", "taskAssignedTo",
                          "isCompleted", "isOverDue"]
    table.close()


def test_creator_list_as_json_matches_the_oracle(oracle):
    n = 200_000
    cfg = oracle.default_cfg(0x15B, n, n_creators=40, n_names=64)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n).generate_payload(cfg, 0)
    who = ot.created_by(17)
    res = table.get_by_creator(who)
    listed = ot.get_by_creator(oracle.MODE_FAKE, who)
    assert res.rows().tolist() == listed.tolist() and res.count > 1000
    for flavor in FLAVORS:
        assert res.to_json(flavor) == ot.to_json(listed, flavor)
    # flags are printed as they are NOW: complete one listed task, mark a batch overdue
    ids = res.ids()
    table.mark_completed(ids[3])
    ot.mark_completed(oracle.MODE_STORE, ids[3])
    table.mark_overdue(tt.MODE_STORE, ids[10:200])
    ot.mark_overdue(oracle.MODE_STORE, ids[10:200])
    assert res.to_json(tt.JSON_RESPONSE) == ot.to_json(listed, tt.JSON_RESPONSE)
    table.close()


# every class of character the two encoders treat differently (written with escapes on purpose)
NASTY = [
    "plain",
    "",
    "quote \" backslash \\ slash / apostrophe ' plus + lt < gt > amp & tick `",
    "controls \x00\x01\x08\t\n\x0b\x0c\r\x1f\x7f end",
    "latin caf\u00e9 \u00fc\u00df  greek \u03b1\u03b2  cjk \u4e2d\u6587  hebrew \u05d0",
    "c1 \u0080\u0085\u009f nbsp \u00a0 ogham \u1680 quads \u2000\u200a ls \u2028 ps \u2029 nnbsp \u202f mmsp \u205f ideo \u3000",
    "soft hyphen \u00ad zwj \u200d bom \ufeff private \ue000\uf8ff nonchar \ufdd0\ufffe\uffff replacement \ufffd",
    "emoji \U0001F600 math \U0001D538 last \U0010FFFF",
    "x" * 700,
    "\u4e2d" * 300 + "\"" * 50,
]


def nasty_table(oracle, n, long_names=False):
    ot = oracle.OracleTable()
    rng = np.random.default_rng(7)
    yesterday = tt.DEFAULT_TODAY - DAY
    for i in range(n):
        name = NASTY[i % len(NASTY)] if not long_names else NASTY[8 + i % 2] + str(i)
        creator = NASTY[(i // 3) % 8] + "@x"
        assignee = NASTY[(i * 7 + 2) % len(NASTY)]
        scale = 10 ** int(rng.integers(0, 8))  # fractions with 0..7 trailing zeros
        created = int(rng.integers(600_000, 739_000)) * DAY + int(rng.integers(0, DAY)) // scale * scale
        kind = [0, KIND_UTC, KIND_LOCAL][i % 3]
        due = (yesterday | [0, 0, KIND_UTC, KIND_LOCAL][i % 4]) if i % 5 else yesterday + 12345678
        ot.add(np.frombuffer(uuid.UUID(int=(int(rng.integers(0, 2**63)) << 64) | i).bytes, dtype=np.uint8).copy(),
               name.encode(), creator.encode(), assignee.encode(), created | kind, due,
               completed=bool(i % 2), overdue=bool(i % 7 == 0))
    return ot


@pytest.mark.parametrize("offset", [0, 120, -330])
def test_escaping_and_date_forms(oracle, offset):
    ot = nasty_table(oracle, 600)
    table = with_payload(ot, load_from_oracle(ot))
    res = table.get_yesterdays_due(tt.MODE_FAKE)  # tick equality: every row due exactly yesterday 00:00, any Kind
    listed = ot.get_yesterdays_due(tt.MODE_FAKE)
    assert res.count == listed.size > 300
    for flavor in FLAVORS:
        got, want = res.to_json(flavor, offset), ot.to_json(listed, flavor, offset)
        assert got == want
        docs = json.loads(got)  # and it is JSON: the strings round-trip
        assert [d["taskName"] for d in docs] == [ot.string_field(int(r), 0).decode() for r in listed]
        assert [d["taskAssignedTo"] for d in docs] == [ot.string_field(int(r), 2).decode() for r in listed]
    table.close()


def test_pieces_larger_than_the_staging_area_are_written_directly(oracle):
    """128 records of > 700 bytes each exceed the 44 KB a CTA stages in shared memory"""
    ot = nasty_table(oracle, 400, long_names=True)
    table = with_payload(ot, load_from_oracle(ot))
    res = table.get_yesterdays_due(tt.MODE_FAKE)
    listed = ot.get_yesterdays_due(tt.MODE_FAKE)
    for flavor in FLAVORS:
        got = res.to_json(flavor)
        assert got == ot.to_json(listed, flavor) and len(got) > 128 * 700
    table.close()


def test_empty_result_and_missing_payload(oracle):
    ot = oracle.OracleTable()
    for i in range(5):  # no TaskName / TaskAssignedTo: the reference's default is string.Empty
        ot.add(np.frombuffer(uuid.UUID(int=i + 1).bytes, dtype=np.uint8).copy(), b"", b"someone@mail.com", b"",
               (700_000 * DAY + i) | KIND_UTC, tt.DEFAULT_TODAY - DAY)
    table = load_from_oracle(ot)  # payload never loaded
    res = table.get_yesterdays_due(tt.MODE_FAKE)
    assert res.to_json() == ot.to_json(ot.get_yesterdays_due(tt.MODE_FAKE))
    none = table.get_by_creator(b"nobody")
    assert none.count == 0 and none.to_json() == b"[]" == ot.to_json([])
    # a row inserted later has no payload either
    new_id = np.frombuffer(uuid.UUID(int=99).bytes, dtype=np.uint8).copy()
    table.insert_task(new_id, 700_001 * DAY | KIND_UTC, tt.DEFAULT_TODAY - DAY, b"someone@mail.com")
    ot.add(new_id, b"", b"someone@mail.com", b"", 700_001 * DAY | KIND_UTC, tt.DEFAULT_TODAY - DAY)
    res = table.get_by_creator(b"someone@mail.com")
    assert res.count == 6 and res.to_json() == ot.to_json(ot.get_by_creator(oracle.MODE_FAKE, b"someone@mail.com"))
    table.close()


def test_malformed_utf8_is_refused(oracle):
    ot = oracle.OracleTable()
    ot.add(np.frombuffer(uuid.UUID(int=1).bytes, dtype=np.uint8).copy(), b"n", b"c", b"a", 700_000 * DAY,
           tt.DEFAULT_TODAY - DAY)
    table = load_from_oracle(ot)
    table.load_payload([b"broken \xe4\xb8"], [b"fine"])
    res = table.get_yesterdays_due(tt.MODE_FAKE)
    with pytest.raises(tt.TaskTableError) as e:
        res.to_json()
    assert e.value.code == tt.TT_ERR_ARG and "UTF-8" in str(e.value)
    table.close()


def test_larger_list_and_device_copy(oracle):
    """2 M rows, ~50 k hits: also copies the text device-to-device (tt_json_copy takes either)"""
    import torch

    n = 2_000_000
    cfg = oracle.default_cfg(0x15C, n, n_creators=10_000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n).generate_payload(cfg, 0)
    res, n_marked, _ = table.sweep(tt.MODE_STORE)  # row order, marks the listed rows
    _, listed, _ = ot.cron_job(tt.MODE_STORE)
    rows = np.sort(listed)
    j = res.json(tt.JSON_RESPONSE)
    buf = torch.zeros(j.size + 1, dtype=torch.uint8, device="cuda")
    j.copy_to(buf.data_ptr())
    got = bytes(buf[: j.size].cpu().numpy())
    assert int(buf[j.size]) == 0 and got == ot.to_json(rows, tt.JSON_RESPONSE) and b'"isOverDue":true' in got
    j.free()
    table.close()


def test_every_unicode_scalar_value(oracle):
    """all 1 112 064 Unicode scalar values (U+0000..U+10FFFF without the surrogates), 16 per TaskName, through
    both encoders: the product's escaping rule must agree with the oracle's on every one of them"""
    scalars = [c for c in range(0x110000) if not 0xD800 <= c <= 0xDFFF]
    names = ["".join(map(chr, scalars[i:i + 16])).encode("utf-8") for i in range(0, len(scalars), 16)]
    n = len(names)
    ot = oracle.OracleTable()
    ids = np.zeros((n, 16), dtype=np.uint8)
    ids[:, 12:] = np.arange(1, n + 1, dtype=">u4").view(np.uint8).reshape(n, 4)
    due = tt.DEFAULT_TODAY - DAY
    for i, name in enumerate(names):
        ot.add(ids[i], name, b"u@x", b"", (700_000 * DAY + i) | KIND_UTC, due)
    table = tt.TaskTable(0).load_soa(ids, np.array([(700_000 * DAY + i) | KIND_UTC for i in range(n)], dtype=np.uint64),
                                     np.full(n, due, dtype=np.uint64), np.zeros(n, dtype=np.uint8), np.zeros(n, dtype=np.uint8),
                                     np.zeros(n, dtype=np.uint32), [b"u@x"])
    table.load_payload(names, [b""] * n)
    res = table.get_yesterdays_due(tt.MODE_FAKE)
    listed = ot.get_yesterdays_due(tt.MODE_FAKE)
    assert res.count == n == listed.size
    for flavor in FLAVORS:
        got, want = res.to_json(flavor), ot.to_json(listed, flavor)
        assert got == want
        if flavor == tt.JSON_RESPONSE:
            # ... and, independently of the oracle (which shares the generated table), with the rule applied to
            # Python's own Unicode database: which BMP code points UnsafeRelaxedJsonEscaping writes as they are
            from test_oracle import expected_relaxed

            pieces = got.decode().split('","taskCreatedBy":"')
            assert len(pieces) == n + 1
            for name, piece in zip(names, pieces):
                assert piece[piece.index('"taskName":"') + 12:] == expected_relaxed(name.decode())
    back = json.loads(got)
    assert back[0]["taskName"] == names[0].decode() and back[-1]["taskName"] == names[-1].decode()
    # and the reader takes every one of them back: K6's documents through K7 print the same text again
    j = res.json(tt.JSON_DOCUMENT)
    again = tt.TaskTable(0)
    again.load_documents(j.device_ptr(), j.offsets(n))
    assert again.get_yesterdays_due(tt.MODE_FAKE).to_json(tt.JSON_DOCUMENT) == j.bytes()
    j.free()
    again.close()
    table.close()


def test_calendar_edges(oracle):
    """every day of the years where the Gregorian rules bite (1, 4, 100, 400, 1600, 1900, 2000, 2024, 2100, 9999)
    and every 97th day of the whole DateTime range, printed by K6 and read back by K7"""
    import datetime

    first = datetime.date(1, 1, 1)
    days = set(range(0, 3_652_059, 97))
    for y in (1, 4, 100, 400, 1600, 1900, 2000, 2024, 2100, 9999):
        d0 = (datetime.date(y, 1, 1) - first).days
        days.update(range(d0, d0 + (366 if y % 4 == 0 and (y % 100 != 0 or y % 400 == 0) else 365)))
    days = sorted(days)
    n = len(days)
    ids = np.zeros((n, 16), dtype=np.uint8)
    ids[:, 12:] = np.arange(1, n + 1, dtype=">u4").view(np.uint8).reshape(n, 4)
    due = tt.DEFAULT_TODAY - DAY
    created = np.array([d * DAY + (d * 7919 % DAY if d % 3 else 0) for d in days], dtype=np.uint64)
    created |= np.array([[0, KIND_UTC, KIND_LOCAL][d % 3] for d in days], dtype=np.uint64)
    ot = oracle.OracleTable()
    for i in range(n):
        ot.add(ids[i], b"", b"u@x", b"", int(created[i]), due)
    table = tt.TaskTable(0).load_soa(ids, created, np.full(n, due, dtype=np.uint64), np.zeros(n, dtype=np.uint8),
                                     np.zeros(n, dtype=np.uint8), np.zeros(n, dtype=np.uint32), [b"u@x"])
    res = table.get_yesterdays_due(tt.MODE_FAKE)
    listed = ot.get_yesterdays_due(tt.MODE_FAKE)
    assert res.count == n
    for offset in (0, 60):
        got = res.to_json(tt.JSON_RESPONSE, offset)
        assert got == ot.to_json(listed, tt.JSON_RESPONSE, offset)
    # spot checks against Python's own calendar
    docs = json.loads(res.to_json(tt.JSON_RESPONSE, 0))
    for k in (0, n // 3, n // 2, n - 1):
        want = (first + datetime.timedelta(days=days[k])).isoformat()
        assert docs[k]["taskCreatedOn"].startswith(f"{want}T"), (docs[k]["taskCreatedOn"], want)
    j = res.json(tt.JSON_DOCUMENT, 60)
    again = tt.TaskTable(0)
    again.load_documents(j.device_ptr(), j.offsets(n), 60)
    assert again.get_yesterdays_due(tt.MODE_FAKE).to_json(tt.JSON_DOCUMENT, 60) == j.bytes()
    j.free()
    again.close()
    table.close()
