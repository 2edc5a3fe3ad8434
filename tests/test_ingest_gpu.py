"""K7 — tt_load_documents: the state store's JSON documents parsed and dictionary-encoded on the device,
against the oracle's restatement of JsonSerializer.Deserialize<TaskModel> (oracle.cpp DeserializeTaskModel).
Equality is checked through everything the table can say about its rows: both sweeps, the flag planes,
creator filters and the JSON text of every row (K6), which carries all eight members."""
import json
import uuid

import numpy as np
import pytest

from helpers import tt

pytestmark = pytest.mark.gpu
DAY = tt.TICKS_PER_DAY


def same_table(oracle, ot, table, creators, offset=0):
    assert len(table) == len(ot)
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        res = table.get_yesterdays_due(mode)
        listed = ot.get_yesterdays_due(mode)
        assert res.rows().tolist() == listed.tolist()
        ids_all, _, _ = ot.export()
        assert (res.ids() == ids_all[listed]).all()
        assert res.to_json(tt.JSON_RESPONSE, offset) == ot.to_json(listed, tt.JSON_RESPONSE, offset)
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    assert (gc == c).all() and (go == o).all()
    seen = 0
    for who in creators:
        res = table.get_by_creator(who)
        listed = ot.get_by_creator(oracle.MODE_STORE, who)  # (the in-memory manager throws on a null TaskCreatedBy)
        assert res.rows().tolist() == listed.tolist()
        for flavor in (tt.JSON_RESPONSE, tt.JSON_DOCUMENT):
            assert res.to_json(flavor, offset) == ot.to_json(listed, flavor, offset)
        seen += res.count
    return seen


@pytest.mark.parametrize("flavor", [tt.JSON_RESPONSE, tt.JSON_DOCUMENT])
def test_documents_written_by_the_reference_load_back(oracle, flavor):
    n = 60_000
    cfg = oracle.default_cfg(0xD0C, n, n_creators=7, n_names=300)
    src = oracle.OracleTable()
    src.add_synth(cfg, 0, n)
    text, off = src.to_documents(np.arange(n), flavor, 120)
    ot = oracle.OracleTable()
    assert ot.add_documents(text, off, 120) == -1
    table = tt.TaskTable(0)
    st = table.load_documents(text, off, 120)
    assert (st.n_docs, st.first_bad, st.text_bytes) == (n, -1, len(text))
    assert (st.n_creators, st.n_names, st.n_assignees) == (7, 300, 50)
    creators = [src.created_by(i) for i in range(40)]
    assert same_table(oracle, ot, table, set(creators), 120) == n  # the 7 creators cover every row
    # the loaded table behaves like any other: a cron tick marks what it lists
    res, n_marked, _ = table.sweep(tt.MODE_STORE)
    _, listed, n_over = ot.cron_job(tt.MODE_STORE)
    assert res.count == listed.size == n_marked == n_over > 500
    assert (table.flags_bytes()[1] == ot.flags()[1]).all()
    table.close()


HAND = [
    # member order, whitespace, unknown members (nested, with braces inside strings), upper-case names
    b' {\r\n "ISCOMPLETED" : true ,\t"extra": {"a": [1, 2.5e3, {"b": "}]\\"{"}], "c": null},\n'
    b'  "taskId":"FBC55B2C-d9fa-405e-AEC8-22e53f4306dd", "TaskName":"order and case", "taskCreatedBy":"a@b",'
    b' "taskCreatedOn":"2023-02-12T00:24:37.7361348Z", "taskDueDate":"2026-11-05", "taskAssignedTo":"x", "isOverDue":false } ',
    # escapes in values and in a member name, surrogate pair, solidus, every short escape
    b'{"taskId":"00000000-0000-0000-0000-000000000002","task\\u004eame":"q\\"b\\\\s\\/\\b\\f\\n\\r\\t\\u00e9\\u4e2d\\ud83d\\ude00 end",'
    b'"taskCreatedBy":"a@b","taskCreatedOn":"2024-02-29T23:59:59","taskDueDate":"2026-11-05T00:00:00","taskAssignedTo":"caf\xc3\xa9 \xe4\xb8\xad \xf0\x9f\x98\x80"}',
    # repeated member: the last value wins; null strings; missing members keep their defaults
    b'{"taskName":"first","taskName":"second","taskCreatedBy":null,"taskAssignedTo":null,"isCompleted":true,"isCompleted":false}',
    # date forms: no seconds, comma fraction, 16 fraction digits, offsets with and without minutes
    b'{"taskId":"00000000-0000-0000-0000-000000000004","taskCreatedBy":"a@b","taskCreatedOn":"2020-01-01T10:30","taskDueDate":"2026-11-05T00:00:00,5"}',
    b'{"taskId":"00000000-0000-0000-0000-000000000005","taskCreatedBy":"c@d","taskCreatedOn":"2020-01-01T10:30:15.1234567890123456","taskDueDate":"2026-11-05T02:00:00+02:00"}',
    b'{"taskId":"00000000-0000-0000-0000-000000000006","taskCreatedBy":"c@d","taskCreatedOn":"2020-01-01T10:30:15.1-05","taskDueDate":"2026-11-04T18:30:00-05:30"}',
    b'{"taskId":"00000000-0000-0000-0000-000000000007","taskCreatedBy":"c@d","taskCreatedOn":"0001-01-01T00:00:00","taskDueDate":"9999-12-31T23:59:59.9999999"}',
    b'{}',
    b' [ {"taskCreatedBy":"a@b","isOverDue":true} , ',
    b',{"taskCreatedBy":"a@b","taskDueDate":"2026-11-05T00:00:00Z"}]',
]


def pack(docs):
    off = np.zeros(len(docs) + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(d) for d in docs])
    return b"".join(docs), off


@pytest.mark.parametrize("offset", [0, 120])
def test_handwritten_documents(oracle, offset):
    text, off = pack(HAND)
    ot = oracle.OracleTable()
    assert ot.add_documents(text, off, offset) == -1
    table = tt.TaskTable(0)
    st = table.load_documents(text, off, offset)
    assert st.n_docs == len(HAND) and st.n_creators == 3  # "a@b", "c@d" and "" (missing; a null shares its slot)
    # document 2's TaskCreatedBy is null: no creator query finds it, not even the empty string
    assert same_table(oracle, ot, table, [b"a@b", b"c@d", b""], offset) == len(HAND) - 1
    doc = json.loads(table.get_by_creator(b"a@b").to_json())
    assert {d["taskName"] for d in doc} >= {"order and case", "q\"b\\s/\b\f\n\r\té中\U0001F600 end", ""}
    # null string members stay null (System.Text.Json keeps them; the initialisers only cover missing members)
    for flavor in (tt.JSON_RESPONSE, tt.JSON_DOCUMENT):
        text2 = table.get_by_id(bytes(16)).to_json(flavor, offset)
        assert text2 == ot.to_json(np.array([2], dtype=np.uint32), flavor, offset)
        row2 = json.loads(text2)[0]
        assert row2["taskName"] == "second" and row2["taskCreatedBy"] is None and row2["taskAssignedTo"] is None
    assert table.get_by_id(bytes(16)).field(tt.FIELD_CREATOR_REF).tolist() == [0xFFFFFFFE]
    # and survive a trip through the device's own writer and reader
    everything = table.get_due_on_or_before(tt.MODE_FAKE, tt.DEFAULT_TODAY + 10**6 * tt.TICKS_PER_DAY)
    j = everything.json(tt.JSON_DOCUMENT, offset)
    again = tt.TaskTable(0)
    again.load_documents(j.bytes(), j.offsets(everything.count), offset)
    assert again.get_due_on_or_before(tt.MODE_FAKE, tt.DEFAULT_TODAY + 10**6 * tt.TICKS_PER_DAY).to_json(tt.JSON_DOCUMENT, offset) == j.bytes()
    again.close()
    table.close()


BAD = [
    (b'{"taskId":"fbc55b2c-d9fa-405e-aec8-22e53f4306d"}', 2),              # 35 characters
    (b'{"taskId":"{fbc55b2c-d9fa-405e-aec8-22e53f4306dd}"}', 2),            # "B" form
    (b'{"taskDueDate":"2026-02-30"}', 3),                                   # no such day
    (b'{"taskDueDate":"2026-11-05T24:00:00"}', 3),
    (b'{"taskDueDate":"2026-11-05 00:00:00"}', 3),                          # blank instead of T
    (b'{"taskDueDate":"2026-11-05T00:00:00.Z"}', 3),                        # empty fraction
    (b'{"taskName":"bad \\x escape"}', 4),
    (b'{"taskName":"lone \\ud83d surrogate"}', 4),
    (b'{"taskName":"raw \x01 control"}', 4),
    (b'{"taskName":"cut \xe4\xb8"}', 4),
    (b'{"isCompleted":"true"}', 5),
    (b'{"taskName":5}', 5),
    (b'{"taskId":null}', 5),
    (b'{"taskName":"x",}', 1),                                              # trailing comma
    (b'{"taskName":"x"} extra', 1),
    (b'["taskName"]', 1),
    (b'{"taskName" "x"}', 1),
    (b'', 1),
]


def test_rejected_documents_leave_the_table_empty(oracle):
    good = b'{"taskCreatedBy":"a@b"}'
    for bad, reason in BAD:
        docs = [good, good, bad, good]
        text, off = pack(docs)
        assert oracle.OracleTable().add_documents(text, off) == 2, bad
        table = tt.TaskTable(0)
        st = table.load_documents(text, off, allow_bad=True)
        assert (st.first_bad, st.first_bad_reason) == (2, reason), bad
        with pytest.raises(tt.TaskTableError):
            table.load_documents(text, off)
        assert len(table) == 0
        st = table.load_documents(*pack([good, good, good]))  # and the table can still be loaded
        assert st.n_docs == 3 and table.get_by_creator(b"a@b").count == 3
        table.close()


def test_offsets_that_do_not_describe_the_text_are_bad_documents(oracle):
    good = b'{"taskCreatedBy":"a@b"}'
    text, off = pack([good] * 4)
    for broken, bad in ((off.copy(), 1), (off.copy(), 2)):
        if bad == 1:
            broken[2] = broken[1] - 3        # runs backwards: documents 1 ends before it starts
        else:
            broken[3] = broken[4] + 1000     # document 2 would end beyond the text
        table = tt.TaskTable(0)
        st = table.load_documents(text, broken, allow_bad=True)
        assert st.first_bad == bad and st.first_bad_reason == 1 and len(table) == 0
        table.close()


def test_empty_load(oracle):
    table = tt.TaskTable(0)
    st = table.load_documents(b"", [0])
    assert st.n_docs == 0 and len(table) == 0 and table.get_yesterdays_due(tt.MODE_STORE).count == 0
    table.close()


def test_device_text_round_trip_at_scale(oracle):
    """every row of a 1.5 M-row table listed, written as documents by K6 on the device, loaded by K7
    straight from device memory, and written again: the two texts must be identical"""
    n = 1_500_000
    cfg = oracle.default_cfg(0xD0D, n, n_creators=20_000, ppm_due_yesterday=1_000_000, ppm_adversarial=0)
    a = tt.TaskTable(0).generate(cfg, 0, n).generate_payload(cfg, 0)
    res = a.get_yesterdays_due(tt.MODE_FAKE)
    assert res.count == n
    j = res.json(tt.JSON_DOCUMENT)
    off = j.offsets(n)
    assert off[0] == 0 and off[-1] == j.size and (np.diff(off.astype(np.int64)) > 160).all()
    import torch

    off_dev = torch.from_numpy(off.astype(np.int64)).cuda()  # offsets may live on the device as well
    b = tt.TaskTable(0)
    st = b.load_documents(j.device_ptr(), (off_dev.data_ptr(), n))
    assert (st.n_docs, st.first_bad, st.n_assignees) == (n, -1, 50) and st.n_creators <= 20_000
    want = j.bytes()
    j.free()
    resb = b.get_yesterdays_due(tt.MODE_FAKE)
    assert resb.count == n and (resb.ids() == res.ids()).all()
    assert resb.to_json(tt.JSON_DOCUMENT) == want
    assert b.count_flags() == a.count_flags()
    ra, ma, _ = a.sweep(tt.MODE_STORE)
    rb, mb, _ = b.sweep(tt.MODE_STORE)
    assert ma == mb > 100_000 and (ra.rows() == rb.rows()).all()
    who = json.loads(want[: off[1]].lstrip(b"[").rstrip(b","))["taskCreatedBy"].encode()
    assert a.get_by_creator(who).rows().tolist() == b.get_by_creator(who).rows().tolist()
    a.close()
    b.close()


def test_documents_out_of_taskid_order_get_a_device_built_index(oracle):
    """a list ordered by TaskCreatedOn is not in TaskId order: lookups by id then go through the index
    tt_load_documents builds with two radix sorts on the device"""
    n = 120_000
    cfg = oracle.default_cfg(0xD0E, n, n_creators=11, ppm_due_yesterday=700_000, ppm_completed=0)
    src = oracle.OracleTable()
    src.add_synth(cfg, 0, n)
    listed = src.get_yesterdays_due(oracle.MODE_STORE)  # ascending TaskCreatedOn
    assert listed.size > 50_000 and (np.diff(listed.astype(np.int64)) < 0).any()
    text, off = src.to_documents(listed, tt.JSON_DOCUMENT)
    ot = oracle.OracleTable()
    assert ot.add_documents(text, off) == -1
    table = tt.TaskTable(0)
    table.load_documents(text, off)
    ids, _, _ = ot.export()
    rng = np.random.default_rng(3)
    pick = rng.choice(len(ot), size=5000, replace=False)
    for r in pick[:50]:
        got = table.get_by_id(ids[r])
        assert got.rows().tolist() == [int(r)]
    unknown = np.frombuffer(uuid.UUID(int=12345).bytes, dtype=np.uint8).copy()
    assert table.get_by_id(unknown).count == 0
    marked, missing = table.mark_overdue(tt.MODE_STORE, np.concatenate([ids[pick], unknown[None, :]]))
    ot.mark_overdue(oracle.MODE_STORE, ids[pick])
    assert (marked, missing) == (5000, 1)
    assert (table.flags_bytes()[1] == ot.flags()[1]).all()
    assert table.mark_completed(ids[pick[7]]) and ot.mark_completed(oracle.MODE_STORE, ids[pick[7]])
    assert (table.flags_bytes()[0] == ot.flags()[0]).all()
    table.close()


def test_documents_from_the_reference_tutorial(oracle):
    """the golden fixture tests/golden/reference_doc_samples.json through K7 and back out through K6"""
    import os

    here = os.path.dirname(os.path.abspath(__file__))
    samples = json.load(open(os.path.join(here, "golden", "reference_doc_samples.json")))
    text, off = pack([s["text"].encode() for s in samples])
    table = tt.TaskTable(0)
    st = table.load_documents(text, off)
    assert st.n_docs == len(samples) and st.first_bad == -1
    ot = oracle.OracleTable()
    assert ot.add_documents(text, off) == -1
    defaults = {"taskId": "00000000-0000-0000-0000-000000000000", "taskName": "", "taskCreatedBy": "",
                "taskCreatedOn": "0001-01-01T00:00:00", "taskDueDate": "0001-01-01T00:00:00", "taskAssignedTo": "",
                "isCompleted": False, "isOverDue": False}
    seen = 0
    for who in {s["members"]["taskCreatedBy"].encode() for s in samples}:
        res = table.get_by_creator(who)
        listed = ot.get_by_creator(oracle.MODE_STORE, who)  # (the in-memory manager throws on a null TaskCreatedBy)
        assert res.rows().tolist() == listed.tolist()
        for row, doc in zip(res.rows(), json.loads(res.to_json())):
            assert doc == {**defaults, **samples[int(row)]["members"]}, samples[int(row)]["source"]
            seen += 1
    assert seen == len(samples)
    table.close()


@pytest.mark.parametrize("env", [{"TT_DICT_SORTED": "1"}, {"TT_DICT_KEY_BITS": "5"}, {}])
def test_both_dictionary_encoders_give_the_same_table(oracle, env, monkeypatch):
    """The hash-table encoder, the sorting encoder it falls back to, and the fall-back itself (hashes cut to 5 bits,
    so that different strings meet in one slot of the table) load the same table."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    n = 40_000
    cfg = oracle.default_cfg(0xD1C, n, n_creators=900, n_names=5_000)
    src = oracle.OracleTable()
    src.add_synth(cfg, 0, n)
    text, off = src.to_documents(np.arange(n), tt.JSON_DOCUMENT, -300)
    ot = oracle.OracleTable()
    assert ot.add_documents(text, off, -300) == -1
    table = tt.TaskTable(0)
    st = table.load_documents(text, off, -300)
    assert (st.n_docs, st.first_bad) == (n, -1)
    assert (st.n_creators, st.n_names, st.n_assignees) == (900, 5_000, 50)
    creators = {src.created_by(i) for i in range(0, n, 997)}
    assert same_table(oracle, ot, table, creators, -300) > 40
    table.close()


@pytest.mark.parametrize("stage", ["0", "4096", "11264"])
def test_parser_staging_through_shared_memory_or_not(oracle, stage, monkeypatch):
    """The parser reads a warp's 32 documents from a copy in shared memory; warps whose documents do not fit (here: a
    4 KB stage, and documents with long names) and TT_PARSE_STAGE=0 read global memory.  Same table either way,
    escapes (unescaped in the copy, written back) included."""
    monkeypatch.setenv("TT_PARSE_STAGE", stage)
    n = 30_000
    cfg = oracle.default_cfg(0xD2C, n, n_creators=40, n_names=2_000)
    src = oracle.OracleTable()
    src.add_synth(cfg, 0, n)
    for flavor in (tt.JSON_DOCUMENT, tt.JSON_RESPONSE):
        text, off = src.to_documents(np.arange(n), flavor, 60)
        ot = oracle.OracleTable()
        assert ot.add_documents(text, off, 60) == -1
        table = tt.TaskTable(0)
        st = table.load_documents(text, off, 60)
        assert (st.n_docs, st.first_bad, st.n_creators, st.n_names) == (n, -1, 40, 2_000)
        creators = {src.created_by(i) for i in range(0, n, 1499)}
        assert same_table(oracle, ot, table, creators, 60) > 20
        table.close()


def test_date_forms_in_documents_of_the_serializers_own_shape(oracle):
    """Documents in exactly the shape JsonSerializer writes take the fast parse kernel, whose date routine reads the
    serializer's own forms (seconds, 0-7 fraction digits, nothing / Z / +hh:mm) at fixed positions and hands every other
    form — and every value out of range — to the general routine.  All of them against the oracle, valid and invalid."""
    rng = np.random.default_rng(5)
    valid, forms = [], []
    for k in range(3000):
        y, mo, d = int(rng.integers(1, 10000)), int(rng.integers(1, 13)), int(rng.integers(1, 29))
        hh, mi, ss = int(rng.integers(0, 24)), int(rng.integers(0, 60)), int(rng.integers(0, 60))
        s = f"{y:04d}-{mo:02d}-{d:02d}T{hh:02d}:{mi:02d}:{ss:02d}"
        nd = int(rng.integers(0, 10))
        if nd:
            s += "." + "".join(str(int(x)) for x in rng.integers(0, 10, nd))
        s += ["", "Z", "+05:30", "-11:45", "+14:00", "+02"][int(rng.integers(0, 6))]
        if k % 7 == 0:
            s = [f"{y:04d}-{mo:02d}-{d:02d}", f"{y:04d}-{mo:02d}-{d:02d}T{hh:02d}:{mi:02d}", s.replace(".", ",", 1)][k % 3]
        forms.append(s)
    edge = ["2024-02-29T23:59:59.9999999Z", "0001-01-01T00:00:00", "9999-12-31T23:59:59.9999999", "2023-12-31T00:00:00.0000001+14:00"]
    forms += edge
    def doc(i, created, due="2026-11-05T00:00:00"):
        return (b'{"taskId":"%08x-0000-4000-8000-000000000000","taskName":"n%d","taskCreatedBy":"c%d@x","taskCreatedOn":"%s",'
                b'"taskDueDate":"%s","taskAssignedTo":"a","isCompleted":%s,"isOverDue":false}'
                % (i, i % 50, i % 9, created.encode(), due.encode(), b"true" if i % 3 == 0 else b"false"))
    docs = [doc(i, f) for i, f in enumerate(forms)] + [doc(len(forms) + i, "2020-05-05T05:05:05", f) for i, f in enumerate(forms[:400])]
    text, off = pack(docs)
    for offset in (0, -480):
        ot = oracle.OracleTable()
        bad = ot.add_documents(text, off, offset)
        if bad != -1:   # the oracle refuses some form (a year 0001 shifted before the calendar's start, say): drop it and go on
            keep = [i for i in range(len(docs)) if i != bad]
            while True:
                t2, o2 = pack([docs[i] for i in keep])
                ot = oracle.OracleTable()
                bad = ot.add_documents(t2, o2, offset)
                if bad == -1:
                    break
                del keep[bad]
            text_o, off_o = t2, o2
        else:
            text_o, off_o = text, off
        table = tt.TaskTable(0)
        st = table.load_documents(text_o, off_o, offset)
        assert st.first_bad == -1 and st.n_docs == len(off_o) - 1
        assert same_table(oracle, ot, table, [b"c%d@x" % k for k in range(9)], offset) == st.n_docs
        table.close()
    # a date out of range in an otherwise perfect document: refused with the date reason, like the oracle does
    wrong = [doc(0, "2021-02-29T00:00:00"), doc(1, "2021-01-01T24:00:00"), doc(2, "2021-13-01T00:00:00Z"), doc(3, "2021-01-01T00:00:60.5")]
    for k, dd in enumerate(wrong):
        text_w, off_w = pack(docs[:5] + [dd] + docs[5:9])
        ot = oracle.OracleTable()
        assert ot.add_documents(text_w, off_w, 0) == 5, k
        table = tt.TaskTable(0)
        st = table.load_documents(text_w, off_w, 0, allow_bad=True)
        assert (st.first_bad, st.first_bad_reason) == (5, 3), k
        table.close()
