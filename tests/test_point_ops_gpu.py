"""The single-row operations of ITasksManager (GetTaskById, CreateNewTask, UpdateTask,
MarkTaskCompleted, DeleteTask) on the device table, against the oracle — and the sweep and the
creator filter after a random mix of them."""
import uuid

import numpy as np
import pytest

from helpers import load_from_oracle, tt

pytestmark = pytest.mark.gpu
TPD = tt.TICKS_PER_DAY
TODAY = tt.DEFAULT_TODAY
YDAY = TODAY - TPD


def same_state(oracle, ot, table, mode):
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    assert (gc == c).all() and (go == o).all()
    assert len(table) == len(ot) and table.live_count() == ot.live_count()
    want = ot.get_yesterdays_due(mode, TODAY)
    got = table.get_yesterdays_due(mode, TODAY).rows()
    assert got.tolist() == want.tolist()


@pytest.mark.parametrize("mode", [tt.MODE_FAKE, tt.MODE_STORE])
def test_random_mix_of_point_operations(oracle, mode):
    rng = np.random.default_rng(17 + mode)
    n0 = 20_000
    cfg = oracle.default_cfg(99, n0, n_creators=40)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n0)
    table = load_from_oracle(ot)
    ids, _, _ = ot.export()
    known = [bytes(ids[i]) for i in range(n0)]
    creators = [ot.created_by(i) for i in range(40)] + [b"new.person@example.org", b"someone.else@example.org"]
    for step in range(300):
        op = rng.integers(0, 6)
        tid = known[int(rng.integers(0, len(known)))] if rng.random() < 0.9 else uuid.uuid4().bytes
        if op == 0:  # CreateNewTask
            new = uuid.uuid4().bytes
            who = creators[int(rng.integers(0, len(creators)))]
            due = int(rng.choice([YDAY, YDAY + 5, TODAY, YDAY | (1 << 62)]))
            created = (TODAY - int(rng.integers(1, 10**9))) | (1 << 62)
            row = ot.add(new, "t", who, "a", created, due)
            assert table.insert_task(new, created, due, who) == row
            known.append(new)
        elif op == 1:  # GetTaskById
            res = table.get_by_id(tid)
            want = ot.get_by_id(mode, tid)
            assert (res.rows().tolist() == [want]) if want >= 0 else res.count == 0
            if want >= 0:
                assert res.ids()[0].tobytes() == tid
        elif op == 2:  # MarkTaskCompleted
            assert table.mark_completed(tid) == bool(ot.mark_completed(mode, tid))
        elif op == 3:  # UpdateTask
            due = int(rng.choice([YDAY, TODAY + TPD, YDAY + 1]))
            want = ot.update_task(mode, tid, "renamed", "moved", due)
            got = table.update_task(tid, due)
            assert got == (want == 1)  # want == 2: the store implementation throws on a missing id
        elif op == 4:  # DeleteTask
            existed = ot.get_by_id(mode, tid) >= 0
            want = ot.delete_task(mode, tid)  # the store version answers true either way
            assert want == (existed if mode == tt.MODE_FAKE else 1)
            assert table.delete_task(tid) == existed
        else:  # GetTasksByCreator
            who = creators[int(rng.integers(0, len(creators)))]
            assert table.get_by_creator(who).rows().tolist() == ot.get_by_creator(mode, who).tolist()
        if step % 50 == 49:
            same_state(oracle, ot, table, mode)
    same_state(oracle, ot, table, mode)
    # the cron tick on the edited table
    threw, listed, n_over = ot.cron_job(mode)
    res, n_marked, gthrew = table.sweep(mode)
    assert gthrew == threw and (res.rows() == np.sort(listed)).all()
    assert n_marked == (n_over if mode == tt.MODE_STORE else 0)
    same_state(oracle, ot, table, mode)
    table.close()


def test_table_grows_from_empty(oracle):
    """CreateNewTask on an empty manager, past one tile of capacity (the columns are regrown)"""
    ot = oracle.OracleTable()
    table = tt.TaskTable(0)
    rng = np.random.default_rng(3)
    ids = []
    for i in range(8192 + 300):
        tid = uuid.UUID(int=(i + 1) << 64 | 0x4000_8000_0000_0000_0000 | i, version=None).bytes if i % 7 else uuid.uuid4().bytes
        who = b"grow%d@example.org" % (i % 13)
        due = YDAY if i % 5 == 0 else TODAY
        created = (TODAY - int(rng.integers(1, 10**9))) | (1 << 62)
        assert table.insert_task(tid, created, due, who) == ot.add(tid, "t", who, "a", created, due)
        ids.append(tid)
    assert len(table) == 8492 and table.live_count() == 8492
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        assert table.get_yesterdays_due(mode).rows().tolist() == ot.get_yesterdays_due(mode).tolist()
    assert table.get_by_creator(b"grow5@example.org").rows().tolist() == ot.get_by_creator(0, b"grow5@example.org").tolist()
    for k in (0, 1, 8191, 8192, 8491):
        assert table.get_by_id(ids[k]).rows().tolist() == [k]
    table.close()


def test_manager_mirror_full_interface(oracle):
    """GpuTasksManager: all eight ITasksManager methods, FAKE semantics (the registered default)"""
    import importlib

    pkg = importlib.import_module("aca-dotnet-workshop_b200")
    clock = [tt.DEFAULT_RUN_AT]
    mgr = pkg.GpuTasksManager([], mode=tt.MODE_FAKE, utc_now=lambda: clock[0] | (1 << 62))
    a = mgr.CreateNewTask("first", "ann@x.io", "bob@x.io", YDAY)
    clock[0] += 10
    b = mgr.CreateNewTask("second", "ann@x.io", "cy@x.io", TODAY)
    clock[0] += 10
    c = mgr.CreateNewTask("third", "dan@x.io", "cy@x.io", YDAY)
    assert [t.TaskName for t in mgr.GetTasksByCreator("ann@x.io")] == ["second", "first"]
    assert mgr.GetTaskById(b).TaskAssignedTo == "cy@x.io" and mgr.GetTaskById(uuid.uuid4()) is None
    assert [t.TaskName for t in mgr.GetYesterdaysDueTasks()] == ["first", "third"]
    assert mgr.MarkTaskCompleted(a) and mgr.GetTaskById(a).IsCompleted and not mgr.MarkTaskCompleted(uuid.uuid4())
    assert mgr.UpdateTask(c, "third!", "eve@x.io", TODAY) and not mgr.UpdateTask(uuid.uuid4(), "x", "y", TODAY)
    assert [t.TaskName for t in mgr.GetYesterdaysDueTasks()] == ["first"]
    assert mgr.GetTaskById(c).TaskName == "third!" and mgr.GetTaskById(c).TaskAssignedTo == "eve@x.io"
    assert mgr.DeleteTask(a) and not mgr.DeleteTask(a)
    assert mgr.GetYesterdaysDueTasks() == [] and [t.TaskName for t in mgr.GetTasksByCreator("ann@x.io")] == ["second"]
    mgr.close()


def test_snapshot_roundtrip(oracle, tmp_path):
    """save -> load into a fresh table: every query answers the same, tombstones and edits included"""
    n = 120_000
    cfg = oracle.default_cfg(55, n, n_creators=300)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    ids, _, _ = ot.export()
    for k in (5, 77, 4000):
        assert table.delete_task(bytes(ids[k])) and ot.delete_task(oracle.MODE_STORE, bytes(ids[k]))
    new = uuid.uuid4().bytes  # lands out of TaskId order: the reloaded table must rebuild its id index
    created = (TODAY - 12345) | (1 << 62)
    assert table.insert_task(new, created, YDAY, b"late@example.org") == ot.add(new, "t", b"late@example.org", "a", created, YDAY)
    table.sweep(tt.MODE_STORE)
    ot.cron_job(oracle.MODE_STORE)
    path = str(tmp_path / "tasks.snap")
    table.save_snapshot(path)
    again = tt.TaskTable(0).load_snapshot(path)
    assert len(again) == len(table) == len(ot) and again.live_count() == ot.live_count()
    assert (again.snapshot_flags()[0] == table.snapshot_flags()[0]).all()
    assert (again.snapshot_flags()[1] == table.snapshot_flags()[1]).all()
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        a, b = again.get_yesterdays_due(mode), table.get_yesterdays_due(mode)
        assert (a.rows() == b.rows()).all() and (a.ids() == b.ids()).all()
        assert a.rows().tolist() == ot.get_yesterdays_due(mode).tolist()
    for who in (ot.created_by(3), b"late@example.org", b"nobody"):
        assert again.get_by_creator(who).rows().tolist() == ot.get_by_creator(oracle.MODE_FAKE, who).tolist()
    assert again.get_by_id(new).rows().tolist() == [n] and again.get_by_id(bytes(ids[77])).count == 0
    assert again.get_by_id(bytes(ids[78])).rows().tolist() == [78]
    with pytest.raises(tt.TaskTableError):
        again.load_snapshot(path)  # not empty any more
    with pytest.raises(tt.TaskTableError):
        tt.TaskTable(0).load_snapshot(str(tmp_path / "missing.snap"))
    table.close()
    again.close()


def test_list_endpoints_follow_single_row_edits(oracle):
    """the device-made response bodies (GpuTasksManager.OverdueTasksGet / TasksGet = tt_result_json) after
    CreateNewTask, UpdateTask, MarkTaskCompleted and DeleteTask, against the oracle doing the same edits
    (FakeTasksManager.cs:32-100) and serialising the same lists"""
    import importlib

    pkg = importlib.import_module("aca-dotnet-workshop_b200")
    guids = [uuid.UUID(int=(k + 1) << 64 | 0x4000_8000_0000_0000_0000 | k) for k in range(8)]
    it = iter(guids)
    clock = [tt.DEFAULT_RUN_AT]
    seed = [pkg.TaskModel(next(it), f"seed {k}", "ann@x.io", (TODAY - 5000 + k) | (1 << 62), YDAY, "bob \"b\" <x>")
            for k in range(3)]
    mgr = pkg.GpuTasksManager(seed, mode=tt.MODE_FAKE, new_guid=lambda: next(it), utc_now=lambda: clock[0] | (1 << 62))
    ot = oracle.OracleTable()
    for t in seed:
        ot.add(t.TaskId.bytes, t.TaskName, t.TaskCreatedBy, t.TaskAssignedTo, t.TaskCreatedOn, t.TaskDueDate)

    def same():
        assert mgr.OverdueTasksGet() == ot.to_json(ot.get_yesterdays_due(oracle.MODE_FAKE), oracle.JSON_RESPONSE)
        for who in ("ann@x.io", "dan@x.io", "nobody"):
            assert mgr.TasksGet(who) == ot.to_json(ot.get_by_creator(oracle.MODE_FAKE, who.encode()), oracle.JSON_RESPONSE)

    same()
    clock[0] += 77
    new = mgr.CreateNewTask("caf\u00e9 & co", "dan@x.io", "eve+1@x.io", YDAY)          # a name the table has not seen
    ot.add(new.bytes, "caf\u00e9 & co", "dan@x.io", "eve+1@x.io", clock[0] | (1 << 62), YDAY)
    same()
    clock[0] += 5
    again = mgr.CreateNewTask("seed 1", "ann@x.io", "bob \"b\" <x>", TODAY)             # strings it already has
    ot.add(again.bytes, "seed 1", "ann@x.io", "bob \"b\" <x>", clock[0] | (1 << 62), TODAY)
    same()
    assert mgr.UpdateTask(seed[0].TaskId, "renamed \\ task", "", YDAY)
    assert ot.update_task(oracle.MODE_FAKE, seed[0].TaskId.bytes, "renamed \\ task", "", YDAY)
    same()
    assert mgr.MarkTaskCompleted(seed[1].TaskId) and ot.mark_completed(oracle.MODE_FAKE, seed[1].TaskId.bytes)
    assert mgr.DeleteTask(seed[2].TaskId) and ot.delete_task(oracle.MODE_FAKE, seed[2].TaskId.bytes)
    same()
    assert b'"taskName":"renamed \\\\ task"' in mgr.TasksGet("ann@x.io") and b'"isCompleted":true' in mgr.TasksGet("ann@x.io")
    mgr.close()


def test_snapshot_carries_the_payload(oracle, tmp_path):
    """TaskName / TaskAssignedTo survive save -> load: the reloaded table prints the same JSON, single-row
    string edits included, and can take further edits (its dictionaries are live again)"""
    n = 50_000
    cfg = oracle.default_cfg(56, n, n_creators=20, n_names=500)
    table = tt.TaskTable(0).generate(cfg, 0, n).generate_payload(cfg, 0)
    who = b"user0000003@mail3.example"
    first = table.get_by_creator(who)
    ids = first.ids()
    assert first.count > 1000 and table.set_task_strings(ids[0], b"edited \"name\"", None)
    want = table.get_by_creator(who).to_json(tt.JSON_DOCUMENT)
    path = str(tmp_path / "payload.snap")
    table.save_snapshot(path)
    again = tt.TaskTable(0).load_snapshot(path)
    assert again.get_by_creator(who).to_json(tt.JSON_DOCUMENT) == want and b"edited \\u0022name\\u0022" in want
    assert again.set_task_strings(ids[1], b"after reload", b"new assignee") and table.set_task_strings(ids[1], b"after reload", b"new assignee")
    assert again.get_by_creator(who).to_json(tt.JSON_RESPONSE) == table.get_by_creator(who).to_json(tt.JSON_RESPONSE)
    # a table without payload still saves and loads (and prints empty strings)
    bare = tt.TaskTable(0).generate(cfg, 0, 1000)
    bare.save_snapshot(path)
    bare2 = tt.TaskTable(0).load_snapshot(path)
    assert bare2.get_by_creator(who).to_json() == bare.get_by_creator(who).to_json()
    for t in (table, again, bare, bare2):
        t.close()


def test_snapshot_is_replaced_atomically_and_checked_on_load(oracle, tmp_path):
    """A save goes to <path>.tmp and is renamed over <path>; a load refuses a file that is cut short, whose bytes
    changed, or whose references point outside their dictionaries, and leaves the table empty and usable; files
    written before the trailer existed (magic TTSNAP01) still load."""
    import os

    n = 20_000
    cfg = oracle.default_cfg(91, n, n_creators=12, n_names=40)
    table = tt.TaskTable(0).generate(cfg, 0, n).generate_payload(cfg, 0)
    path = str(tmp_path / "t.snap")
    table.save_snapshot(path)
    assert not os.path.exists(path + ".tmp")
    good = open(path, "rb").read()
    assert good[:8] == b"TTSNAP02" and good[-24:-16] == b"TTSNEND2"
    want = table.get_yesterdays_due(tt.MODE_STORE).ids()
    # a save that cannot finish (the directory of the temporary file is gone) leaves the old file alone
    with pytest.raises(tt.TaskTableError):
        table.save_snapshot(str(tmp_path / "no" / "such" / "dir.snap"))
    assert open(path, "rb").read() == good

    def refused(blob, code):
        bad = str(tmp_path / "bad.snap")
        open(bad, "wb").write(blob)
        t2 = tt.TaskTable(0)
        with pytest.raises(tt.TaskTableError) as e:
            t2.load_snapshot(bad)
        assert e.value.code == code and len(t2) == 0
        t2.load_snapshot(path)  # the table is still empty and takes the good file
        assert (t2.get_yesterdays_due(tt.MODE_STORE).ids() == want).all()
        t2.close()

    refused(good[: len(good) // 2], tt.TT_ERR_STATE)                     # cut short
    refused(good[:-24], tt.TT_ERR_STATE)                                  # trailer missing
    flipped = bytearray(good)
    flipped[len(good) // 3] ^= 0x40
    refused(bytes(flipped), tt.TT_ERR_STATE)                              # one bit changed: checksum
    # an old-format file (no trailer, so no checksum) with a creator_ref beyond the dictionary: caught by the check
    # of the references on the device
    head = 56
    np_rows = int.from_bytes(good[16:24], "little")
    ref_off = head + np_rows * 4 + 2 * (np_rows // 8) + np_rows * 8 + np_rows * 16 + np_rows * 8 + np_rows * 8
    old = bytearray(b"TTSNAP01" + good[8:-24])
    t3 = tt.TaskTable(0).load_snapshot(_write(tmp_path / "old.snap", bytes(old)))
    assert (t3.get_yesterdays_due(tt.MODE_STORE).ids() == want).all()
    t3.close()
    old[ref_off + 4 * 7: ref_off + 4 * 7 + 4] = (1000).to_bytes(4, "little")
    refused(bytes(old), tt.TT_ERR_ARG)
    table.close()


def _write(path, blob):
    open(path, "wb").write(blob)
    return str(path)


def test_mark_overdue_after_creates_with_fresh_guids(oracle):
    """The reference's everyday flow: CreateNewTask with Guid.NewGuid() ids (which break the table's TaskId
    order), then the cron job — GET the list, POST it back to MarkOverdueTasks (TasksStoreManager.cs:141-149).
    tt_mark_overdue has to find the new rows too (it rebuilds the TaskId index), pass over deleted rows, and
    report ids it does not hold."""
    rng = np.random.default_rng(5)
    n0 = 30_000
    cfg = oracle.default_cfg(1234, n0, n_creators=30)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n0)
    table = load_from_oracle(ot)
    fresh = []
    for i in range(400):
        tid = uuid.UUID(int=int.from_bytes(rng.bytes(16), "big"), version=4).bytes
        created = (TODAY - int(rng.integers(1, 10**9))) | (1 << 62)
        due = YDAY if i % 2 == 0 else TODAY
        who = b"creator%d@example.org" % (i % 7)
        assert table.insert_task(tid, created, due, who) == ot.add(tid, "t", who, "a", created, due)
        fresh.append(tid)
    # one of the new tasks is deleted again: its row keeps the id as a tombstone
    gone = fresh[10]
    assert table.delete_task(gone) and ot.delete_task(tt.MODE_STORE, gone)
    listed = ot.get_yesterdays_due(tt.MODE_STORE)
    res = table.get_yesterdays_due(tt.MODE_STORE)
    assert res.rows().tolist() == listed.tolist() and (listed >= n0).sum() == 199
    ids = res.ids()
    unknown = uuid.uuid4().bytes
    post = np.concatenate([ids, np.frombuffer(unknown + gone, dtype=np.uint8).reshape(2, 16)])
    marked, missing = table.mark_overdue(tt.MODE_STORE, post)
    threw, want_missing = ot.mark_overdue(tt.MODE_STORE, post)
    assert not threw and (marked, missing) == (len(listed), 2) and want_missing == 2
    # (the oracle's store upserted two documents for the unknown ids; they are not rows of the list)
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    assert (gc == c).all() and (go == o).all()
    # a second insert drops the index again; marking still works
    extra = uuid.uuid4().bytes
    table.insert_task(extra, TODAY | (1 << 62), YDAY, b"late@example.org")
    assert table.mark_overdue(tt.MODE_STORE, np.frombuffer(extra, dtype=np.uint8)) == (1, 0)
    assert table.get_yesterdays_due(tt.MODE_STORE).count == 0
    table.close()


def test_manager_mirror_mark_overdue_upserts_unknown_tasks(oracle):
    """MarkOverdueTasks is an upsert of the whole record (SaveStateAsync, TasksStoreManager.cs:147): a listed task
    the store does not hold is created from the caller's copy, with IsOverDue = true."""
    import importlib

    pkg = importlib.import_module("aca-dotnet-workshop_b200")
    mgr = pkg.GpuTasksManager([], mode=tt.MODE_STORE)
    a = mgr.CreateNewTask("kept", "ann@x.io", "bob@x.io", YDAY)
    listed = mgr.GetYesterdaysDueTasks()
    assert [t.TaskId for t in listed] == [a]
    ghost = pkg.TaskModel(uuid.uuid4(), "ghost", "zed@x.io", (TODAY - 77) | (1 << 62), YDAY, "amy@x.io", True, False)
    mgr.MarkOverdueTasks(listed + [ghost])
    assert mgr.last_mark_missing == 1 and ghost.IsOverDue
    back = mgr.GetTaskById(ghost.TaskId)
    assert back == ghost and mgr.GetTaskById(a).IsOverDue
    assert [t.TaskName for t in mgr.GetTasksByCreator("zed@x.io")] == ["ghost"]
    assert mgr.GetYesterdaysDueTasks() == []
    mgr.table.close()
