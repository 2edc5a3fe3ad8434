"""GPU parity of GetTasksByCreator: hash scan + exact string confirm against the oracle."""
import numpy as np
import pytest

from helpers import load_from_oracle, oracle_columns, tt

pytestmark = pytest.mark.gpu


def mix64(z):
    M = (1 << 64) - 1
    z = (z + 0x9E3779B97F4A7C15) & M
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M
    return z ^ (z >> 31)


def hash64(b):
    h = mix64(0x7461736B73 ^ len(b))
    for i in range(0, len(b), 8):
        h = mix64(h ^ int.from_bytes(b[i:i + 8].ljust(8, b"\0"), "little"))
    return h


def colliding_pair():
    """two distinct 16-byte ASCII strings with the same tt_hash64 (the chain is invertible)"""
    a = b"alice@x.example!"
    h0 = mix64(0x7461736B73 ^ 16)
    ha = mix64(h0 ^ int.from_bytes(a[:8], "little"))
    a2 = int.from_bytes(a[8:], "little")
    for k in range(1, 1 << 20):
        b1 = b"bob%05d" % k
        hb = mix64(h0 ^ int.from_bytes(b1, "little"))
        b2 = (a2 ^ ha ^ hb).to_bytes(8, "little")
        if all(0x20 <= c < 0x7F for c in b2):
            b = b1 + b2
            assert hash64(a) == hash64(b) and a != b
            return a, b
    raise AssertionError("no printable collision found")


def test_hash_function_matches_library(oracle):
    # the library hashes the dictionary with tt_hash64; a query for an existing creator must hit
    assert hash64(b"") == mix64(0x7461736B73)


def test_creator_filter_matches_oracle(oracle):
    n = 400_000
    cfg = oracle.default_cfg(0x5EED0003, n, n_creators=4000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    ids_all, created, _ = ot.export()
    rng = np.random.default_rng(4)
    queries = [ot.created_by(int(r)) for r in rng.integers(0, n, size=200)]
    queries += [b"nobody@nowhere.example", b"", b"user0000001@mail1.examplE", b"User0000001@mail1.example",
                b"user0000001@mail1.example ", b"user0000001@mail1.exampl"]
    for q in queries:
        want = ot.get_by_creator(oracle.MODE_FAKE, q)
        assert (want == ot.get_by_creator(oracle.MODE_STORE, q)).all()
        res = table.get_by_creator(q)
        rows = res.rows()
        assert rows.tolist() == want.tolist(), q
        assert (res.ids() == ids_all[rows]).all()
        res.free()
    table.close()


def test_creator_batch_matches_oracle(oracle):
    """64+ creators in one table pass: duplicates, absent names, near misses, and a big batch"""
    n = 400_000
    cfg = oracle.default_cfg(0x5EED0003, n, n_creators=4000, ppm_created_dup=100_000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    ids_all, _, _ = ot.export()
    rng = np.random.default_rng(5)
    for nq in (3, 64, 700):
        queries = [ot.created_by(int(r)) for r in rng.integers(0, n, size=nq - 2)]
        queries += [b"nobody@nowhere.example", queries[0]]  # an absent creator and a repeated one
        results = table.get_by_creators(queries)
        st = table.last_op_stats()
        assert len(results) == nq and st.queries == nq and st.scan_ms > 0
        total = 0
        for q, res in zip(queries, results):
            want = ot.get_by_creator(oracle.MODE_FAKE, q)
            rows = res.rows()
            assert rows.tolist() == want.tolist(), q
            assert (res.ids() == ids_all[rows]).all()
            total += rows.size
        assert st.matches == sum(ot.get_by_creator(oracle.MODE_FAKE, q).size for q in set(queries))
        for res in results:
            res.free()
    # a batch whose total result is too large for the single-CTA ordering path
    few = tt.TaskTable(0).generate(oracle.default_cfg(13, 200_000, n_creators=5, ppm_created_dup=200_000), 0, 200_000)
    ot2 = oracle.OracleTable()
    ot2.add_synth(oracle.default_cfg(13, 200_000, n_creators=5, ppm_created_dup=200_000), 0, 200_000)
    qs = [ot2.created_by(k) for k in range(20)]
    for q, res in zip(qs, few.get_by_creators(qs)):
        assert res.rows().tolist() == ot2.get_by_creator(oracle.MODE_FAKE, q).tolist()
    table.close()
    few.close()


def test_colliding_queries_in_one_batch(oracle):
    a, b = colliding_pair()
    TPD = tt.TICKS_PER_DAY
    ot = oracle.OracleTable()
    who = [a, b, a, b"carol@x.example", b, a]
    for i, w in enumerate(who):
        ot.add(bytes([i + 1] * 16), "n", w, "z", (tt.DEFAULT_TODAY - TPD + i) | (1 << 62), tt.DEFAULT_TODAY - TPD)
    table = load_from_oracle(ot)
    ra, rb, rc = table.get_by_creators([a, b, b"carol@x.example"])
    assert ra.rows().tolist() == [5, 2, 0] and rb.rows().tolist() == [4, 1] and rc.rows().tolist() == [3]
    table.close()


def test_creator_ties_keep_source_order(oracle):
    n = 60_000
    cfg = oracle.default_cfg(12, n, n_creators=7, ppm_created_dup=300_000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    for k in range(7):
        q = ot.created_by(k)
        want = ot.get_by_creator(oracle.MODE_FAKE, q)
        got = table.get_by_creator(q).rows()
        assert got.size > 1000 and got.tolist() == want.tolist()
    table.close()


def test_forced_hash_collision_is_rejected_by_the_string_compare(oracle):
    a, b = colliding_pair()
    TPD = tt.TICKS_PER_DAY
    ot = oracle.OracleTable()
    who = [a, b, a, b"carol@x.example", b, a]
    for i, w in enumerate(who):
        ot.add(bytes([i + 1] * 16), "n", w, "z", (tt.DEFAULT_TODAY - TPD + i) | (1 << 62), tt.DEFAULT_TODAY - TPD)
    table = load_from_oracle(ot)
    assert table.get_by_creator(a).rows().tolist() == ot.get_by_creator(oracle.MODE_FAKE, a).tolist() == [5, 2, 0]
    assert table.get_by_creator(b).rows().tolist() == ot.get_by_creator(oracle.MODE_FAKE, b).tolist() == [4, 1]
    table.close()


def test_seed_rows_config1(oracle):
    """BASELINE.json config 1: the ten FakeTasksManager seed rows"""
    now = tt.DEFAULT_TODAY + 9 * 3600 * 10_000_000 + 1234567
    ot = oracle.OracleTable()
    ot.seed_fake(now)
    table = load_from_oracle(ot)
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        res, n_marked, threw = table.sweep(mode, tt.DEFAULT_TODAY, now)
        assert (res.count, n_marked, threw) == (0, 0, False)
    assert table.get_by_creator("tjoudeh@bitoftech.net").rows().tolist() == list(range(9, -1, -1))
    assert table.get_by_creator("Tjoudeh@bitoftech.net").count == 0
    table.close()


def test_manager_mirror_roundtrip(oracle):
    """GpuTasksManager (the ITasksManager mirror): unfused cron tick == fused tick == oracle"""
    import importlib
    import uuid

    pkg = importlib.import_module("aca-dotnet-workshop_b200")
    n = 20_000
    cfg = oracle.default_cfg(21, n, n_creators=50)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    cols = oracle_columns(ot)
    tasks = [pkg.TaskModel(uuid.UUID(bytes=cols["task_id"][i].tobytes()), f"Task number: {i}",
                           cols["creators"][cols["creator_ref"][i]].decode(), int(cols["created_raw"][i]),
                           int(cols["due_raw"][i]), f"assignee{i % 50}@mail.com", bool(cols["completed"][i]),
                           bool(cols["overdue"][i])) for i in range(n)]
    mgr = pkg.GpuTasksManager(tasks, mode=tt.MODE_STORE)
    want = ot.get_yesterdays_due(oracle.MODE_STORE)
    got = mgr.GetYesterdaysDueTasks()
    assert [t.TaskName for t in got] == [f"Task number: {r}" for r in want.tolist()]
    assert all(not t.IsCompleted and not t.IsOverDue for t in got)
    n_over = mgr.CheckOverDueTasksJob()
    _, listed, want_over = ot.cron_job(oracle.MODE_STORE)
    assert n_over == want_over == listed.size
    assert (mgr.table.flags_bytes()[1] == ot.flags()[1]).all()
    by = mgr.GetTasksByCreator(tasks[0].TaskCreatedBy)
    assert [t.TaskName for t in by] == [f"Task number: {r}" for r in ot.get_by_creator(0, tasks[0].TaskCreatedBy).tolist()]
    fake = pkg.GpuTasksManager(tasks[:5000], mode=tt.MODE_FAKE)
    with pytest.raises(tt.NotImplementedInReference):
        fake.CheckOverDueTasksJob()
    mgr.close()
    fake.close()
