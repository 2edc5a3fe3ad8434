"""The task table sharded by TaskId range behind the C ABI: tt_shards_* (one process drives
G shards) and tt_exchange_* (one exchange object per shard; here also across two processes through the CUDA
IPC handle).  Everything runs on ONE GPU — several shard tables and the root's receive buffer all live on
device 0 — so the driver's single-GPU test run exercises the whole path: global row numbers, TaskIds, the
cross-shard OrderBy(TaskCreatedOn) with ties, per-shard flag planes, the sharded creator filter and the
single-row methods, each against the CPU oracle's answer over the UNSHARDED table."""
import os
import subprocess
import sys
import uuid

import numpy as np
import pytest

from helpers import oracle_columns, tt

# how long the root's device-side wait lasts before it reports a missing shard (read once by the library,
# before the first exchange is made; pytest imports every test module before it runs any test)
os.environ.setdefault("TT_SHARD_TIMEOUT_MS", "1500")

pytestmark = pytest.mark.gpu
TPD = tt.TICKS_PER_DAY
TODAY = tt.DEFAULT_TODAY
YDAY = TODAY - TPD


def synth_pair(oracle, n, shards, seed=0x5A4D, **kw):
    kw.setdefault("n_creators", 300)
    cfg = oracle.default_cfg(seed, n, **kw)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    st = tt.ShardedTable([0] * shards).generate(cfg, n)
    return cfg, ot, st


def planes(st, shards):
    comp, over = [], []
    for g in range(shards):
        table, first = st.shard(g)
        c, o = table.flags_bytes()
        assert first == sum(len(x) for x in comp)
        comp.append(c)
        over.append(o)
    return np.concatenate(comp), np.concatenate(over)


@pytest.mark.parametrize("shards", [1, 2, 4])
def test_sharded_manager_matches_the_unsharded_oracle(oracle, shards):
    n = 1_000_003  # ragged: shard sizes differ, the last tile of every shard is partial
    cfg, ot, st = synth_pair(oracle, n, shards, ppm_created_dup=200_000)  # 20 % duplicated TaskCreatedOn: ties
    assert len(st) == n
    ids_all, created_all, _ = ot.export()
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        want = ot.get_yesterdays_due(mode)
        res = st.get_yesterdays_due(mode)
        assert res.count == len(want) > 10_000
        rows = res.rows()
        assert rows.dtype == np.uint64 and (rows == want).all(), "global rows in the reference's order"
        assert (res.ids() == ids_all[want]).all()
        if mode == tt.MODE_STORE:  # the order really is by TaskCreatedOn, and there are ties to get right
            ticks = created_all[rows.astype(np.int64)] & ((1 << 62) - 1)
            assert (np.diff(ticks.astype(np.int64)) >= 0).all() and (np.diff(ticks.astype(np.int64)) == 0).sum() > 20
        counts = res.shard_counts()[:shards]
        assert sum(counts) == res.count and all(c > 0 for c in counts)
        res.free()
    # the cron tick: FAKE lists and throws, STORE lists, marks, and a second tick finds nothing
    threw, listed, _ = ot.cron_job(tt.MODE_FAKE)
    res, n_marked, gthrew = st.sweep(tt.MODE_FAKE)
    assert gthrew == threw and n_marked == 0 and (res.rows() == np.sort(listed)).all()
    threw, listed, n_over = ot.cron_job(tt.MODE_STORE)
    res, n_marked, gthrew = st.sweep(tt.MODE_STORE)
    assert not gthrew and not threw and n_marked == n_over == len(listed)
    assert (res.rows() == np.sort(listed)).all() and (res.ids() == ids_all[np.sort(listed)]).all()
    c, o = ot.flags()
    gc, go = planes(st, shards)
    assert (gc == c).all() and (go == o).all(), "per-shard flag planes, end to end"
    res2, n2, _ = st.sweep(tt.MODE_STORE)
    assert res2.count == 0 and n2 == 0
    st.close()


def test_sharded_creator_filter_and_backlog(oracle):
    n, shards = 600_000, 3
    cfg, ot, st = synth_pair(oracle, n, shards, seed=77, n_creators=50, ppm_created_dup=300_000)
    ids_all, _, _ = ot.export()
    for who in (ot.created_by(5), ot.created_by(123_456), b"nobody@example.org", ot.created_by(5).upper()):
        want = ot.get_by_creator(tt.MODE_FAKE, who)
        assert (want == ot.get_by_creator(tt.MODE_STORE, who)).all()
        res = st.get_by_creator(who)
        assert res.count == len(want)
        assert (res.rows() == want).all() and (res.ids() == ids_all[want]).all()
    assert len(ot.get_by_creator(tt.MODE_FAKE, ot.created_by(5))) > 5_000  # far beyond one CTA's bitonic list
    # the generalised predicate goes through the same exchange (a much larger list: the slots are regrown)
    want, _, _ = ot.backlog(tt.MODE_STORE)
    res = st.get_due_on_or_before(tt.MODE_STORE)
    assert res.count == len(want) > 50_000 and (res.rows() == want).all()
    st.close()


def test_sharded_single_row_methods_and_mark_overdue(oracle):
    """Tasks arrive through CreateNewTask with random Guids: each goes to the shard that owns its id range."""
    rng = np.random.default_rng(11)
    shards = 4
    st = tt.ShardedTable([0] * shards)
    ot = oracle.OracleTable()
    tids = []
    for i in range(3000):
        tid = uuid.UUID(int=int.from_bytes(rng.bytes(16), "big"), version=4).bytes
        created = (TODAY - int(rng.integers(1, 10**12))) | (1 << 62)  # distinct creation times
        due = YDAY if i % 3 == 0 else TODAY
        who = b"user%d@example.org" % (i % 11)
        g, _ = st.insert_task(tid, created, due, who)
        assert g == tid[0] * shards // 256  # equal-width ranges of the id space
        ot.add(tid, "t", who, "a", created, due)
        tids.append(tid)
    assert len(st) == 3000
    ids_all, created_all, _ = ot.export()
    for k in (0, 7, 2999):
        r = st.get_by_id(tids[k])
        assert r.count == 1 and r.ids()[0].tobytes() == tids[k]
    assert st.get_by_id(uuid.uuid4().bytes).count == 0
    assert st.mark_completed(tids[3]) and ot.mark_completed(tt.MODE_STORE, tids[3])
    assert st.update_task(tids[4], TODAY) and ot.update_task(tt.MODE_STORE, tids[4], "n", "a", TODAY) == 1
    assert st.delete_task(tids[6]) and ot.delete_task(tt.MODE_STORE, tids[6])
    assert not st.mark_completed(uuid.uuid4().bytes)
    # STORE list: ordered by TaskCreatedOn over all shards; row numbers differ from the oracle's (its list is
    # in insertion order, the shards are in TaskId-range order), so compare the TaskIds
    want = ot.get_yesterdays_due(tt.MODE_STORE)
    res = st.get_yesterdays_due(tt.MODE_STORE)
    assert res.count == len(want) and (res.ids() == ids_all[want]).all()
    for who in (b"user3@example.org", b"user10@example.org"):
        want = ot.get_by_creator(tt.MODE_STORE, who)
        res = st.get_by_creator(who)
        assert (res.ids() == ids_all[want]).all()
    # MarkOverdueTasks(list): ids are routed to their shards; unknown and deleted ones are reported missing
    listed = ot.get_yesterdays_due(tt.MODE_STORE)
    post = np.concatenate([ids_all[listed], np.frombuffer(uuid.uuid4().bytes + tids[6], dtype=np.uint8).reshape(2, 16)])
    assert st.mark_overdue(tt.MODE_STORE, post) == (len(listed), 2)
    assert st.get_yesterdays_due(tt.MODE_STORE).count == 0
    with pytest.raises(tt.NotImplementedInReference):
        st.mark_overdue(tt.MODE_FAKE, post)
    st.close()


def test_load_soa_partitions_by_equal_width_id_ranges(oracle):
    n, shards = 50_000, 3
    cfg = oracle.default_cfg(9, n, n_creators=20, ppm_created_dup=0)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    cols = oracle_columns(ot)
    rng = np.random.default_rng(2)
    cols["task_id"] = np.frombuffer(rng.bytes(16 * n), dtype=np.uint8).reshape(n, 16).copy()  # random ids, any order
    st = tt.ShardedTable([0] * shards).load_soa(**cols)
    sizes = [len(st.shard(g)[0]) for g in range(shards)]
    bound = [(g << 64) // shards for g in range(shards)]  # lowest top-64-bit value shard g owns
    owner = [max(g for g in range(shards) if bound[g] <= int.from_bytes(bytes(i[:8]), "big")) for i in cols["task_id"]]
    want_sizes = [owner.count(g) for g in range(shards)]
    assert sizes == want_sizes and sum(sizes) == n
    want = ot.get_yesterdays_due(tt.MODE_STORE)  # created order; creation times are distinct here
    res = st.get_yesterdays_due(tt.MODE_STORE)
    assert (res.ids() == cols["task_id"][want]).all()
    st.close()


def test_exchange_objects_in_one_process(oracle):
    """The lower layer, as bench.py drives it: explicit exchange objects, asynchronous ticks, a collect that
    waits on the epoch words, views that expire, a slot that is too small, a shard that never ticks."""
    n, world = 400_000, 2
    cfg = oracle.default_cfg(31, n, n_creators=10)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    bounds = [0, n // 2, n]
    tables = [tt.TaskTable(0).generate(cfg, bounds[g], bounds[g + 1] - bounds[g]) for g in range(world)]
    tiles = max((len(t) + 8191) // 8192 for t in tables)
    root, handle = tt.Exchange.create_root(0, world, 16384, tiles, tt.EXCHANGE_IDS)
    assert len(handle) == 64 and any(handle) and root.slot_bytes % 256 == 0
    ex = [root, root.attach_local(0, 1)]
    want = np.sort(ot.get_yesterdays_due(tt.MODE_STORE))
    ids_all, _, _ = ot.export()
    flags = tt.SWEEP_LIST_ONLY | tt.SWEEP_ASYNC | tt.SWEEP_NO_TIMING
    for tick in range(3):  # the slots are reused tick after tick
        for g in (1, 0):
            ex[g].tick(tables[g], tt.MODE_STORE, flags)
        res = root.collect(tt.ORDER_ROWS, bounds[:world])
        for g in range(world):
            st, _ = tables[g].sweep_wait()
            assert st.n_stored == res.shard_counts()[g]
        assert root.epoch == ex[1].epoch == tick + 1
        assert (res.rows() == want).all() and (res.ids() == ids_all[want]).all()
        root.release()
        res.free()
    # a view that was not read before the next collect says so instead of returning the wrong list
    for g in range(world):
        ex[g].tick(tables[g], tt.MODE_STORE, tt.SWEEP_LIST_ONLY)
    stale = root.collect(tt.ORDER_ROWS | tt.COLLECT_RELEASE, bounds[:world])
    for g in range(world):
        ex[g].tick(tables[g], tt.MODE_STORE, tt.SWEEP_LIST_ONLY)
    fresh = root.collect(tt.ORDER_ROWS | tt.COLLECT_RELEASE, bounds[:world])
    with pytest.raises(tt.TaskTableError) as e:
        stale.rows()
    assert e.value.code == tt.TT_ERR_STATE
    assert (fresh.rows() == want).all()
    # ordering by TaskCreatedOn needs an exchange that carries the keys
    with pytest.raises(tt.TaskTableError):
        root.collect(tt.ORDER_CREATED_ASC, bounds[:world])
    # a shard that does not tick: the root's wait gives up (TT_SHARD_TIMEOUT_MS) instead of hanging the GPU
    ex[0].tick(tables[0], tt.MODE_STORE, tt.SWEEP_LIST_ONLY)
    with pytest.raises(tt.TaskTableError) as e:
        root.collect(tt.ORDER_ROWS, bounds[:world])
    assert e.value.code == tt.TT_ERR_STATE and "shard 1" in str(e.value)
    for x in (ex[1], root):
        x.close()
    # slots too small for the hits: TT_ERR_CAPACITY from the tick, a row-ordered subset in the slot
    small, _ = tt.Exchange.create_root(0, 1, 1024, tiles, 0)
    st, rc = small.tick(tables[0], tt.MODE_STORE, tt.SWEEP_LIST_ONLY, allow_capacity=True)
    assert rc == tt.TT_ERR_CAPACITY and 0 < st.n_stored <= 1024 < st.n_listed
    part = small.collect(tt.ORDER_ROWS, [0])
    rows = part.rows()
    assert len(rows) == st.n_stored and np.isin(rows, want).all() and (np.diff(rows.astype(np.int64)) > 0).all()
    with pytest.raises(tt.TaskTableError):
        part.ids()  # this exchange carries no TaskIds
    small.close()
    for t in tables:
        t.close()


CHILD = r"""
import importlib, sys
sys.path.insert(0, sys.argv[1])
pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable
seed, n, first, rows, cap, tiles, flags, dup = (int(x) for x in sys.argv[2:10])
handle = bytes.fromhex(sys.stdin.readline().strip())
cfg = tt.default_cfg(seed, n, n_creators=10, ppm_created_dup=dup)
table = tt.TaskTable(0).generate(cfg, first, rows)
ex = tt.Exchange.attach(0, 1, 2, cap, tiles, flags, handle)
for line in sys.stdin:                      # one tick per line from the parent
    if line.strip() == "quit":
        break
    st, _ = ex.tick(table, int(line), tt.SWEEP_MARK)
    print(st.n_stored, st.n_marked, flush=True)
ex.close()
table.close()
"""


def test_exchange_across_two_processes(oracle, tmp_path):
    """Shard 1 lives in ANOTHER process and stores into the root's buffer through the CUDA IPC mapping, exactly as
    one rank per GPU does under torchrun; both shards are on device 0 here.  Marking ticks, TaskIds and the
    cross-shard order included; the second tick must wait for the root's release of the first."""
    n, seed = 500_000, 4711
    cfg = oracle.default_cfg(seed, n, n_creators=10, ppm_created_dup=100_000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    half = n // 2
    table = tt.TaskTable(0).generate(cfg, 0, half)
    cap, tiles, flags = 32768, (n - half + 8191) // 8192, tt.EXCHANGE_IDS | tt.EXCHANGE_CREATED
    root, handle = tt.Exchange.create_root(0, 2, cap, tiles, flags)
    script = tmp_path / "child.py"
    script.write_text(CHILD)
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    child = subprocess.Popen([sys.executable, str(script), repo, str(seed), str(n), str(half), str(n - half), str(cap), str(tiles),
                              str(flags), "100000"], stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True)
    try:
        child.stdin.write(handle.hex() + "\n")
        ids_all, _, _ = ot.export()
        want = ot.get_yesterdays_due(tt.MODE_STORE)  # TaskCreatedOn order, ties in row order
        threw, listed, n_over = ot.cron_job(tt.MODE_STORE)
        child.stdin.write(f"{tt.MODE_STORE}\n")
        child.stdin.flush()
        st0, _ = root.tick(table, tt.MODE_STORE, tt.SWEEP_MARK)
        theirs = [int(x) for x in child.stdout.readline().split()]
        res = root.collect(tt.ORDER_CREATED_ASC, [0, half])
        assert res.count == len(want) == st0.n_stored + theirs[0] and st0.n_marked + theirs[1] == n_over
        assert (res.rows() == want).all() and (res.ids() == ids_all[want]).all()
        res.free()
        root.release()
        # second tick: everything is marked already, both slots report zero hits
        child.stdin.write(f"{tt.MODE_STORE}\n")
        child.stdin.flush()
        root.tick(table, tt.MODE_STORE, tt.SWEEP_MARK)
        assert [int(x) for x in child.stdout.readline().split()] == [0, 0]
        res = root.collect(tt.ORDER_ROWS | tt.COLLECT_RELEASE, [0, half])
        assert res.count == 0
        # third tick with FAKE semantics (lists, never marks): row order across the two processes
        child.stdin.write(f"{tt.MODE_FAKE}\n")
        child.stdin.flush()
        root.tick(table, tt.MODE_FAKE, tt.SWEEP_MARK)
        child.stdout.readline()
        res = root.collect(tt.ORDER_ROWS, [0, half])
        fake = ot.get_yesterdays_due(tt.MODE_FAKE)
        assert (res.rows() == fake).all() and (res.ids() == ids_all[fake]).all()
        root.release()
        child.stdin.write("quit\n")
        child.stdin.flush()
        assert child.wait(timeout=60) == 0
    finally:
        if child.poll() is None:
            child.kill()
    root.close()
    table.close()


def test_sharded_manager_with_empty_shards(oracle):
    """Two tasks in a four-shard manager: two shards hold nothing, yet every list call still gets an answer from
    each of them (an empty shard publishes "no hits" for the tick)."""
    st = tt.ShardedTable([0] * 4)
    assert st.get_yesterdays_due(tt.MODE_STORE).count == 0 and st.sweep(tt.MODE_STORE)[0].count == 0
    a = bytes([0x10]) + bytes(15)   # shard 0
    b = bytes([0xF0]) + bytes(15)   # shard 3
    assert st.insert_task(b, (TODAY - 5) | (1 << 62), YDAY, b"x@y")[0] == 3
    assert st.insert_task(a, (TODAY - 9) | (1 << 62), YDAY, b"x@y")[0] == 0
    res = st.get_yesterdays_due(tt.MODE_STORE)   # created order: a (older) first, although b was inserted first
    assert res.count == 2 and [bytes(i) for i in res.ids()] == [a, b] and res.rows().tolist() == [0, 1]
    assert res.shard_counts()[:4] == [1, 0, 0, 1]
    assert [bytes(i) for i in st.get_by_creator(b"x@y").ids()] == [b, a]   # descending
    res, n_marked, _ = st.sweep(tt.MODE_STORE)
    assert res.count == 2 and n_marked == 2 and st.get_yesterdays_due(tt.MODE_STORE).count == 0
    st.close()


@pytest.mark.parametrize("ppm", [300, 120_000, 1_000_000])
def test_exchange_at_every_density(oracle, ppm):
    """From a hit every few thousand rows (most tiles empty) to every row a hit (pieces beyond the staging area take the
    direct path): three shards of ragged sizes, two ticks into the same slots."""
    n, world = 530_000, 3
    cfg = oracle.default_cfg(77 + ppm, n, n_creators=10, ppm_due_yesterday=ppm, ppm_adversarial=0, ppm_completed=0, ppm_overdue_past=0)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    want = np.sort(ot.get_yesterdays_due(tt.MODE_STORE))
    assert len(want) > 100
    bounds = [0, 170_001, 390_000, n]
    tables = [tt.TaskTable(0).generate(cfg, bounds[g], bounds[g + 1] - bounds[g]) for g in range(world)]
    tiles = max((len(t) + 8191) // 8192 for t in tables)
    cap = (max(len(t) for t in tables) + 7) // 8 * 8
    ids_all, _, _ = ot.export()
    root, _ = tt.Exchange.create_root(0, world, cap, tiles, tt.EXCHANGE_IDS)
    ex = [root] + [root.attach_local(0, g) for g in range(1, world)]
    for tick in range(2):
        for g in range(world):
            ex[g].tick(tables[g], tt.MODE_STORE, tt.SWEEP_LIST_ONLY)
        res = root.collect(tt.ORDER_ROWS, bounds[:world])
        assert res.count == len(want) and (res.rows() == want).all() and (res.ids() == ids_all[want]).all(), tick
        root.release()
    for x in ex[1:] + [root]:
        x.close()
    for t in tables:
        t.close()
