"""BASELINE.json configs 3, 4 and 5 at their full sizes, against the streaming oracle (rows are generated on the
fly on the host, nothing is stored): order-free summaries per query — count, XOR of the TaskIds, sum / min / max
of the row numbers — plus the ordering property checked on the device results themselves.  TT_FULL_ROWS /
TT_CREATOR_ROWS shrink the tables for a quick run."""
import os

import numpy as np
import pytest

from helpers import be_u64, tt, xor_reduce

pytestmark = pytest.mark.gpu
TICKS = (1 << 62) - 1


def creator_name(k):
    return b"user%07d@mail%d.example" % (k, k % 97)


def query_set(rng, n_creators, n_present):
    """present creators, 10 absent ones, 10 that differ only in case or in one byte."""
    ks = rng.integers(0, n_creators, size=n_present)
    present = [creator_name(int(k)) for k in ks]
    absent = [b"absent%d@nowhere.example" % i for i in range(10)]
    near = []
    for i, k in enumerate(ks[:10]):
        s = bytearray(creator_name(int(k)))
        if i % 2 == 0:
            s[0:4] = b"USER"            # same letters, other case: ordinal comparison says different
        else:
            s[4] = ord("1")             # one byte off: creator numbers have 7 digits and start with 0, so nobody has this name
        near.append(bytes(s))
    return present + absent + near


def check_results(results, table, want, n_present):
    hits = 0
    for q, (res, w) in enumerate(zip(results, want)):
        assert res.count == w.count, q
        if q >= n_present:
            assert res.count == 0, "absent and near-miss creators have no tasks"
            continue
        rows, ids = res.rows(), res.ids()
        assert (xor_reduce(be_u64(ids, False)), xor_reduce(be_u64(ids, True))) == (w.xor_id_hi, w.xor_id_lo), q
        assert int(rows.astype(np.uint64).sum()) == w.sum_rows and rows.min() == w.min_row and rows.max() == w.max_row, q
        # OrderByDescending(TaskCreatedOn), stable: ticks never increase, equal ticks keep row order
        ticks = (res.field(tt.FIELD_CREATED_RAW) & np.uint64(TICKS)).astype(np.int64)
        d = np.diff(ticks)
        assert (d <= 0).all() and (np.diff(rows.astype(np.int64))[d == 0] > 0).all(), q
        hits += res.count
    return hits


def test_config4_creator_filter_100m_rows_1m_creators(oracle):
    """GetTasksByCreator (FakeTasksManager.cs:68-72, TasksStoreManager.cs:54-69) over 100 M tasks of 1 M creators:
    1 000 present + 10 absent + 10 near-miss queries, as one batch and (a sample) one by one."""
    n = int(os.environ.get("TT_CREATOR_ROWS", 100_000_000))
    n_creators = 1_000_000
    cfg = oracle.default_cfg(0x5EED0003, n, n_creators=n_creators)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    queries = query_set(np.random.default_rng(4), n_creators, 1000)
    want = oracle.stream_creators(cfg, 0, n, queries)
    results = table.get_by_creators(queries)
    hits = check_results(results, table, want, 1000)
    assert hits > 50 * n // 1_000_000  # about n / n_creators tasks per creator
    for q in (0, 1, 999, 1000, 1010, 1011):  # the single-query path (its own kernel instantiation)
        single = table.get_by_creator(queries[q])
        assert single.rows().tolist() == results[q].rows().tolist() and (single.ids() == results[q].ids()).all()
    table.close()


def test_config5_mixed_batch_1b_rows(oracle):
    """The mixed cron batch on the 1 B-task table: the STORE-mode tick, then 64 creator filters back to back."""
    n = int(os.environ.get("TT_FULL_ROWS", 1_000_000_000))
    cfg = oracle.default_cfg(0x5EED0002, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    want = oracle.stream_sweep(cfg, 0, n, tt.MODE_STORE)
    res, n_marked, threw = table.sweep(tt.MODE_STORE)
    assert (res.count, n_marked, threw) == (want.n_listed, want.n_marked, False)
    assert res.digest() == (want.xor_id_hi, want.xor_id_lo, want.sum_rows)
    assert table.count_flags() == (want.completed, want.overdue_after)
    res.free()
    rng = np.random.default_rng(5)
    queries = [creator_name(int(k)) for k in rng.integers(0, 1_000_000, size=64)]
    wantc = oracle.stream_creators(cfg, 0, n, queries)
    results = table.get_by_creators(queries)
    assert check_results(results, table, wantc, 64) > 32 * n // 1_000_000
    table.close()


def test_config3_one_table_over_shards_full_size(oracle):
    """BASELINE.json config 3's table (1 B tasks, TaskId-range shards) through tt_shards_* with all shards on this
    GPU: the cron tick's list (count, XOR of TaskIds, sum of GLOBAL rows) and the per-shard planes' popcounts."""
    n = int(os.environ.get("TT_FULL_ROWS", 1_000_000_000))
    shards = 4
    cfg = oracle.default_cfg(0x5EED0002, n)
    st = tt.ShardedTable([0] * shards).generate(cfg, n)
    want = oracle.stream_sweep(cfg, 0, n, tt.MODE_STORE)
    res, n_marked, threw = st.sweep(tt.MODE_STORE)
    assert (res.count, n_marked, threw) == (want.n_listed, want.n_marked, False)
    assert res.digest() == (want.xor_id_hi, want.xor_id_lo, want.sum_rows)
    comp = over = 0
    for g in range(shards):
        c, o = st.shard(g)[0].count_flags()
        comp, over = comp + c, over + o
    assert (comp, over) == (want.completed, want.overdue_after)
    res.free()
    res2, n2, _ = st.sweep(tt.MODE_STORE)
    assert (res2.count, n2) == (0, 0)
    st.close()
