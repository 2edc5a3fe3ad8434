"""GPU parity: the CUDA sweep (through the C ABI) against the CPU restatement on the same rows.

Bit-exact bar: identical hit rows in identical order, identical TaskIds, identical flag
planes afterwards, identical marked counts and identical error behaviour, in both reference
modes (FAKE = FakeTasksManager, STORE = TasksStoreManager).
"""
import numpy as np
import pytest

from helpers import be_u64, load_from_oracle, tt, xor_reduce

pytestmark = pytest.mark.gpu

TPD = tt.TICKS_PER_DAY
TODAY = tt.DEFAULT_TODAY
RUN_AT = tt.DEFAULT_RUN_AT
MODES = [tt.MODE_FAKE, tt.MODE_STORE]


def check_sweep(oracle, ot, table, mode, today=TODAY, run_at=RUN_AT, offset_min=0):
    ids_all, _, _ = ot.export()
    threw, listed, n_over = ot.cron_job(mode, today, run_at, offset_min)
    res, n_marked, gthrew = table.sweep(mode, today, run_at)
    rows = res.rows()
    # the fused sweep lists in row order; the reference's STORE list is ordered by TaskCreatedOn
    # (checked separately through get_yesterdays_due) — same set, and here same order after sorting
    assert res.count == listed.size
    assert (np.sort(rows) == np.sort(listed)).all()
    assert (rows[:-1] < rows[1:]).all()
    assert (res.ids() == ids_all[rows]).all()
    assert gthrew == threw
    assert n_marked == (n_over if mode == tt.MODE_STORE else 0)
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    assert (gc == c).all() and (go == o).all()
    assert table.count_flags() == (int(c.sum()), int(o.sum()))
    res.free()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("n", [0, 1, 31, 33, 8191, 8192, 8193, 100_000, 1_000_003])
def test_sweep_matches_oracle_generated(oracle, mode, n):
    cfg = oracle.default_cfg(0x5EED0001, max(n, 1), n_creators=1000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    assert len(table) == n
    # the device generator produced the same rows as the host one
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    assert (gc == c).all() and (go == o).all()
    check_sweep(oracle, ot, table, mode)
    # second tick the same day finds nothing new in STORE mode; FAKE mode lists the same rows again
    check_sweep(oracle, ot, table, mode)
    table.close()


@pytest.mark.parametrize("mode", MODES)
def test_sweep_matches_oracle_loaded_shard(oracle, mode):
    """rows of a middle shard, loaded through tt_load_soa instead of generated on the device"""
    n = 50_000
    cfg = oracle.default_cfg(0x5EED0002, 4 * n, n_creators=300)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 2 * n, n)
    table = load_from_oracle(ot)
    check_sweep(oracle, ot, table, mode)
    gen = tt.TaskTable(0).generate(cfg, 2 * n, n)
    a, b = table.get_yesterdays_due(tt.MODE_FAKE), gen.get_yesterdays_due(tt.MODE_FAKE)
    assert (a.rows() == b.rows()).all() and (a.ids() == b.ids()).all()
    table.close()
    gen.close()


@pytest.mark.parametrize("mode", MODES)
def test_get_yesterdays_due_order(oracle, mode):
    n = 300_000
    cfg = oracle.default_cfg(77, n, n_creators=100, ppm_created_dup=200_000)  # many TaskCreatedOn ties
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    want = ot.get_yesterdays_due(mode, TODAY)
    res = table.get_yesterdays_due(mode, TODAY)
    rows = res.rows()
    assert rows.size == want.size and (rows == want).all()  # same order, ties included
    ids_all, created, _ = ot.export()
    assert (res.ids() == ids_all[rows]).all()
    assert (res.field(tt.FIELD_CREATED_RAW) == created[rows]).all()
    # GetYesterdaysDueTasks does not modify the table
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    assert (gc == c).all() and (go == o).all()
    table.close()


def build_kat(oracle):
    from test_oracle import build_kat as bk

    return bk(oracle)


@pytest.mark.parametrize("mode", MODES)
def test_hand_built_cases(oracle, mode):
    ot = build_kat(oracle)
    table = load_from_oracle(ot)
    want = ot.get_yesterdays_due(mode, TODAY)
    got = table.get_yesterdays_due(mode, TODAY).rows()
    assert got.tolist() == want.tolist() == ([0, 2, 3, 4, 5, 8, 9] if mode == tt.MODE_FAKE else [8, 0, 9])
    check_sweep(oracle, ot, table, mode)
    table.close()


@pytest.mark.parametrize("mode", MODES)
def test_utc_local_skew_marks_nothing(oracle, mode):
    # local zone ahead of UTC: at local 00:05 the UTC date is still yesterday
    ot = build_kat(oracle)
    table = load_from_oracle(ot)
    run_at = TODAY + 5 * 60 * 10_000_000 - 2 * 3600 * 10_000_000
    check_sweep(oracle, ot, table, mode, TODAY, run_at, 120)
    assert table.count_flags()[1] == 1
    table.close()


def test_unfused_get_filter_mark_equals_fused(oracle):
    n = 200_000
    cfg = oracle.default_cfg(5, n, n_creators=100)
    fused = tt.TaskTable(0).generate(cfg, 0, n)
    unfused = tt.TaskTable(0).generate(cfg, 0, n)
    res, n_marked, _ = fused.sweep(tt.MODE_STORE)
    lst = unfused.get_yesterdays_due(tt.MODE_STORE)
    due = lst.field(tt.FIELD_DUE_RAW)
    keep = (RUN_AT // TPD) > ((due & np.uint64((1 << 62) - 1)) // np.uint64(TPD))  # cron predicate on the host
    ids = lst.ids()[keep]
    marked, missing = unfused.mark_overdue(tt.MODE_STORE, ids)
    assert (marked, missing) == (n_marked, 0) and n_marked == res.count > 0
    assert (fused.snapshot_flags()[1] == unfused.snapshot_flags()[1]).all()
    fused.close()
    unfused.close()


def test_mark_overdue_by_id(oracle):
    ot = build_kat(oracle)  # ids are not in TaskId order? they are; shuffle to exercise the index
    cols = __import__("helpers").oracle_columns(ot)
    perm = np.array([3, 0, 7, 10, 1, 9, 2, 8, 4, 6, 5])
    shuffled = {k: (v[perm] if isinstance(v, np.ndarray) else v) for k, v in cols.items()}
    table = tt.TaskTable(0).load_soa(**shuffled)
    from test_oracle import gid

    ids = np.stack([np.frombuffer(gid(i), dtype=np.uint8) for i in (2, 7, 7, 99)])
    assert table.mark_overdue(tt.MODE_STORE, ids) == (3, 1)
    _, o = table.flags_bytes()
    want = np.array([0, 1, 0, 0, 0, 1, 1, 0, 0, 0, 0], dtype=np.uint8)[perm]
    assert (o == want).all()
    with pytest.raises(tt.NotImplementedInReference):
        table.mark_overdue(tt.MODE_FAKE, ids)
    table.close()


@pytest.mark.parametrize("mode", MODES)
def test_every_row_hits(oracle, mode):
    """worst case for the compaction: all rows due yesterday and pending"""
    n = 3 * 8192 + 17
    cfg = oracle.default_cfg(3, n, n_creators=10, ppm_due_yesterday=1_000_000, ppm_adversarial=0, ppm_completed=0)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    check_sweep(oracle, ot, table, mode)
    assert table.count_flags()[1] == (n if mode == tt.MODE_STORE else 0)
    table.close()


def test_no_row_hits(oracle):
    n = 100_000
    cfg = oracle.default_cfg(3, n, n_creators=10, ppm_due_yesterday=0, ppm_adversarial=0)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    for mode in MODES:
        res, n_marked, threw = table.sweep(mode)
        assert (res.count, n_marked, threw) == (0, 0, False)
    table.close()


def test_device_sweep_capacity_and_resume(oracle):
    import torch

    n = 500_000
    cfg = oracle.default_cfg(9, n, n_creators=10)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    _, listed, _ = ot.cron_job(tt.MODE_STORE)
    want = np.sort(listed)
    ids_all, _, _ = ot.export()
    table = tt.TaskTable(0).generate(cfg, 0, n)
    cap = want.size // 3 + 5
    rows = torch.zeros(want.size + 16, dtype=torch.int32, device="cuda")
    ids = torch.zeros((want.size + 16, 16), dtype=torch.uint8, device="cuda")
    # a listing sweep that does not fit changes nothing and reports the size it needs
    st, rc = table.sweep_device(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), ids.data_ptr(), cap,
                                allow_capacity=True)
    assert rc == tt.TT_ERR_CAPACITY and st.n_listed == want.size and 0 < st.n_stored <= cap and st.n_marked == 0
    part = rows[: st.n_stored].cpu().numpy().view(np.uint32)
    assert (part[:-1] < part[1:]).all() and np.isin(part, want).all()  # a row-ordered subset
    assert table.count_flags()[1] == int(ot.flags()[1].sum()) - want.size  # nothing marked yet
    # a marking sweep that does not fit marks what it stored; calling again yields the rest
    got, passes = 0, 0
    while True:
        st, rc = table.sweep_device(tt.MODE_STORE, tt.SWEEP_MARK, rows.data_ptr() + 4 * got, ids.data_ptr() + 16 * got,
                                    cap, allow_capacity=True)
        passes += 1
        run = rows[got: got + st.n_stored].cpu().numpy().view(np.uint32)
        assert (run[:-1] < run[1:]).all()
        assert (ids[got: got + st.n_stored].cpu().numpy() == ids_all[run]).all()
        got += st.n_stored
        assert st.n_marked == st.n_stored
        if rc == tt.TT_OK:
            assert st.resume_row == n
            break
        assert rc == tt.TT_ERR_CAPACITY and st.n_listed > st.n_stored
    assert passes >= 3 and got == want.size
    assert (np.sort(rows[:got].cpu().numpy().view(np.uint32)) == want).all()
    _, o = ot.flags()
    assert (table.flags_bytes()[1] == o).all()
    # count-only form: no output buffers
    table2 = tt.TaskTable(0).generate(cfg, 0, n)
    st, _ = table2.sweep_device(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, None, None, 0)
    assert st.n_listed == want.size and st.n_stored == 0
    table.close()
    table2.close()


@pytest.mark.parametrize("mode", MODES)
def test_owned_result_grows_when_the_first_guess_is_too_small(oracle, mode):
    """> n/16 hits: tt_sweep has to regrow its result (and, after marking, merge two runs)"""
    n = 2_000_000
    cfg = oracle.default_cfg(31, n, n_creators=10, ppm_due_yesterday=400_000, ppm_completed=100_000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    check_sweep(oracle, ot, table, mode)
    table.close()


def test_argument_errors():
    table = tt.TaskTable(0)
    with pytest.raises(tt.TaskTableError) as e:
        table.sweep(tt.MODE_STORE, TODAY + 1)  # not a midnight
    assert e.value.code == tt.TT_ERR_ARG
    with pytest.raises(tt.TaskTableError):
        table.sweep(7)
    res, n_marked, threw = table.sweep(tt.MODE_STORE)  # empty table
    assert (res.count, n_marked, threw) == (0, 0, False)
    table.close()


@pytest.mark.parametrize("mode", MODES)
def test_full_size_properties_1b(oracle, mode):
    """BASELINE.json's 1 B-row table: size-independent properties against the streaming oracle —
    hit count, XOR of the hit TaskIds, sum of hit rows, popcount of both planes after the tick,
    and idempotence of a second tick."""
    import os

    n = int(os.environ.get("TT_FULL_ROWS", 1_000_000_000))
    cfg = oracle.default_cfg(0x5EED0002, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    want = oracle.stream_sweep(cfg, 0, n, mode)
    res, n_marked, threw = table.sweep(mode)
    assert res.count == want.n_listed and threw == bool(want.threw)
    assert n_marked == (want.n_marked if mode == tt.MODE_STORE else 0)
    assert res.digest() == (want.xor_id_hi, want.xor_id_lo, want.sum_rows)
    assert table.count_flags() == (want.completed, want.overdue_after)
    res.free()
    res2, n2, _ = table.sweep(mode)
    assert (res2.count, n2) == ((0, 0) if mode == tt.MODE_STORE else (want.n_listed, 0))
    table.close()


def test_compact_hit_list_roundtrip(oracle):
    """tt_pack_hits / tt_expand_hits: 2-byte in-tile offsets + per-tile counts rebuild the row list"""
    import torch

    n = 700_001
    cfg = oracle.default_cfg(41, n, n_creators=10)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    cap = 40_000
    rows = torch.zeros(cap, dtype=torch.int32, device="cuda")
    st, _ = table.sweep_device(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), None, cap)
    nt_cap = (n + 8191) // 8192
    off16 = torch.zeros(cap, dtype=torch.int16, device="cuda")
    counts = torch.zeros(nt_cap, dtype=torch.int32, device="cuda")
    nt = table.pack_hits(rows.data_ptr(), st.n_stored, off16.data_ptr(), counts.data_ptr())
    torch.cuda.synchronize()
    assert nt == nt_cap and int(counts.sum()) == st.n_stored
    want = rows[: st.n_stored].cpu().numpy().view(np.uint32)
    assert (off16[: st.n_stored].cpu().numpy().view(np.uint16) == want % 8192).all()
    assert (np.bincount(want // 8192, minlength=nt) == counts.cpu().numpy()).all()
    back = torch.full((cap,), -1, dtype=torch.int32, device="cuda")
    scratch = torch.empty(nt + nt // 1024 + 8, dtype=torch.int32, device="cuda")
    tt.expand_hits(0, counts.data_ptr(), nt, off16.data_ptr(), scratch.data_ptr(), back.data_ptr())
    torch.cuda.synchronize()
    assert (back[: st.n_stored].cpu().numpy().view(np.uint32) == want).all()
    table.close()


def _ordered_list_of(oracle, created_ticks):
    """A table whose rows are all due yesterday with the given creation times -> (oracle's ordered rows, table)."""
    n = len(created_ticks)
    cfg = oracle.default_cfg(5, n, n_creators=3, ppm_due_yesterday=1_000_000, ppm_adversarial=0, ppm_completed=0,
                             ppm_overdue_past=0)
    src = oracle.OracleTable()
    src.add_synth(cfg, 0, n)
    ids, _, due = src.export()
    created = np.asarray(created_ticks, dtype=np.uint64) | np.uint64(1 << 62)  # Kind = Utc
    ot = oracle.OracleTable()
    for i in range(n):
        ot.add(ids[i].tobytes(), "t", "c", "a", int(created[i]), int(due[i]))
    table = tt.TaskTable(0).load_soa(ids, created, due, np.zeros(n, np.uint8), np.zeros(n, np.uint8), np.zeros(n, np.uint32),
                                     [b"c"])
    return ot.get_yesterdays_due(tt.MODE_STORE, TODAY), table


def test_order_by_created_on_wide_keys_with_near_ties(oracle):
    """Creation times that differ in many bits are sorted on a prefix of them (24 bits for 60 000 keys); groups that share a prefix and
    came in out of order (here: bursts of tasks 1-200 ticks apart, shuffled) are put right afterwards."""
    rng = np.random.default_rng(11)
    base = 638_000_000_000_000_000
    span = 1 << 52
    n_bursts, n = 3_000, 60_000
    starts = rng.integers(0, span, n_bursts, dtype=np.int64)
    ticks = np.concatenate([rng.integers(0, span, n - 8 * n_bursts, dtype=np.int64),
                            (starts[:, None] + rng.integers(0, 200, (n_bursts, 8), dtype=np.int64)).ravel(), [0, span - 1]])
    ticks[:5000] = ticks[5000:10000]  # exact ties too
    rng.shuffle(ticks)
    want, table = _ordered_list_of(oracle, base + ticks)
    res = table.get_yesterdays_due(tt.MODE_STORE, TODAY)
    assert (res.rows() == want).all()
    bits, passes, fixed, fell_back = table.last_order_info()
    assert (bits, passes, fell_back) == (52, 3, False) and fixed > 1000  # 60 000 keys: a 24-bit prefix
    table.close()


def test_order_by_created_falls_back_on_long_groups(oracle):
    """Hundreds of distinct creation times inside one prefix, out of order: ordinary passes finish the sort."""
    rng = np.random.default_rng(12)
    base = 638_000_000_000_000_000
    span = 1 << 50
    ticks = np.concatenate([rng.integers(0, span, 20_000, dtype=np.int64), 12345 + rng.permutation(700), [0, span - 1]])
    rng.shuffle(ticks)
    want, table = _ordered_list_of(oracle, base + ticks)
    res = table.get_yesterdays_due(tt.MODE_STORE, TODAY)
    assert (res.rows() == want).all()
    bits, passes, _, fell_back = table.last_order_info()
    assert bits == 50 and fell_back and passes == 3 + 4 + 3
    table.close()


def test_order_by_created_in_creation_order_needs_no_fixing(oracle):
    """A table filled in creation order (dense timestamps, many ties) has no group out of order."""
    ticks = 638_000_000_000_000_000 + np.cumsum(np.random.default_rng(13).integers(0, 3, 50_000)).astype(np.int64)
    ticks[-1] += 1 << 45
    want, table = _ordered_list_of(oracle, ticks)
    res = table.get_yesterdays_due(tt.MODE_STORE, TODAY)
    assert (res.rows() == want).all() and (want == np.arange(len(ticks))).all()
    bits, passes, fixed, fell_back = table.last_order_info()
    assert (bits, passes, fixed, fell_back) == (46, 3, 0, False)
    table.close()
