"""tt_sweep_static — the sweep for hosts that must not allocate device memory.

Some callers of this library cannot let it call cudaMalloc: a server that has given the whole GPU
to its own memory pool, or a request path that cannot afford cudaMalloc's device-wide
synchronisation.  For them the library keeps the sweep's working memory in a fixed device array.
This test calls that entry point through the C ABI twice in a row (the array keeps the previous
call's contents, so the second call proves it is re-initialised) and requires bit-identical output
to the ordinary entry point tt_sweep_device: same rows, same order, same TaskIds, same counts, same
flag planes.  The file name makes pytest collect it last.
"""
import numpy as np
import pytest

from helpers import tt

pytestmark = pytest.mark.gpu


def make_table(oracle, n, seed=0x57A71C):
    cfg = oracle.default_cfg(seed, n, n_creators=100)
    return tt.TaskTable(0).generate(cfg, 0, n)


def fetch(rows, ids, n):
    return rows[:n].cpu().numpy().view(np.uint32).copy(), ids[:n].cpu().numpy().copy()


def test_static_sweep_matches_the_ordinary_entry_point_twice(oracle):
    import torch

    n = 3_000_000
    cap = 200_000
    ref_table, table = make_table(oracle, n), make_table(oracle, n)
    rows = torch.zeros(cap, dtype=torch.int32, device="cuda")
    ids = torch.zeros((cap, 16), dtype=torch.uint8, device="cuda")

    # what the existing entry point answers
    st0, _ = ref_table.sweep_device(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), ids.data_ptr(), cap)
    want_rows, want_ids = fetch(rows, ids, st0.n_stored)
    assert st0.n_listed == st0.n_stored > 50_000

    # two listing calls in a row: neither changes the table, both must give the same list
    for call in range(2):
        rows.zero_()
        ids.zero_()
        st, rc = table.sweep_static(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), ids.data_ptr(), cap)
        got_rows, got_ids = fetch(rows, ids, st.n_stored)
        assert rc == tt.TT_OK and (st.n_listed, st.n_stored, st.n_marked) == (st0.n_listed, st0.n_stored, 0), call
        assert (got_rows == want_rows).all() and (got_ids == want_ids).all(), call

    # a marking call, then one more: the second finds nothing and leaves the planes alone
    stm, _ = ref_table.sweep_device(tt.MODE_STORE, tt.SWEEP_MARK, rows.data_ptr(), ids.data_ptr(), cap)
    st1, _ = table.sweep_static(tt.MODE_STORE, tt.SWEEP_MARK, rows.data_ptr(), ids.data_ptr(), cap)
    got_rows, got_ids = fetch(rows, ids, st1.n_stored)
    assert (st1.n_listed, st1.n_stored, st1.n_marked) == (stm.n_listed, stm.n_stored, stm.n_marked)
    assert (got_rows == want_rows).all() and (got_ids == want_ids).all()
    st2, _ = table.sweep_static(tt.MODE_STORE, tt.SWEEP_MARK, rows.data_ptr(), ids.data_ptr(), cap)
    assert (st2.n_listed, st2.n_stored, st2.n_marked) == (0, 0, 0)
    assert (table.snapshot_flags()[1] == ref_table.snapshot_flags()[1]).all()
    assert table.count_flags() == ref_table.count_flags()

    # FAKE semantics and the count-only form go through the same array
    a, _ = ref_table.sweep_device(tt.MODE_FAKE, tt.SWEEP_LIST_ONLY, None, None, 0)
    b, _ = table.sweep_static(tt.MODE_FAKE, tt.SWEEP_LIST_ONLY, None, None, 0)
    assert a.n_listed == b.n_listed > st0.n_listed
    ref_table.close()
    table.close()


def test_static_sweep_honours_the_backlog_flag(oracle):
    """TT_SWEEP_DUE_ON_OR_BEFORE through the fixed-array entry point: the same list, marks and refusals as
    tt_sweep_device gives for that flag (it was silently ignored before round 2)."""
    import torch

    n, cap = 2_000_000, tt.STATIC_MAX_HITS
    ref_table, table = make_table(oracle, n, 4242), make_table(oracle, n, 4242)
    rows = torch.zeros(cap, dtype=torch.int32, device="cuda")
    ids = torch.zeros((cap, 16), dtype=torch.uint8, device="cuda")
    flags = tt.SWEEP_LIST_ONLY | tt.SWEEP_DUE_ON_OR_BEFORE
    exact, _ = ref_table.sweep_device(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, None, None, 0)
    a, _ = ref_table.sweep_device(tt.MODE_STORE, flags, rows.data_ptr(), ids.data_ptr(), cap)
    want_rows, want_ids = fetch(rows, ids, a.n_stored)
    assert a.n_stored > 4 * exact.n_listed  # the backlog is far larger than the exact match
    for call in range(2):
        rows.zero_()
        ids.zero_()
        b, _ = table.sweep_static(tt.MODE_STORE, flags, rows.data_ptr(), ids.data_ptr(), cap)
        got_rows, got_ids = fetch(rows, ids, b.n_stored)
        assert (b.n_listed, b.n_stored, b.n_marked) == (a.n_listed, a.n_stored, 0), call
        assert (got_rows == want_rows).all() and (got_ids == want_ids).all(), call
    mflags = tt.SWEEP_MARK | tt.SWEEP_DUE_ON_OR_BEFORE
    am, _ = ref_table.sweep_device(tt.MODE_STORE, mflags, rows.data_ptr(), None, cap)
    bm, _ = table.sweep_static(tt.MODE_STORE, mflags, rows.data_ptr(), None, cap)
    assert (bm.n_listed, bm.n_stored, bm.n_marked) == (am.n_listed, am.n_stored, am.n_marked) and bm.n_marked == a.n_stored
    assert (table.snapshot_flags()[1] == ref_table.snapshot_flags()[1]).all()
    fa, _ = ref_table.sweep_device(tt.MODE_FAKE, flags, None, None, 0)
    fb, _ = table.sweep_static(tt.MODE_FAKE, flags, None, None, 0)
    assert fa.n_listed == fb.n_listed > a.n_listed
    # a marking backlog sweep with the clock behind yesterday is refused by both
    for call in (ref_table.sweep_device, table.sweep_static):
        with pytest.raises(tt.TaskTableError) as e:
            call(tt.MODE_STORE, mflags, rows.data_ptr(), None, cap, run_at=tt.DEFAULT_TODAY - 2 * tt.TICKS_PER_DAY)
        assert e.value.code == tt.TT_ERR_ARG
    ref_table.close()
    table.close()


def test_static_sweep_capacity_behaviour_matches(oracle):
    import torch

    n, cap = 1_000_000, 5_000
    ref_table, table = make_table(oracle, n, 77), make_table(oracle, n, 77)
    rows = torch.zeros(cap, dtype=torch.int32, device="cuda")
    a, rca = ref_table.sweep_device(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), None, cap, allow_capacity=True)
    b, rcb = table.sweep_static(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), None, cap, allow_capacity=True)
    assert rca == rcb == tt.TT_ERR_CAPACITY and a.n_listed == b.n_listed and 0 < b.n_stored <= cap
    ref_table.close()
    table.close()


def test_static_sweep_refuses_sizes_beyond_its_fixed_array(oracle):
    import torch

    table = make_table(oracle, 100_000)
    rows = torch.zeros(16, dtype=torch.int32, device="cuda")
    with pytest.raises(tt.TaskTableError) as e:
        table.sweep_static(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), None, tt.STATIC_MAX_HITS + 1)
    assert e.value.code == tt.TT_ERR_ARG and "at most" in str(e.value)
    table.close()
    big = make_table(oracle, tt.STATIC_MAX_ROWS + 1)
    with pytest.raises(tt.TaskTableError) as e:
        big.sweep_static(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), None, 16)
    assert e.value.code == tt.TT_ERR_ARG and "at most" in str(e.value)
    big.close()
