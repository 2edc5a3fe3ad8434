"""The generalised predicate (an extension with no reference counterpart): "due yesterday
or earlier" instead of "due exactly yesterday at 00:00:00".  Checked against the oracle's statement of the
same LINQ clause (oracle.cpp orc_backlog); everything around the predicate — flag tests, order, the cron
handler's filter, marking, FAKE mode's throw — is the reference's."""
import numpy as np
import pytest

from helpers import load_from_oracle, tt

pytestmark = pytest.mark.gpu
DAY = tt.TICKS_PER_DAY


def synth_pair(oracle, n, seed=0xBAC, **kw):
    cfg = oracle.default_cfg(seed, n, n_creators=20, **kw)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    return ot, tt.TaskTable(0).generate(cfg, 0, n)


@pytest.mark.parametrize("mode", [tt.MODE_FAKE, tt.MODE_STORE])
def test_listing(oracle, mode):
    ot, table = synth_pair(oracle, 400_000)
    want, _, _ = ot.backlog(mode)
    res = table.get_due_on_or_before(mode)
    assert res.count == want.size > 80_000  # about half the due dates lie in the past
    assert (res.rows() == want).all()       # FAKE: list order; STORE: ascending TaskCreatedOn, ties by row
    ids_all, _, _ = ot.export()
    assert (res.ids() == ids_all[want]).all()
    exact = table.get_yesterdays_due(mode)
    assert np.isin(exact.rows(), res.rows()).all() and exact.count < res.count
    assert table.count_flags() == (int(ot.flags()[0].sum()), int(ot.flags()[1].sum()))  # listing changes nothing
    table.close()


def test_tick_marks_the_whole_backlog_and_is_idempotent(oracle):
    ot, table = synth_pair(oracle, 300_000)
    res, n_marked, threw = table.sweep_backlog(tt.MODE_STORE)
    want, n_over, othrew = ot.backlog(tt.MODE_STORE, mark=True)
    assert (threw, othrew) == (False, False)
    assert res.count == want.size == n_marked == n_over > 50_000
    assert (res.rows() == np.sort(want)).all()  # the fused tick lists in row order, like tt_sweep
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    assert (gc == c).all() and (go == o).all()
    res2, n2, _ = table.sweep_backlog(tt.MODE_STORE)
    assert (res2.count, n2) == (0, 0)
    # and the exact sweep finds nothing left either: its rows were part of the backlog
    assert table.sweep(tt.MODE_STORE)[0].count == 0
    table.close()


def test_fake_mode_lists_then_throws_and_changes_nothing(oracle):
    ot, table = synth_pair(oracle, 50_000)
    before = table.count_flags()
    res, n_marked, threw = table.sweep_backlog(tt.MODE_FAKE)
    want, n_over, othrew = ot.backlog(tt.MODE_FAKE, mark=True)
    assert threw and othrew and n_marked == 0 and (res.rows() == want).all() and n_over == want.size
    assert table.count_flags() == before
    table.close()


def test_edge_rows_time_of_day_kind_padding_and_tombstones(oracle):
    """rows the exact match never lists: a time of day, a Kind suffix, a date long past; plus rows that must
    never match: due today or later, deleted tasks, the padding behind the last row"""
    import uuid

    y = tt.DEFAULT_TODAY - DAY
    dues = [y, y + 1, y | (1 << 62), (y + 5 * 36_000_000_000) | (2 << 62), y - 400 * DAY, 1 * DAY, 0,
            tt.DEFAULT_TODAY, tt.DEFAULT_TODAY + 1, tt.DEFAULT_TODAY + 30 * DAY, (y + DAY - 1), y - 1]
    ot = oracle.OracleTable()
    for i, d in enumerate(dues):
        ot.add(np.frombuffer(uuid.UUID(int=i + 1).bytes, dtype=np.uint8).copy(), b"n", b"c@d", b"a",
               (700_000 * DAY + (len(dues) - i)) | (1 << 62), d, completed=(i == 4), overdue=False)
    table = load_from_oracle(ot)
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        want, _, _ = ot.backlog(mode)
        assert table.get_due_on_or_before(mode).rows().tolist() == want.tolist()
    assert ot.backlog(tt.MODE_FAKE)[0].tolist() == [0, 1, 2, 3, 4, 5, 6, 10, 11]
    ids_all, _, _ = ot.export()
    table.delete_task(ids_all[5])
    ot.delete_task(oracle.MODE_STORE, ids_all[5])
    res, n_marked, _ = table.sweep_backlog(tt.MODE_STORE)
    want, n_over, _ = ot.backlog(tt.MODE_STORE, mark=True)
    assert res.rows().tolist() == sorted(want.tolist()) == [0, 1, 2, 3, 6, 10, 11] and n_marked == n_over == 7
    assert (table.flags_bytes()[1] == ot.flags()[1]).all()
    table.close()


def test_device_form_capacity_and_clock_checks(oracle):
    import torch

    ot, table = synth_pair(oracle, 200_000)
    want, _, _ = ot.backlog(tt.MODE_STORE)
    cap = want.size + 10
    rows = torch.zeros(cap, dtype=torch.int32, device="cuda")
    ids = torch.zeros((cap, 16), dtype=torch.uint8, device="cuda")
    flags = tt.SWEEP_LIST_ONLY | tt.SWEEP_DUE_ON_OR_BEFORE
    st, _ = table.sweep_device(tt.MODE_STORE, flags, rows.data_ptr(), ids.data_ptr(), cap)
    assert st.n_listed == st.n_stored == want.size
    assert (rows[: st.n_stored].cpu().numpy().view(np.uint32) == np.sort(want)).all()
    # capacity is granted tile by tile (8192 rows, about 2000 hits each here)
    st, rc = table.sweep_device(tt.MODE_STORE, flags, rows.data_ptr(), ids.data_ptr(), 9000, allow_capacity=True)
    assert rc == tt.TT_ERR_CAPACITY and st.n_listed == want.size and 0 < st.n_stored <= 9000
    # a marking range sweep needs runAt past yesterday: the handler's test would otherwise differ per row
    with pytest.raises(tt.TaskTableError):
        table.sweep_device(tt.MODE_STORE, tt.SWEEP_MARK | tt.SWEEP_DUE_ON_OR_BEFORE, rows.data_ptr(), ids.data_ptr(), cap,
                           run_at=tt.DEFAULT_TODAY - 2 * DAY)
    table.close()
