"""tt_tick_host — the cron tick for a caller whose task state and results live in host memory: flag planes
up, hits down, pipelined over chunks of the table.  Must equal tt_restore_flags + tt_sweep + tt_result_rows /
tt_result_ids on a twin table, whatever the chunking and however small the staging buffers are."""
import numpy as np
import pytest

from helpers import tt

pytestmark = pytest.mark.gpu


def twin(oracle, n, **kw):
    cfg = oracle.default_cfg(0x71C, n, n_creators=10, **kw)
    return tt.TaskTable(0).generate(cfg, 0, n), tt.TaskTable(0).generate(cfg, 0, n)


@pytest.mark.parametrize("flags", [tt.SWEEP_MARK, tt.SWEEP_LIST_ONLY, tt.SWEEP_MARK | tt.SWEEP_DUE_ON_OR_BEFORE])
@pytest.mark.parametrize("chunks", [1, 3, 8, 64])
def test_tick_host_equals_the_unpipelined_calls(oracle, flags, chunks):
    import torch

    n = 1_000_003
    ref, table = twin(oracle, n)
    comp, over = ref.snapshot_flags()
    # the state "arrives" from the host: start the device planes from garbage to prove they are replaced
    junk = np.full_like(comp, 0xA5)
    table.restore_flags(junk, junk)
    if flags & tt.SWEEP_DUE_ON_OR_BEFORE:
        want, n_marked, _ = ref.sweep_backlog(tt.MODE_STORE)
    elif flags & tt.SWEEP_MARK:
        want, n_marked, _ = ref.sweep(tt.MODE_STORE)
    else:
        want, n_marked = ref.get_yesterdays_due(tt.MODE_FAKE), 0
    mode = tt.MODE_STORE if flags else tt.MODE_FAKE
    want_rows, want_ids = np.sort(want.rows()), None
    order = np.argsort(want.rows(), kind="stable")
    want_ids = want.ids()[order]
    cap = want.count + 100
    h_c = torch.from_numpy(comp.copy()).pin_memory()
    h_o = torch.from_numpy(over.copy()).pin_memory()
    h_rows = torch.full((cap,), -1, dtype=torch.int32).pin_memory()
    h_ids = torch.zeros((cap, 16), dtype=torch.uint8).pin_memory()
    for tick in range(2):
        st, rc = table.tick_host(mode, flags, h_c.data_ptr(), h_o.data_ptr(), h_rows.data_ptr(), h_ids.data_ptr(), cap, chunks)
        assert rc == tt.TT_OK and st.n_listed == st.n_stored == want.count and st.n_marked == n_marked
        assert (h_rows[: st.n_stored].numpy().view(np.uint32) == want_rows).all()
        assert (h_ids[: st.n_stored].numpy() == want_ids).all()
        assert int(h_rows[st.n_stored]) == -1
        assert (table.snapshot_flags()[1] == ref.snapshot_flags()[1]).all() and (table.snapshot_flags()[0] == comp).all()
        # the second tick starts from the same host state again and must give the same answer
    ref.close()
    table.close()


def test_more_hits_than_the_staging_buffers(oracle):
    """20 % of the rows hit: a chunk's hits exceed the staging buffer (1/16 of its rows), so a listing tick
    regrows it and a marking tick takes the chunk in pieces"""
    import torch

    n = 1_500_000
    for flags in (tt.SWEEP_LIST_ONLY, tt.SWEEP_MARK):
        ref, table = twin(oracle, n, ppm_due_yesterday=400_000, ppm_completed=500_000)
        want, n_marked, _ = ref.sweep(tt.MODE_STORE) if flags else (ref.get_yesterdays_due(tt.MODE_STORE), 0, False)
        want_rows = np.sort(want.rows())
        assert want.count > 250_000
        cap = want.count
        h_rows = torch.zeros(cap, dtype=torch.int32).pin_memory()
        st, rc = table.tick_host(tt.MODE_STORE, flags, None, None, h_rows.data_ptr(), None, cap, 4)
        assert rc == tt.TT_OK and st.n_stored == want.count and st.n_marked == n_marked
        assert (h_rows.numpy().view(np.uint32) == want_rows).all()
        assert table.count_flags() == ref.count_flags()
        ref.close()
        table.close()


def test_host_buffers_too_small(oracle):
    import torch

    ref, table = twin(oracle, 400_000)
    want, n_marked, _ = ref.sweep(tt.MODE_STORE)
    cap = want.count // 2
    h_rows = torch.zeros(cap, dtype=torch.int32).pin_memory()
    st, rc = table.tick_host(tt.MODE_STORE, tt.SWEEP_MARK, None, None, h_rows.data_ptr(), None, cap, 4, allow_capacity=True)
    assert rc == tt.TT_ERR_CAPACITY and st.n_listed == want.count and st.n_stored == cap and st.n_marked == n_marked
    assert (h_rows.numpy().view(np.uint32) == np.sort(want.rows())[:cap]).all()
    assert table.count_flags() == ref.count_flags()
    ref.close()
    table.close()
