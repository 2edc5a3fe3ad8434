"""tt_sweep_push — the sweep whose placement kernel also stores the compact hit list (the form
shards exchange).  On one GPU the push target is local memory: the kernel is the
same one that stores into the root GPU's mapped buffer over NVLink (tests/test_multigpu_gpu.py
runs that on 2 GPUs).  Everything must be bit-identical to tt_sweep_device + tt_pack_hits, which
tests/test_sweep_gpu.py checks against the oracle."""
import numpy as np
import pytest

from helpers import tt

pytestmark = pytest.mark.gpu


def run_both(oracle, n, cap, flags, **cfg_kw):
    import torch

    cfg = oracle.default_cfg(0xB0B, n, n_creators=50, **cfg_kw)
    ref, table = tt.TaskTable(0).generate(cfg, 0, n), tt.TaskTable(0).generate(cfg, 0, n)
    nt = (n + 8191) // 8192
    mk = lambda *shape, dt=torch.int32: torch.full(shape, -1, dtype=dt, device="cuda")
    rows0, ids0 = mk(cap), mk(cap, 16, dt=torch.uint8)
    st0, rc0 = ref.sweep_device(tt.MODE_STORE, flags, rows0.data_ptr(), ids0.data_ptr(), cap, allow_capacity=True)
    off0, cnt0 = mk(cap, dt=torch.int16), mk(nt)
    ref.pack_hits(rows0.data_ptr(), st0.n_stored, off0.data_ptr(), cnt0.data_ptr())
    rows1, ids1, off1, cnt1 = mk(cap), mk(cap, 16, dt=torch.uint8), mk(cap + 8, dt=torch.int16), mk(nt)
    total = torch.zeros(1, dtype=torch.int64, device="cuda")
    before = table.count_flags()
    st1, rc1 = table.sweep_push(tt.MODE_STORE, flags, rows1.data_ptr(), ids1.data_ptr(), cap, off1.data_ptr(),
                                cnt1.data_ptr(), total.data_ptr(), allow_capacity=True)
    torch.cuda.synchronize()
    if rc0 == tt.TT_ERR_CAPACITY:
        # which tiles fit depends on the order the CTAs finish in, so two runs may store different
        # row-ordered subsets: check this run against itself and the totals against the other
        k = st1.n_stored
        assert rc1 == rc0 and st1.n_listed == st0.n_listed and 0 < k <= cap and int(total.item()) == k
        assert st1.n_marked == (k if flags == tt.SWEEP_MARK else 0)
        r = rows1[:k].cpu().numpy().view(np.uint32)
        assert (r[:-1] < r[1:]).all()
        assert (off1[:k].cpu().numpy().view(np.uint16) == r % 8192).all()
        assert (np.bincount(r // 8192, minlength=nt) == cnt1.cpu().numpy()).all()
        assert bool((off1[k:] == -1).all()) and bool((rows1[k:] == -1).all()), "wrote past the end of the list"
        assert table.count_flags() == (before[0], before[1] + st1.n_marked)
        for t in (ref, table):
            t.close()
        return st1
    k = st0.n_stored
    assert rc1 == rc0 and (st1.n_listed, st1.n_stored, st1.n_marked) == (st0.n_listed, k, st0.n_marked)
    assert int(total.item()) == k
    assert torch.equal(rows1[:k], rows0[:k]) and torch.equal(ids1[:k], ids0[:k])
    assert torch.equal(off1[:k], off0[:k]) and torch.equal(cnt1, cnt0)
    assert bool((off1[k:] == -1).all()) and bool((rows1[k:] == -1).all()), "wrote past the end of the list"
    assert table.count_flags() == ref.count_flags()
    assert (table.snapshot_flags()[1] == ref.snapshot_flags()[1]).all()
    # push only: no local copy of the result
    table2 = tt.TaskTable(0).generate(cfg, 0, n)
    off2, cnt2 = mk(cap + 8, dt=torch.int16), mk(nt)
    st2, rc2 = table2.sweep_push(tt.MODE_STORE, flags, None, None, cap, off2.data_ptr(), cnt2.data_ptr(),
                                 total.data_ptr(), allow_capacity=True)
    torch.cuda.synchronize()
    assert rc2 == rc0 and st2.n_stored == k and torch.equal(off2[:k], off0[:k]) and torch.equal(cnt2, cnt0)
    for t in (ref, table, table2):
        t.close()
    return st0


@pytest.mark.parametrize("flags", [tt.SWEEP_LIST_ONLY, tt.SWEEP_MARK])
def test_push_matches_sweep_then_pack(oracle, flags):
    st = run_both(oracle, 1_234_567, 60_000, flags)
    assert st.n_stored > 20_000


def test_push_dense_groups_take_the_direct_path(oracle):
    """every row hits: 8 tiles hold 65 536 hits, far beyond the 4096 staged offsets per group"""
    st = run_both(oracle, 5 * 8192 + 100, 50_000, tt.SWEEP_MARK, ppm_due_yesterday=1_000_000, ppm_adversarial=0,
                  ppm_completed=0)
    assert st.n_stored == 5 * 8192 + 100


def test_push_mixed_density_and_ragged_tail(oracle):
    st = run_both(oracle, 8192 * 19 + 1, 120_000, tt.SWEEP_LIST_ONLY, ppm_due_yesterday=400_000, ppm_completed=100_000)
    assert st.n_stored > 40_000


def test_push_capacity_overflow_matches(oracle):
    st = run_both(oracle, 900_000, 9_000, tt.SWEEP_MARK)
    assert 0 < st.n_stored < st.n_listed


def test_push_empty_result(oracle):
    run_both(oracle, 100_000, 1_000, tt.SWEEP_MARK, ppm_due_yesterday=0, ppm_adversarial=0)


def test_push_rejects_bad_targets(oracle):
    import torch

    cfg = oracle.default_cfg(1, 10_000, n_creators=5)
    table = tt.TaskTable(0).generate(cfg, 0, 10_000)
    buf = torch.zeros(4096, dtype=torch.int16, device="cuda")
    cnt = torch.zeros(8, dtype=torch.int32, device="cuda")
    tot = torch.zeros(1, dtype=torch.int64, device="cuda")
    with pytest.raises(tt.TaskTableError):  # misaligned off16
        table.sweep_push(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, None, None, 1000, buf.data_ptr() + 2, cnt.data_ptr(),
                         tot.data_ptr())
    with pytest.raises(tt.TaskTableError):  # null buffer
        table.sweep_push(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, None, None, 1000, buf.data_ptr(), None, tot.data_ptr())
    table.close()


def test_push_into_a_peer_buffer(oracle):
    """tt_peer_alloc hands out zeroed device memory with an exportable handle; used here as the push
    target of its own GPU (opening the handle is a cross-process operation: tests/test_multigpu_gpu.py).
    The slot layout is the one PushExchange uses: [count 256 B][tile counts][offsets]."""
    import torch

    n, cap = 300_000, 20_000
    nt = (n + 8191) // 8192
    o_cnt, o_off = 256, 256 + ((nt * 4 + 255) // 256) * 256
    ptr, handle = tt.peer_alloc(0, o_off + cap * 2)
    assert ptr and ptr % 256 == 0 and len(handle) == 64 and any(handle)
    cfg = oracle.default_cfg(77, n, n_creators=5)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    rows = torch.zeros(cap, dtype=torch.int32, device="cuda")
    st, _ = table.sweep_push(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), None, cap, ptr + o_off, ptr + o_cnt, ptr)
    back = torch.full((cap,), -1, dtype=torch.int32, device="cuda")
    scratch = torch.empty(nt + nt // 1024 + 8, dtype=torch.int32, device="cuda")
    tt.expand_hits(0, ptr + o_cnt, nt, ptr + o_off, scratch.data_ptr(), back.data_ptr())
    torch.cuda.synchronize()
    assert st.n_stored > 5_000 and torch.equal(back[: st.n_stored], rows[: st.n_stored])
    table.close()
    tt.peer_free(0, ptr)
