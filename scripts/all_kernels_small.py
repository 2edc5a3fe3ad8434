"""Every kernel of the library once, on a small table whose row count is not a tile multiple, each
answer checked against the oracle: a one-minute end-to-end check (python scripts/all_kernels_small.py).
(compute-sanitizer is closed on the GPU pool, so bad accesses have to show up as wrong answers.)"""
import importlib
import os
import sys
import uuid

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable
from oracle import oracle as orc  # noqa: E402


def main():
    n = 3 * 8192 + 777
    cfg = orc.default_cfg(0x5A17, n, n_creators=40, n_names=200)
    ot = orc.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n).generate_payload(cfg, 0)
    ids_all, _, _ = ot.export()
    # listings, both modes, exact and generalised predicate; JSON of each in both encoders
    for mode in (tt.MODE_FAKE, tt.MODE_STORE):
        for get, want in ((table.get_yesterdays_due, ot.get_yesterdays_due(mode)),
                          (table.get_due_on_or_before, ot.backlog(mode)[0])):
            res = get(mode)
            assert (res.rows() == want).all() and (res.ids() == ids_all[want]).all()
            for flavor in (tt.JSON_RESPONSE, tt.JSON_DOCUMENT):
                assert res.to_json(flavor, 60) == ot.to_json(want, flavor, 60)
    # creator filters: one, a batch, an absent one
    who = [ot.created_by(i) for i in range(5)] + [b"nobody@nowhere"]
    for w, r in zip(who, table.get_by_creators(who)):
        assert r.rows().tolist() == ot.get_by_creator(orc.MODE_FAKE, w).tolist()
    assert table.get_by_creator(who[0]).rows().tolist() == ot.get_by_creator(orc.MODE_FAKE, who[0]).tolist()
    # point operations
    new_id = np.frombuffer(uuid.UUID(int=7).bytes, dtype=np.uint8).copy()
    table.insert_task(new_id, 700_000 * tt.TICKS_PER_DAY, tt.DEFAULT_TODAY - tt.TICKS_PER_DAY, b"fresh@creator")
    ot.add(new_id, b"", b"fresh@creator", b"", 700_000 * tt.TICKS_PER_DAY, tt.DEFAULT_TODAY - tt.TICKS_PER_DAY)
    assert table.get_by_id(new_id).rows().tolist() == [n]
    assert table.mark_completed(ids_all[5]) and ot.mark_completed(orc.MODE_STORE, ids_all[5])
    table.update_task(ids_all[6], tt.DEFAULT_TODAY - tt.TICKS_PER_DAY)
    ot.update_task(orc.MODE_STORE, ids_all[6], "", "", tt.DEFAULT_TODAY - tt.TICKS_PER_DAY)
    assert table.delete_task(ids_all[9]) and ot.delete_task(orc.MODE_STORE, ids_all[9])
    marked, missing = table.mark_overdue(tt.MODE_STORE, ids_all[100:140])
    ot.mark_overdue(orc.MODE_STORE, ids_all[100:140])
    assert (marked, missing) == (40, 0)
    # device-resident sweep with the fused push, then the static variant, then the owned tick
    cap = 4096
    rows = torch.zeros(cap, dtype=torch.int32, device="cuda")
    ids = torch.zeros((cap, 16), dtype=torch.uint8, device="cuda")
    off16 = torch.zeros(cap + 8, dtype=torch.int16, device="cuda")
    nt = (len(table) + 8191) // 8192 + 1
    counts = torch.zeros(nt + 8, dtype=torch.int32, device="cuda")
    total = torch.zeros(1, dtype=torch.int64, device="cuda")
    st, _ = table.sweep_push(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), ids.data_ptr(), cap, off16.data_ptr(),
                             counts.data_ptr(), total.data_ptr())
    want = np.sort(ot.get_yesterdays_due(orc.MODE_STORE))
    assert st.n_stored == want.size == int(total.item())
    assert (rows[: st.n_stored].cpu().numpy().view(np.uint32) == want).all()
    back = torch.zeros(cap, dtype=torch.int32, device="cuda")
    scratch = torch.zeros(nt + nt // 1024 + 8, dtype=torch.int32, device="cuda")
    ntiles = table.device_view().n_padded // 8192
    tt.expand_hits(0, counts.data_ptr(), ntiles, off16.data_ptr(), scratch.data_ptr(), back.data_ptr())
    torch.cuda.synchronize()
    assert (back[: st.n_stored].cpu().numpy().view(np.uint32) == want).all()
    st2, _ = table.sweep_static(tt.MODE_STORE, tt.SWEEP_LIST_ONLY, rows.data_ptr(), ids.data_ptr(), cap)
    assert st2.n_stored == want.size
    res, n_marked, _ = table.sweep(tt.MODE_STORE)
    _, listed, n_over = ot.cron_job(orc.MODE_STORE)
    assert res.count == listed.size == n_marked == n_over
    c, o = ot.flags()
    gc, go = table.flags_bytes()
    live = np.ones(len(ot), dtype=bool)
    live[9] = False  # the deleted row is a tombstone on the device
    assert (gc[live] == c[live]).all() and (go[live] == o[live]).all()
    # documents: written by K6, read back by K7
    all_rows = table.get_due_on_or_before(tt.MODE_FAKE)
    j = all_rows.json(tt.JSON_DOCUMENT)
    off = j.offsets(all_rows.count)
    again = tt.TaskTable(0)
    stats = again.load_documents(j.device_ptr(), off)
    assert stats.n_docs == all_rows.count and stats.first_bad == -1
    assert again.get_due_on_or_before(tt.MODE_FAKE).to_json(tt.JSON_DOCUMENT) == j.bytes()
    j.free()
    again.close()
    table.close()
    print("all_kernels_small ok")


if __name__ == "__main__":
    main()
