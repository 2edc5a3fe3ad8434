"""Run under torchrun: every rank sweeps its TaskId-range shard on its own GPU, the hit counts
and hit ids are exchanged over NCCL, and rank 0 checks the global answer against the streaming
oracle (count, XOR of TaskIds, sum of global rows, ascending order)."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    n_total = int(sys.argv[1]) if len(sys.argv) > 1 else 64_000_000
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    plan = pkg.ShardPlan(n_total, world, rank)
    cfg = pkg.default_cfg(0x5EED0002, n_total)
    table = pkg.TaskTable(local).generate(cfg, plan.first_row, plan.n_rows)
    cap = int(plan.n_rows * 0.032) + 65536
    d_rows = torch.empty(cap, dtype=torch.int32, device=dev)
    d_ids = torch.empty((cap, 16), dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    st, _ = table.sweep_device(tt.MODE_STORE, tt.SWEEP_MARK, d_rows.data_ptr(), d_ids.data_ptr(), cap,
                               stream=stream.cuda_stream)
    counts, glob, gids = pkg.exchange_hits(d_rows, st.n_stored, plan, ids=d_ids)
    # the persistent-buffer exchange bench.py uses (peer copy when the mapping works) must agree
    ex = pkg.HitExchange(plan, cap, dev, with_ids=True)
    ex.rows_all[: st.n_stored] = d_rows[: st.n_stored]
    ex.ids_all[: st.n_stored] = d_ids[: st.n_stored]
    counts2 = ex.run(st.n_stored)
    torch.cuda.synchronize()
    same = counts2 == counts
    if rank == 0:
        same = same and bool((ex.global_rows(counts2) == glob).all()) and bool((ex.ids_all[: sum(counts2)] == gids).all())
        print("exchange transport:", ex.transport, "agrees" if same else "DIFFERS")
    assert same
    # the compact exchange (2-byte offsets + per-tile counts), expanded again on rank 0
    ex.enable_compact()
    ex.tune_send_cap(st.n_stored)
    table.pack_hits(d_rows.data_ptr(), st.n_stored, ex.off16_all.data_ptr(), ex.tile_counts_all.data_ptr(),
                    stream=stream.cuda_stream)
    cdev = ex.run_fixed_compact(st.n_stored)
    torch.cuda.synchronize()
    if rank == 0:
        scratch = torch.empty(ex.n_tiles + ex.n_tiles // 1024 + 8, dtype=torch.int32, device=dev)

        def expand(tile_counts, off16, n_hits):
            out = torch.empty(max(n_hits, 1), dtype=torch.int32, device=dev)
            tt.expand_hits(local, tile_counts.data_ptr(), ex.n_tiles, off16.data_ptr(), scratch.data_ptr(), out.data_ptr(),
                           stream.cuda_stream)
            torch.cuda.synchronize()
            return out[:n_hits]

        g2 = ex.global_rows_compact(cdev, expand)
        ok2 = cdev.tolist() == counts and bool((g2 == glob).all())
        print("compact exchange", "agrees" if ok2 else "DIFFERS")
        assert ok2
    # the fused exchange behind the C ABI (tt_exchange_*): a fresh copy of the shard, ticked straight into rank 0's
    # buffer with TaskIds and creation times; rank 0 waits for the epoch words on the device
    table_p = pkg.TaskTable(local).generate(cfg, plan.first_row, plan.n_rows)
    sx = pkg.ShardExchange(plan.first_row, plan.n_rows, cap, local, tt.EXCHANGE_IDS | tt.EXCHANGE_CREATED)
    p_rows = torch.empty(sx.cap, dtype=torch.int32, device=dev)
    p_ids = torch.empty((sx.cap, 16), dtype=torch.uint8, device=dev)
    for tick in range(3):  # the later ticks find nothing: the slots must be overwritten, not appended to
        dist.barrier()
        stp, _ = sx.tick(table_p, tt.MODE_STORE, tt.SWEEP_MARK | tt.SWEEP_ASYNC, stream=stream.cuda_stream,
                         d_rows=p_rows.data_ptr(), d_ids=p_ids.data_ptr())
        ok3 = True
        if rank == 0:
            res = sx.collect(tt.ORDER_ROWS if tick != 1 else tt.ORDER_CREATED_ASC)
            if tick == 0:
                ok3 = (res.shard_counts()[:world] == counts and bool((torch.from_numpy(res.rows().astype(np.int64)) == glob.cpu()).all())
                       and bool((torch.from_numpy(res.ids()) == gids.cpu()).all()))
            else:
                ok3 = res.count == 0
            res.free()
            sx.release()
        stp, _ = table_p.sweep_wait()
        if tick == 0:
            ok3 = ok3 and (stp.n_stored == st.n_stored and bool((p_rows[: stp.n_stored] == d_rows[: st.n_stored]).all())
                           and bool((p_ids[: stp.n_stored] == d_ids[: st.n_stored]).all()))
        else:
            ok3 = ok3 and stp.n_listed == 0
        if rank == 0:
            print("fused exchange (tt_exchange_*) tick", tick, "agrees" if ok3 else "DIFFERS")
        assert ok3
    assert table_p.count_flags() == table.count_flags()
    # the cross-shard OrderBy(TaskCreatedOn) on an unmarked copy: same hits, creation times ascending
    table_q = pkg.TaskTable(local).generate(cfg, plan.first_row, plan.n_rows)
    dist.barrier()
    sx.tick(table_q, tt.MODE_STORE, tt.SWEEP_LIST_ONLY)
    if rank == 0:
        res = sx.collect(tt.ORDER_CREATED_ASC)
        rows_o = res.rows().astype(np.int64)
        okq = res.count == sum(counts) and bool((np.sort(rows_o) == glob.cpu().numpy()).all())
        idmap = {int(r): i for i, r in enumerate(glob.cpu().numpy())}
        okq = okq and bool((res.ids() == gids.cpu().numpy()[[idmap[int(r)] for r in rows_o]]).all())
        print("cross-shard order", "agrees" if okq else "DIFFERS")
        assert okq
        res.free()
        sx.release()
    sx.close()
    table_q.close()
    table_p.close()
    ok = True
    if rank == 0:
        from oracle import oracle as orc

        want = orc.stream_sweep(orc.default_cfg(0x5EED0002, n_total), 0, n_total, orc.MODE_STORE)
        g = glob.cpu().numpy()
        ids = gids.cpu().numpy()
        hi = int(np.bitwise_xor.reduce(np.ascontiguousarray(ids[:, :8]).view(">u8").ravel().astype(np.uint64)))
        lo = int(np.bitwise_xor.reduce(np.ascontiguousarray(ids[:, 8:]).view(">u8").ravel().astype(np.uint64)))
        ok = (sum(counts) == want.n_listed == g.size and (g[:-1] < g[1:]).all()
              and int(g.astype(np.uint64).sum() & np.uint64(2**64 - 1)) == want.sum_rows
              and (hi, lo) == (want.xor_id_hi, want.xor_id_lo))
        print("multigpu_check", "OK" if ok else "MISMATCH", "world", world, "counts", counts, "listed", want.n_listed)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, 0)
    # per-shard flag planes against the oracle's popcount for that shard
    from oracle import oracle as orc

    mine = orc.stream_sweep(orc.default_cfg(0x5EED0002, n_total), plan.first_row, plan.n_rows, orc.MODE_STORE)
    good = table.count_flags() == (mine.completed, mine.overdue_after) and st.n_listed == mine.n_listed
    allgood = torch.tensor([1 if good else 0], device=dev)
    dist.all_reduce(allgood, op=dist.ReduceOp.MIN)
    table.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 and allgood.item() == 1 else 1)


if __name__ == "__main__":
    main()
