"""Host logic of the multi-GPU path on CPU: TaskId-range shard plan and the count / id exchange
over torch.distributed (gloo, world_size 2 and 3), checked against the oracle's global answer."""
import importlib
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("aca-dotnet-workshop_b200")


def test_shard_rows_cover_the_table_exactly():
    for n in (0, 1, 7, 1000, 10**9 + 7):
        for w in (1, 2, 3, 4, 8):
            edges = [pkg.shard_rows(n, w, r) for r in range(w)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(edges[i][1] == edges[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1


def test_taskid_ranges_follow_row_ranges(oracle):
    """rows are in TaskId order, so a row range is a TaskId range"""
    n = 4000
    cfg = oracle.default_cfg(3, n)
    t = oracle.OracleTable()
    t.add_synth(cfg, 0, n)
    ids, _, _ = t.export()
    keys = [bytes(r) for r in ids]
    for w in (2, 4):
        hi_prev = b""
        for r in range(w):
            a, b = pkg.shard_rows(n, w, r)
            assert min(keys[a:b]) > hi_prev
            hi_prev = max(keys[a:b])


def _worker(rank, world, port, n_total, with_ids, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as orc

    cfg = orc.default_cfg(0x5EED0002, n_total, n_creators=100)
    plan = pkg.ShardPlan(n_total, world, rank)
    # each rank sweeps its own shard (here: the oracle stands in for the device sweep)
    shard = orc.OracleTable()
    shard.add_synth(cfg, plan.first_row, plan.n_rows)
    _, listed, _ = shard.cron_job(orc.MODE_STORE)
    local = np.sort(listed)
    ids_all, _, _ = shard.export()
    pad = 17  # buffers are larger than the hit count, like the device buffers
    rows_t = torch.zeros(local.size + pad, dtype=torch.int32)
    rows_t[: local.size] = torch.from_numpy(local.astype(np.int32))
    ids_t = torch.zeros((local.size + pad, 16), dtype=torch.uint8)
    ids_t[: local.size] = torch.from_numpy(ids_all[local])
    counts, glob, gids = pkg.exchange_hits(rows_t, int(local.size), plan, ids=ids_t if with_ids else None)
    # the persistent-buffer form used by bench.py must give the same answer
    ex = pkg.HitExchange(plan, 6000, torch.device("cpu"), with_ids=with_ids)  # same capacity on every rank
    ex.rows_all[: local.size] = rows_t[: local.size]
    if with_ids:
        ex.ids_all[: local.size] = ids_t[: local.size]
    counts2 = ex.run(int(local.size))
    assert counts2 == counts
    assert ex.tune_send_cap(int(local.size)) < 6000
    fixed = ex.run_fixed(int(local.size))
    # compact transport (the packing itself is a device kernel; here the same encoding in numpy)
    ex.enable_compact()
    tc = np.bincount(local // 8192, minlength=ex.n_tiles).astype(np.int32)
    ex.tile_counts_all[: ex.n_tiles] = torch.from_numpy(tc)
    ex.off16_all[: local.size] = torch.from_numpy((local % 8192).astype(np.int16))
    cdev = ex.run_fixed_compact(int(local.size))
    if rank == 0:
        def expand(tile_counts, off16, n_hits):
            t = np.repeat(np.arange(ex.n_tiles), tile_counts.numpy())
            return torch.from_numpy((t * 8192 + off16.numpy().astype(np.int64)[:n_hits]).astype(np.int64))
        assert cdev.tolist() == counts and (ex.global_rows_compact(cdev, expand) == glob).all()
        assert fixed.tolist() == counts and (ex.global_rows_fixed(fixed) == glob).all()
        ex.rows_all[: local.size] = rows_t[: local.size]  # run_fixed left the exact layout untouched only in slot 0
    if rank == 0:
        assert (ex.global_rows(counts2) == glob).all()
        if with_ids:
            assert (ex.ids_all[: sum(counts2)] == gids).all()
        q.put((counts, glob.numpy(), gids.numpy() if gids is not None else None))
    else:
        assert glob is None
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,with_ids", [(2, True), (3, False)])
def test_exchange_over_gloo_matches_global_oracle(oracle, world, with_ids):
    n_total = 30_000
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 500) + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_total, with_ids, q)) for r in range(world)]
    for p in procs:
        p.start()
    counts, glob, gids = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cfg = oracle.default_cfg(0x5EED0002, n_total, n_creators=100)
    whole = oracle.OracleTable()
    whole.add_synth(cfg, 0, n_total)
    _, listed, _ = whole.cron_job(oracle.MODE_STORE)
    want = np.sort(listed)
    assert sum(counts) == want.size and len(counts) == world
    assert (glob == want.astype(np.int64)).all()  # shards are ordered by TaskId range => globally ascending rows
    if with_ids:
        ids_all, _, _ = whole.export()
        assert (gids == ids_all[want]).all()


def test_push_slot_layout_is_aligned_and_disjoint():
    """the slot a shard pushes into: count, per-tile counts and 2-byte offsets never overlap and the
    offsets start 16-byte aligned (tt_sweep_push stores them as 16-byte vectors)"""
    for n_tiles, cap in ((1, 8), (37, 40_000), (122_071, 32_065_536)):
        o_cnt, o_off, total = pkg.push_slot_layout(n_tiles, cap)
        assert o_cnt >= 8 and o_off >= o_cnt + 4 * n_tiles and total >= o_off + 2 * cap
        assert o_cnt % 256 == 0 and o_off % 256 == 0 and total % 256 == 0
