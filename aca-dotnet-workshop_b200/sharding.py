"""TaskId-range sharding and the cross-shard exchange.

The table is sorted by TaskId, so an equal-width TaskId range is a contiguous block of rows:
shard g of G owns global rows [g*N/G, (g+1)*N/G).  Every GPU sweeps its own shard with no
data-path collective; afterwards only the per-shard hit counts (all-gather, 8 bytes per rank)
and the hit identifiers (variable-length gather to the root) cross NVLink.  The exchange is
written against torch.distributed so the same code runs over NCCL on GPUs and over gloo in
the CPU tests.
"""
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class ShardPlan:
    n_total: int
    world: int
    rank: int

    @property
    def first_row(self):
        return shard_rows(self.n_total, self.world, self.rank)[0]

    @property
    def n_rows(self):
        a, b = shard_rows(self.n_total, self.world, self.rank)
        return b - a


def shard_rows(n_total, world, rank):
    """[first, last) global rows of shard `rank`: equal-width split of the TaskId-ordered table."""
    return (n_total * rank) // world, (n_total * (rank + 1)) // world


def exchange_hits(local_rows, count, plan, ids=None, dst=0, group=None):
    """Gathers every shard's hit list on rank `dst`.

    local_rows : int32/uint32 tensor, first `count` entries valid (shard-local row numbers)
    ids        : optional uint8 tensor [*, 16] with the TaskIds of the same hits
    Returns (counts, global_rows, ids_all): counts is a python list on every rank; the other two
    are tensors on `dst` (None elsewhere): global row numbers (int64, ascending — shards are
    ordered by TaskId range) and the concatenated TaskIds.
    """
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = local_rows.device
    mine = torch.tensor([count], dtype=torch.int64, device=dev)
    allc = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(allc, mine, group=group)
    counts = allc.tolist()
    if world == 1:
        rows64 = local_rows[:count].to(torch.int64) & 0xFFFFFFFF
        return counts, rows64 + plan.first_row, (ids[:count] if ids is not None else None)
    offs = [0]
    for c in counts:
        offs.append(offs[-1] + c)
    rows32 = local_rows.view(torch.int32)
    out_rows = out_ids = None
    ops = []
    if rank == dst:
        out_rows = torch.empty(offs[-1], dtype=torch.int32, device=dev)
        out_rows[offs[rank]:offs[rank + 1]] = rows32[:count]
        if ids is not None:
            out_ids = torch.empty((offs[-1], 16), dtype=torch.uint8, device=dev)
            out_ids[offs[rank]:offs[rank + 1]] = ids[:count]
        for r in range(world):
            if r == dst or counts[r] == 0:
                continue
            ops.append(dist.P2POp(dist.irecv, out_rows[offs[r]:offs[r + 1]], r, group))
            if ids is not None:
                ops.append(dist.P2POp(dist.irecv, out_ids[offs[r]:offs[r + 1]], r, group))
    elif count > 0:
        ops.append(dist.P2POp(dist.isend, rows32[:count].contiguous(), dst, group))
        if ids is not None:
            ops.append(dist.P2POp(dist.isend, ids[:count].contiguous(), dst, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    if rank != dst:
        return counts, None, None
    # shard-local -> global row numbers
    glob = out_rows.to(torch.int64) & 0xFFFFFFFF
    for r in range(world):
        glob[offs[r]:offs[r + 1]] += shard_rows(plan.n_total, world, r)[0]
    return counts, glob, out_ids
