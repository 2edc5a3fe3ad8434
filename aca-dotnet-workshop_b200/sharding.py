"""TaskId-range sharding under one process per GPU.

The table is sorted by TaskId, so an equal-width TaskId range is a contiguous block of rows:
shard g of G owns global rows [g*N/G, (g+1)*N/G).  The exchange itself lives behind the C ABI
(include/tasktable.h, tt_exchange_* / tt_shards_*): ShardExchange below only does the rendezvous.
HitExchange / exchange_hits are the earlier NCCL forms (count all-gather + variable-length gather
with grouped send/recv), kept as the fallback for boxes where CUDA IPC is not permitted and
because they run over gloo in the CPU tests.
"""
import os
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


class HitExchange:
    """Persistent buffers for the cross-shard exchange, so the timed path allocates nothing.

    After a shard's sweep, `run` all-gathers the per-shard hit counts (8 bytes per rank) and
    gathers the hit row numbers (uint32, shard-local) and optionally the TaskIds on rank `dst`
    with grouped send/recv at the exclusive-scan offsets of the counts (NCCL has no gather-v).
    On `dst` the local hits are expected to be in `rows_all[:count]` / `ids_all[:count]` already
    (rank dst's offset is 0 when dst == 0), so the sweep can write them in place.
    """

    def __init__(self, plan, cap_per_rank, device, with_ids=False, dst=0, group=None, root_slots=True):
        self.plan, self.cap, self.dst, self.group, self.with_ids = plan, cap_per_rank, dst, group, with_ids
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        # root_slots=False: only this shard's own output buffers (the fused exchange has its own receive buffer)
        n = cap_per_rank * (self.world if self.rank == dst and root_slots else 1)
        self.rows_all = torch.empty(n, dtype=torch.int32, device=device)
        self.ids_all = torch.empty((n, 16), dtype=torch.uint8, device=device) if with_ids else None
        self._mine = torch.zeros(1, dtype=torch.int64, device=device)
        self._counts = torch.zeros(self.world, dtype=torch.int64, device=device)
        self._host = torch.zeros(1, dtype=torch.int64)
        if device.type == "cuda":
            self._host = self._host.pin_memory()
        self._flag = torch.zeros(1, dtype=torch.int32, device=device)
        self.parts = int(os.environ.get("TT_EXCHANGE_PARTS", "1"))
        self.send_cap = cap_per_rank  # entries every shard sends in run_fixed (see tune_send_cap)
        self.peer_rows = self.peer_ids = None
        self.transport = "nccl send/recv"
        # Opt-in only: on the 2 x B200 box measured in round 1 the mapped copy ran at PCIe speed
        # (2.96 ms per 100 MB against 0.63 ms for NCCL send/recv), so send/recv stays the default.
        if device.type == "cuda" and self.world > 1 and os.environ.get("TT_EXCHANGE_PEER", "0") == "1":
            self._map_root_buffers()

    def _map_root_buffers(self):
        """Maps the root's receive buffers into every other process (CUDA IPC through torch's own
        tensor sharing), so a shard can store its hit ids straight into the root's HBM over
        NVLink with one device-to-device copy instead of a NCCL send/recv pair.  Falls back to
        send/recv if the mapping cannot be set up; every rank takes the same decision."""
        from torch.multiprocessing.reductions import reduce_tensor

        ok = 1
        payload = [None]
        try:
            if self.rank == self.dst:
                payload[0] = (reduce_tensor(self.rows_all), reduce_tensor(self.ids_all) if self.ids_all is not None else None)
            dist.broadcast_object_list(payload, src=self.dst, group=self.group)
            if self.rank != self.dst:
                (fn, a), ids = payload[0]
                self.peer_rows = fn(*a)
                if ids is not None:
                    self.peer_ids = ids[0](*ids[1])
                probe = torch.zeros(4, dtype=torch.int32, device=self.rows_all.device)
                self.peer_rows[self.cap * self.rank: self.cap * self.rank + 4].copy_(probe)  # touch the mapping
                torch.cuda.synchronize()
        except Exception:
            ok = 0
        agree = torch.tensor([ok], dtype=torch.int32, device=self.rows_all.device)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=self.group)
        if int(agree.item()) == 1:
            self.transport = "peer copy over NVLink (CUDA IPC mapping of the root's buffers)"
        else:
            self.peer_rows = self.peer_ids = None

    def run(self, count):
        """-> list of per-shard counts (every rank).  On dst, rows_all / ids_all hold all hits."""
        if self.world == 1:
            return [count]
        assert self.dst == 0, "in-place layout assumes the root is rank 0"
        self._host[0] = count
        self._mine.copy_(self._host, non_blocking=True)
        dist.all_gather_into_tensor(self._counts, self._mine, group=self.group)
        counts = self._counts.tolist()
        offs = [0]
        for c in counts:
            offs.append(offs[-1] + c)
        if self.transport.startswith("peer"):
            # every shard stores its ids at its offset in the root's buffers, then a one-word
            # all-reduce (stream-ordered after the copy) tells the root that everything has landed
            if self.rank != self.dst and count > 0:
                self.peer_rows[offs[self.rank]:offs[self.rank] + count].copy_(self.rows_all[:count], non_blocking=True)
                if self.with_ids:
                    self.peer_ids[offs[self.rank]:offs[self.rank] + count].copy_(self.ids_all[:count], non_blocking=True)
            dist.all_reduce(self._flag, group=self.group)
            return counts
        # every payload goes as `parts` send/recv pairs of one grouped launch, so NCCL can spread a
        # large message over several channels
        parts = self.parts

        def pieces(lo, hi):
            step = max(1, -(-(hi - lo) // parts))
            return [(a, min(a + step, hi)) for a in range(lo, hi, step)]

        ops = []
        if self.rank == self.dst:
            for r in range(self.world):
                if r == self.dst or counts[r] == 0:
                    continue
                for a, b in pieces(offs[r], offs[r + 1]):
                    ops.append(dist.P2POp(dist.irecv, self.rows_all[a:b], r, self.group))
                    if self.with_ids:
                        ops.append(dist.P2POp(dist.irecv, self.ids_all[a:b], r, self.group))
        elif count > 0:
            for a, b in pieces(0, count):
                ops.append(dist.P2POp(dist.isend, self.rows_all[a:b], self.dst, self.group))
                if self.with_ids:
                    ops.append(dist.P2POp(dist.isend, self.ids_all[a:b], self.dst, self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return counts

    def tune_send_cap(self, observed_max_count, slack=1.1):
        """Collective.  Shrinks the fixed message of run_fixed from the buffer capacity to
        `slack` x the largest hit count any shard has seen so far (call it after warm-up)."""
        want = torch.tensor([int(observed_max_count * slack) + 1024], dtype=torch.int64, device=self.rows_all.device)
        if self.world > 1:
            dist.all_reduce(want, op=dist.ReduceOp.MAX, group=self.group)
        self.send_cap = min(self.cap, int(want.item()))
        return self.send_cap

    def run_fixed(self, count):
        """Single-phase variant: every shard sends its whole `cap`-sized buffer into the root's slot
        r*cap and its count into counts[r] in ONE grouped send/recv, so no count all-gather and no
        host synchronisation precede the transfer (the price: cap instead of count entries travel).
        Returns the device tensor of counts on the root (None elsewhere); rows of shard r are
        rows_all[r*cap : r*cap + counts[r]].  A shard whose count exceeds send_cap raises."""
        if count > self.send_cap:
            raise RuntimeError(f"{count} hits exceed the agreed message size {self.send_cap}; use run() or tune_send_cap")
        if self.world == 1:
            self._counts[0] = count
            return self._counts
        assert self.dst == 0
        self._host[0] = count
        self._mine.copy_(self._host, non_blocking=True)
        ops = []
        if self.rank == self.dst:
            self._counts[0:1].copy_(self._mine)
            for r in range(1, self.world):
                ops.append(dist.P2POp(dist.irecv, self._counts[r:r + 1], r, self.group))
                ops.append(dist.P2POp(dist.irecv, self.rows_all[r * self.cap:r * self.cap + self.send_cap], r, self.group))
                if self.with_ids:
                    ops.append(dist.P2POp(dist.irecv, self.ids_all[r * self.cap:r * self.cap + self.send_cap], r, self.group))
        else:
            ops.append(dist.P2POp(dist.isend, self._mine, self.dst, self.group))
            ops.append(dist.P2POp(dist.isend, self.rows_all[:self.send_cap], self.dst, self.group))
            if self.with_ids:
                ops.append(dist.P2POp(dist.isend, self.ids_all[:self.send_cap], self.dst, self.group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        return self._counts if self.rank == self.dst else None

    # ---- compact form: 2 bytes per hit (in-tile offsets) + 4 bytes per tile (hit counts) ------------
    def enable_compact(self, tile_rows=8192):
        """Allocates the buffers of run_fixed_compact.  Every rank must call it."""
        rows_max = max(shard_rows(self.plan.n_total, self.world, r)[1] - shard_rows(self.plan.n_total, self.world, r)[0]
                       for r in range(self.world))
        self.tile_rows = tile_rows
        self.n_tiles = max(1, -(-rows_max // tile_rows))
        k = self.world if self.rank == self.dst else 1
        dev = self.rows_all.device
        self.off16_all = torch.zeros(k * self.cap, dtype=torch.int16, device=dev)
        self.tile_counts_all = torch.zeros(k * self.n_tiles, dtype=torch.int32, device=dev)
        return self

    def run_fixed_compact(self, count):
        """Like run_fixed, but what travels is the compact form produced by TaskTable.pack_hits into
        off16_all[:count] / tile_counts_all[:n_tiles]: send_cap 2-byte offsets + n_tiles counts per shard."""
        if count > self.send_cap:
            raise RuntimeError(f"{count} hits exceed the agreed message size {self.send_cap}")
        if self.world == 1:
            self._counts[0] = count
            return self._counts
        self._host[0] = count
        self._mine.copy_(self._host, non_blocking=True)
        ops, nt, cap = [], self.n_tiles, self.cap
        wire = self.off16_all.view(torch.uint8)  # NCCL has no 16-bit integer type: the offsets travel as bytes
        if self.rank == self.dst:
            self._counts[0:1].copy_(self._mine)
            for r in range(1, self.world):
                ops.append(dist.P2POp(dist.irecv, self._counts[r:r + 1], r, self.group))
                ops.append(dist.P2POp(dist.irecv, self.tile_counts_all[r * nt:(r + 1) * nt], r, self.group))
                ops.append(dist.P2POp(dist.irecv, wire[2 * r * cap:2 * (r * cap + self.send_cap)], r, self.group))
        else:
            ops.append(dist.P2POp(dist.isend, self._mine, self.dst, self.group))
            ops.append(dist.P2POp(dist.isend, self.tile_counts_all[:nt], self.dst, self.group))
            ops.append(dist.P2POp(dist.isend, wire[:2 * self.send_cap], self.dst, self.group))
        for w in dist.batch_isend_irecv(ops):
            w.wait()
        return self._counts if self.rank == self.dst else None

    def global_rows_compact(self, counts, expand):
        """dst only: ascending global row numbers from the compact layout.  `expand(tile_counts, off16,
        n_hits)` must return the shard-local rows as an int tensor (tasktable.expand_hits on a GPU)."""
        parts = []
        for r, c in enumerate(counts.tolist()):
            local = expand(self.tile_counts_all[r * self.n_tiles:(r + 1) * self.n_tiles],
                           self.off16_all[r * self.cap:r * self.cap + c], c)
            parts.append((local.to(torch.int64) & 0xFFFFFFFF) + shard_rows(self.plan.n_total, self.world, r)[0])
        return torch.cat(parts) if parts else torch.empty(0, dtype=torch.int64)

    def global_rows_fixed(self, counts):
        """dst only: ascending global row numbers from the fixed-slot layout of run_fixed."""
        cs = counts.tolist()
        parts = []
        for r, c in enumerate(cs):
            seg = self.rows_all[r * self.cap: r * self.cap + c].to(torch.int64) & 0xFFFFFFFF
            parts.append(seg + shard_rows(self.plan.n_total, self.world, r)[0])
        return torch.cat(parts) if parts else torch.empty(0, dtype=torch.int64)

    def global_rows(self, counts):
        """dst only: global row numbers (int64, ascending) from the gathered shard-local ones."""
        total = sum(counts)
        glob = self.rows_all[:total].to(torch.int64) & 0xFFFFFFFF
        off = 0
        for r, c in enumerate(counts):
            glob[off:off + c] += shard_rows(self.plan.n_total, self.world, r)[0]
            off += c
        return glob


def push_slot_layout(n_tiles, cap):
    """Byte layout of one shard's slot in the root's receive buffer for the fused exchange:
    [hit count, u64, in its own 256 B][per-tile counts, u32 x n_tiles][in-tile offsets, u16 x cap],
    every part 256-byte aligned.  -> (off_tile_count, off_off16, slot_bytes)."""
    up = lambda x: (x + 255) // 256 * 256
    o_cnt = 256
    o_off = o_cnt + up(4 * n_tiles)
    return o_cnt, o_off, o_off + up(2 * cap)


class ShardExchange:
    """One rank's end of the fused cross-shard exchange: tasktable.Exchange (C ABI: tt_exchange_*) plus the
    rendezvous it needs under one-process-per-GPU — the root's 64-byte buffer handle is broadcast, every rank
    learns the shards' first rows and the largest tile count.  That is ALL torch.distributed does here: a
    tick stores its result into the root's slot from inside the placement kernel (over NVLink when the
    slot is on another GPU), publishes an epoch word, and the root waits for the words on the device; no
    collective runs in the tick.  Collective constructor: every rank must create it."""

    def __init__(self, first_row, n_rows, cap_per_shard, device_index, flags=0, group=None, tile_rows=8192):
        from . import tasktable as tt

        self._tt, self.group, self.device_index = tt, group, device_index
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        dev = torch.device("cuda", device_index)
        mine = torch.tensor([first_row, -(-n_rows // tile_rows), (cap_per_shard + 7) // 8 * 8], dtype=torch.int64, device=dev)
        every = [torch.zeros(3, dtype=torch.int64, device=dev) for _ in range(self.world)]
        if self.world > 1:
            dist.all_gather(every, mine, group=group)
        else:
            every = [mine]
        self.first_rows = [int(x[0].item()) for x in every]
        self.max_tiles = max(1, max(int(x[1].item()) for x in every))
        self.cap = max(int(x[2].item()) for x in every)  # shards need not be the same size: the slots fit the largest
        payload = [None]
        self.ex = None
        if self.rank == 0:
            try:
                self.ex, payload[0] = tt.Exchange.create_root(device_index, self.world, self.cap, self.max_tiles, flags)
            except tt.TaskTableError as e:  # the other ranks are waiting in the broadcast: tell them
                payload[0] = str(e)
        if self.world > 1:
            dist.broadcast_object_list(payload, src=0, group=group)
        if not isinstance(payload[0], bytes):
            raise RuntimeError(f"root could not allocate the receive buffer: {payload[0]}")
        if self.rank != 0:
            self.ex = tt.Exchange.attach(device_index, self.rank, self.world, self.cap, self.max_tiles, flags, payload[0])

    def tick(self, table, mode, flags, **kw):
        """This shard's cron tick (Exchange.tick).  -> (SweepStats, rc)."""
        return self.ex.tick(table, mode, flags, **kw)

    def collect(self, order=0):
        """root: every shard's list as one GResult (waits on the device for the shards' epoch words)."""
        assert self.rank == 0
        return self.ex.collect(order, self.first_rows)

    def release(self):
        assert self.rank == 0
        self.ex.release()

    def close(self):
        if self.ex is None:
            return
        if self.world > 1:
            dist.barrier(group=self.group)  # nobody may still be storing into the buffer
        if self.rank != 0:
            self.ex.close()
        if self.world > 1:
            dist.barrier(group=self.group)  # the mappings are gone before the root frees the memory
        if self.rank == 0:
            self.ex.close()
        self.ex = None


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
