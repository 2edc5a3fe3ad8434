#!/usr/bin/env python
"""bench.py — overdue-sweep throughput (tasks/s) on N B200s, with roofline, e2e and CPU baseline.

One "step" is one cron tick (GetYesterdaysDueTasks -> runAt.Date > due.Date -> MarkOverdueTasks,
STORE semantics) over the whole task table.  `value` is timed with the table resident in HBM
(tt_sweep_device); `e2e` goes through the reference-facing call tt_sweep with host buffers on
both sides.  With N > 1 every rank owns one TaskId-range shard, sweeps it with no data-path
collective, then the hit counts are all-gathered and the hit row ids gathered on rank 0 (NCCL).

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference        # the reference's CPU path (oracle port), host cores
"""
import argparse
import importlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "overdue_sweep_tasks_per_sec"
UNIT = "tasks/s"
SEED = 0x5EED0002
BYTES_PER_TASK = 4.25  # u32 due_code + completed bit + overdue bit      
BYTES_PER_HIT = 36.0   # u32 row out + 16 B TaskId in + 16 B TaskId out  


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=1_000_000_000, help="rows per GPU (weak) or in total (strong)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--exchange", default="rows", choices=["rows", "ids"],
                    help="what the shards' placement kernels store into rank 0's buffer over NVLink (N > 1, tt_exchange_*): "
                         "rows = the compact row lists (2-byte in-tile offsets + per-tile counts), each shard keeping its "
                         "TaskIds; ids = the 16-byte TaskIds as well")
    ap.add_argument("--root-share", type=float, default=0.0,
                    help="N > 1: rank 0's shard as a fraction of an equal share (rank 0 also collects every shard's list, "
                         "so its TaskId range can be made smaller while the other ranks split the rest equally); 0 = 1.0, equal "
                         "ranges: measured on 8 B200s, a smaller root shard does not shorten the tick (1.54 / 1.58 / 1.59 / "
                         "1.63 ms at 1.0 / 0.8 / 0.7 / 0.6) — the peers' placement kernels, held to the root's NVLink "
                         "intake, are the critical path, and rank 0 only waits longer for them")
    ap.add_argument("--no-extra", action="store_true", help="N > 1: skip the strong-scaling / mixed / ids-at-root passes")
    ap.add_argument("--cpu-sample-rows", type=int, default=10_000_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=0, help="row ranges tt_tick_host pipelines over (0 = the library's 8)")
    ap.add_argument("--workload", default="sweep", choices=["sweep", "sweep10m", "creator", "creator-threads", "mixed", "json", "ingest"],
                    help="sweep = the headline (BASELINE.json metric); the others are the remaining configs of "
                         "SURVEY.md 8(d): sweep10m = config 2, creator = config 4, mixed = config 5 (per GPU)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons while the timed region runs (NVML, 2 ms period)."""

    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h) if hasattr(
                    nv, "nvmlDeviceGetCurrentClocksEventReasons") else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for bit, name in self.REASONS.items():
                    if mask & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self._nv:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._thread:
            self._thread.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# --------------------------------------------------------------------------------------------------
def cpu_reference_run(sample_rows, steps, warmup):
    """The reference's CPU path, restated (oracle/oracle.cpp): List<TaskModel> of heap objects,
    sequential scan on one thread like LINQ-to-objects, STORE semantics, whole cron tick."""
    from oracle import oracle as orc

    cfg = orc.default_cfg(SEED, 1_000_000_000)
    ot = orc.OracleTable()
    ot.add_synth(cfg, 0, sample_rows)
    ot.build_keys()  # the state store's key index exists before the tick
    c0, o0 = ot.flags()
    total, hits = 0.0, 0
    for i in range(warmup + steps):
        ot.set_all_flags(c0, o0)
        t = time.perf_counter()
        _, listed, _ = ot.cron_job(orc.MODE_STORE)
        dt = time.perf_counter() - t
        if i >= warmup:
            total += dt
            hits = listed.size
    value = sample_rows * steps / total
    # context: the same tick over SoA columns with every host core (not the reference's layout)
    cores = os.cpu_count() or 1
    soa = orc.SoaTable(cfg, 0, min(4 * sample_rows, 64_000_000), threads=cores)
    best = float("inf")
    for _ in range(3):
        soa.reset()
        t = time.perf_counter()
        soa.sweep()
        best = min(best, time.perf_counter() - t)
    par = soa.n / best
    soa.close()
    ot.close()
    return {"value": value, "unit": UNIT, "cores": 1, "kind": "port",
            "sample": f"{sample_rows} synthetic tasks (first rows of the 1B table), AoS List<TaskModel>, "
                      f"STORE-mode cron tick x{steps}, {hits} hits/tick; 1 thread because LINQ-to-objects is sequential",
            "ms_per_step": 1e3 * total / steps,
            "parallel_soa": {"value": par, "unit": UNIT, "cores": cores,
                             "note": "best-effort SoA scan on all host cores, context only"}}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 10))
    base = cpu_reference_run(args.cpu_sample_rows, steps, min(args.warmup, 1))
    line = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": min(args.warmup, 1), "ms_per_step": base["ms_per_step"], "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": {"workload": "overdue sweep (cron tick), STORE semantics, config-5 mix; CPU sample of the 1B-task table",
                       "sample_rows": args.cpu_sample_rows},
            "cpu_baseline": {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "parallel_soa")},
            "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "dotnet is unavailable on this box: the reference's C# LINQ path is timed as its C++ restatement"}
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------
class ShardRun:
    """One rank's shard of a synthetic table plus its end of the cross-shard exchange (C ABI: tt_exchange_*).
    torch.distributed is used for the rendezvous only (the 64-byte handle of the root's buffer, barriers
    between steps, max-over-ranks of the timings): the tick itself contains no collective."""

    def __init__(self, pkg, torch, dist, dev, local, rank, world, n_total, first, n_rows, ex_flags, local_ids=True):
        tt = pkg.tasktable
        self.tt, self.torch, self.dist, self.rank, self.world, self.n_total, self.n_rows = tt, torch, dist, rank, world, n_total, n_rows
        self.table = pkg.TaskTable(local).generate(pkg.default_cfg(SEED, n_total), first, n_rows)
        self.nbytes = (n_rows + 7) // 8
        self.pristine = torch.empty(self.nbytes, dtype=torch.uint8, device=dev)
        tt._check(tt.lib().tt_snapshot_flags(self.table._h, None, self.pristine.data_ptr()))
        self.cap = (int(n_rows * 0.032) + 65536 + 7) // 8 * 8
        self.d_rows = torch.empty(self.cap, dtype=torch.int32, device=dev)
        self.d_ids = torch.empty((self.cap, 16), dtype=torch.uint8, device=dev) if local_ids else None
        self.stream = torch.cuda.Stream(device=dev)  # non-default: its handle is what the C ABI launches on
        self.sx = pkg.ShardExchange(first, n_rows, self.cap, local, ex_flags) if world > 1 else None
        self.ex = self.sx.ex if self.sx else None
        self.first_rows = self.sx.first_rows if self.sx else [first]

    def reset(self):
        self.tt._check(self.tt.lib().tt_restore_flags(self.table._h, None, self.pristine.data_ptr()))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def step(self, extra_flags):
        """Queues this rank's tick; on the root also collects every shard's list.  -> collect ms (root) or None."""
        tt = self.tt
        flags = tt.SWEEP_MARK | tt.SWEEP_ASYNC | extra_flags
        ids = self.d_ids.data_ptr() if self.d_ids is not None else None
        if self.ex is None:
            self.table.sweep_device(tt.MODE_STORE, flags, self.d_rows.data_ptr(), ids, self.cap, stream=self.stream.cuda_stream)
            return None
        # every shard keeps its own row list (and TaskIds); the root's placement kernel writes its rows straight
        # into the end-to-end list the collect builds, so it needs no second copy
        own_rows = None if self.rank == 0 else self.d_rows.data_ptr()
        self.ex.tick(self.table, tt.MODE_STORE, flags, stream=self.stream.cuda_stream, d_rows=own_rows, d_ids=ids)
        if self.rank == 0:
            res = self.ex.collect(tt.ORDER_ROWS | tt.COLLECT_RELEASE, self.first_rows)
            self.collected = res.count
            res.free()
            return self.ex.last_collect_ms()
        return None

    def timed(self, steps, warmup, clocks_for=None):
        """-> dict(ms_per_step (max over ranks), hits (all shards), launches, kernel/sweep/gather medians of an
        instrumented second pass, collect_ms on the root)."""
        torch, tt = self.torch, self.tt
        for _ in range(warmup):
            self.reset()
            self.barrier()
            self.step(tt.SWEEP_NO_TIMING)
            self.table.sweep_wait()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        hits = launches = 0
        collect_ms = []
        self.barrier()

        def loop():
            nonlocal hits, launches
            for i in range(steps):
                self.reset()     # untimed: the overdue plane goes back to its pre-tick state
                self.barrier()
                ev[i][0].record(self.stream)
                c = self.step(tt.SWEEP_NO_TIMING)
                ev[i][1].record(self.stream)
                st, _ = self.table.sweep_wait()
                hits = st.n_listed
                launches += st.launches + (2 if c is not None else 0)
                if c is not None:
                    collect_ms.append(c)
            self.barrier()

        if clocks_for is not None:
            with clocks_for:
                loop()
        else:
            loop()
        step_ms = [a.elapsed_time(b) for a, b in ev]
        # per-kernel breakdown: a second, instrumented pass (the library records CUDA events between its
        # kernels on the stream they run on); the headline steps above carry none
        kms, sms, gms = [], [], []
        for i in range(steps):
            self.reset()
            self.barrier()
            self.step(0)
            st, _ = self.table.sweep_wait()
            kms.append(st.kernel_ms)
            sms.append(st.sweep_ms)
            gms.append(st.gather_ms)
        self.barrier()
        dev = self.pristine.device
        total = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
        hits_all = torch.tensor([hits], dtype=torch.int64, device=dev)
        if self.world > 1:
            self.dist.all_reduce(total, op=self.dist.ReduceOp.MAX)
            self.dist.all_reduce(hits_all, op=self.dist.ReduceOp.SUM)
        med = lambda xs: sorted(xs)[len(xs) // 2] if xs else None
        return {"total_ms": float(total.item()), "ms_per_step": float(total.item()) / steps, "hits": int(hits_all.item()),
                "hits_rank0": hits, "launches": launches, "kernel_ms": med(kms), "sweep_ms": med(sms), "gather_ms": med(gms),
                "collect_ms": med(collect_ms), "step_ms_rank0": med(step_ms),
                "collected": getattr(self, "collected", None)}

    def close(self):
        if self.sx is not None:
            self.torch.cuda.synchronize()
            self.sx.close()
        self.table.close()


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    pkg = importlib.import_module("aca-dotnet-workshop_b200")
    tt = pkg.tasktable

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    share = args.root_share if args.root_share > 0 else 1.0

    def shard_of(total, r):
        """[first, last) rows of rank r: TaskId-range shards; rank 0's range is `share` of an equal one (whole tiles)."""
        if world == 1:
            return 0, total
        root = int(total / world * share) // 8192 * 8192
        rest = total - root
        cut = lambda k: root + rest * k // (world - 1)   # k = 0 .. world - 1 over the other ranks
        return (0, root) if r == 0 else (cut(r - 1), cut(r))

    n_total = args.rows * world if args.scaling == "weak" else args.rows
    first, last = shard_of(n_total, rank)
    n_rows = last - first
    with_root_ids = args.exchange == "ids"
    run = ShardRun(pkg, torch, dist, dev, local, rank, world, n_total, first, n_rows,
                   tt.EXCHANGE_IDS if with_root_ids else 0, local_ids=not with_root_ids)
    clocks = ClockSampler(local)
    head = run.timed(args.steps, args.warmup, clocks_for=clocks)
    value = n_total * args.steps / (head["total_ms"] * 1e-3)
    table, cap, nbytes, hits = run.table, run.cap, run.nbytes, head["hits_rank0"]

    def barrier():
        run.barrier()

    # ---- e2e: the reference-facing call with host buffers on both sides ----
    # tt_tick_host: both flag planes come from pinned host memory, the listed rows + TaskIds go back to
    # pinned host memory; the library pipelines upload / kernels / download over chunks of the table.
    # Beside it, the same work as separate calls (tt_restore_flags, tt_sweep, tt_result_rows / _ids).
    e2e = None
    if not args.no_e2e:
        h_comp = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        h_over = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        tt._check(tt.lib().tt_snapshot_flags(table._h, h_comp.data_ptr(), None))
        h_over.copy_(run.pristine)
        h_rows = torch.empty(cap, dtype=torch.int32).pin_memory()
        h_ids = torch.empty((cap, 16), dtype=torch.uint8).pin_memory()
        e_steps = max(1, min(args.steps, 10))

        def pipelined():
            st, _ = table.tick_host(tt.MODE_STORE, tt.SWEEP_MARK, h_comp.data_ptr(), h_over.data_ptr(), h_rows.data_ptr(),
                                    h_ids.data_ptr(), cap, args.e2e_chunks)
            return st.n_stored

        def sequential():
            tt._check(tt.lib().tt_restore_flags(table._h, h_comp.data_ptr(), h_over.data_ptr()))
            res, n_marked, _ = table.sweep(tt.MODE_STORE)
            tt._check(tt.lib().tt_result_rows(res._h, h_rows.data_ptr()))
            tt._check(tt.lib().tt_result_ids(res._h, h_ids.data_ptr()))
            cnt = res.count
            res.free()
            return cnt

        def timed(fn):
            total, cnt = 0.0, 0
            for i in range(2 + e_steps):
                barrier()
                t0 = time.perf_counter()
                cnt = fn()
                barrier()
                if i >= 2:
                    total += time.perf_counter() - t0
            et = torch.tensor([total], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(et, op=dist.ReduceOp.MAX)
            return n_total * e_steps / float(et.item()), cnt

        v_pipe, cnt = timed(pipelined)
        assert cnt == hits, (cnt, hits)
        v_seq, cnt2 = timed(sequential)
        assert cnt2 == hits, (cnt2, hits)
        e2e = {"value": v_pipe, "unit": UNIT, "h2d_bytes_per_step": 2 * nbytes, "d2h_bytes_per_step": cnt * 20,
               "steps": e_steps, "chunks": args.e2e_chunks or 8,
               "what": "tt_tick_host per rank: pinned host flag planes -> device, the tick, hit rows + TaskIds -> pinned host, "
                       "the three stages pipelined over chunks of the table on three streams",
               "sequential": {"value": v_seq, "what": "the same bytes as separate calls: tt_restore_flags + tt_sweep + "
                                                      "tt_result_rows + tt_result_ids"}}

    # ---- N > 1: the other driver-visible numbers, measured in the same run after the headline ----
    extra = {}
    if world > 1 and not args.no_extra:
        run.close()
        x_steps, x_warm = max(3, min(args.steps, 10)), max(3, min(args.warmup, 3))
        if args.scaling == "weak" and not with_root_ids:
            # the same weak-scaled tick with the 16-byte TaskIds ALSO delivered into the root's slots (fused into K3g)
            r2 = ShardRun(pkg, torch, dist, dev, local, rank, world, n_total, first, n_rows, tt.EXCHANGE_IDS, local_ids=False)
            m = r2.timed(x_steps, x_warm)
            extra["ids_at_root_ms_per_step"] = m["ms_per_step"]
            extra["ids_at_root_tasks_per_sec"] = n_total * x_steps / (m["total_ms"] * 1e-3)
            r2.close()
        # BASELINE.json config 3: ONE table of args.rows tasks sharded by TaskId range over the N GPUs
        s_first, s_last = shard_of(args.rows, rank)
        r3 = ShardRun(pkg, torch, dist, dev, local, rank, world, args.rows, s_first, s_last - s_first, 0)
        m = r3.timed(x_steps, x_warm)
        extra["strong_ms_per_step"] = m["ms_per_step"]
        extra["strong_tasks_per_sec"] = args.rows * x_steps / (m["total_ms"] * 1e-3)
        extra["strong_n_total"] = args.rows
        extra["strong_hits"] = m["hits"]
        extra["strong_collect_ms"] = m["collect_ms"]
        # BASELINE.json config 5: the tick, then 64 creator filters as one batch on every shard and the per-creator
        # match counts summed over the shards (the one step of this path with a real reduction: NCCL all-reduce)
        rng = np.random.default_rng(5)
        names = [("user%07d@mail%d.example" % (k, k % 97)).encode() for k in rng.integers(0, 1_000_000, size=64)]
        counts_dev = torch.zeros(64, dtype=torch.int64, device=dev)
        counts_host = torch.zeros(64, dtype=torch.int64).pin_memory()
        walls = []
        for i in range(x_warm + x_steps):
            r3.reset()
            r3.barrier()
            t0 = time.perf_counter()
            r3.step(tt.SWEEP_NO_TIMING)
            r3.table.sweep_wait()
            rs = r3.table.get_by_creators(names)
            for k, r in enumerate(rs):
                counts_host[k] = r.count
            counts_dev.copy_(counts_host, non_blocking=True)
            dist.all_reduce(counts_dev)
            r3.barrier()
            dt = time.perf_counter() - t0
            for r in rs:
                r.free()
            if i >= x_warm:
                walls.append(dt * 1e3)
        w = torch.tensor([sorted(walls)[len(walls) // 2]], dtype=torch.float64, device=dev)
        dist.all_reduce(w, op=dist.ReduceOp.MAX)
        extra["mixed_ms"] = float(w.item())
        extra["mixed_creator_matches"] = int(counts_dev.sum().item())
        r3.close()
        run = None

    if rank == 0:
        peak, peak_src = peaks()
        kms, sms, gms = head["kernel_ms"], head["sweep_ms"], head["gather_ms"]
        # Dominant kernel = K1 sweep_kernel: it reads 4.25 B/task and writes one u32 row per hit.
        # K3g (row order + TaskId gather) reads 4 B + 16 B and writes 4 B + 16 B per hit.
        # The whole step is SURVEY.md 8(d)'s figure: 4.25 B/task + 36 B/hit (+ the 8 B/hit of the
        # intermediate list, which is not counted).
        sweep_bytes = n_rows * BYTES_PER_TASK + hits * 4.0
        place_bytes = hits * 40.0
        step_bytes = n_rows * BYTES_PER_TASK + hits * BYTES_PER_HIT
        achieved = sweep_bytes / (sms * 1e-3) / 1e9
        step_gbs = step_bytes / (kms * 1e-3) / 1e9
        place_gbs = place_bytes / (gms * 1e-3) / 1e9 if gms > 0 else None
        ms_per_step = head["ms_per_step"]
        step_frac = step_bytes / (ms_per_step * 1e-3) / 1e9 / peak  # from the timed step, not from the kernel sum
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "sweep_traffic.json")
        if os.path.exists(tpath):
            try:
                tj = json.load(open(tpath))
                if tj.get("rows") == n_rows:
                    traffic = tj.get("dram_bytes_per_launch")
            except Exception:
                pass
        exchange = "none" if world == 1 else ("fused push of compact rows + TaskIds" if with_root_ids else "fused push of compact rows")
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {
                "workload": f"overdue sweep (cron tick, STORE semantics) over a synthetic {n_total}-task table, "
                            "config-5 mix: 5% due yesterday, 1% adversarial inexact, 50% completed",
                "rows_per_gpu": n_total // world, "rows_rank0": n_rows, "root_share": share if world > 1 else None,
                "n_total": n_total, "hits_per_step": head["hits"],
                "sharding": ("TaskId ranges; rank 0, which also collects every shard's list, owns root_share of an equal "
                             "range, the other ranks split the rest equally") if world > 1 else "single shard",
                "exchange": exchange,
                "l2": f"inputs {n_rows * BYTES_PER_TASK / 1e6:.0f} MB per step vs 126 MB L2; overdue plane rewritten between steps",
                "reset": "overdue plane restored between steps outside the timed region (per-step CUDA events summed)",
                "timed_region": "per step, CUDA events on the launching stream around the queued tick (tt_sweep_device / "
                                "tt_exchange_tick with TT_SWEEP_ASYNC) and, on rank 0 at N > 1, the collect (device-side wait "
                                "for every shard's epoch word + rebuilding the row numbers); max over ranks",
            },
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "step_frac": step_frac, "step_achieved": step_frac * peak,
                         "traffic": traffic,
                         "traffic_source": "profiles/sweep_traffic.json: dram__bytes_read + dram__bytes_write of one ncu --set full "
                                           "capture of this kernel at this size (not re-measured in this run)",
                         "kernel": "tt::sweep_kernel<true,true,256,3,false> (K1, dominant)",
                         "kernel_ms": sms, "algorithmic_bytes": sweep_bytes,
                         "bytes_per_unit": "4.25 B/task + 4 B/hit", "peak_source": peak_src,
                         "step": {"kernels": "K1 sweep_kernel + K3g place_group_kernel + sweep_tail_kernel", "ms": kms,
                                  "algorithmic_bytes": step_bytes, "bytes_per_unit": "4.25 B/task + 36 B/hit",
                                  "achieved": step_gbs, "frac": step_gbs / peak,
                                  "note": "frac here divides by the kernels' own time (instrumented pass); step_frac above "
                                          "divides by ms_per_step"},
                         "place_kernels": {"ms": gms, "achieved": place_gbs,
                                           "frac": (place_gbs / peak) if place_gbs else None,
                                           "note": "random 16 B TaskId gathers cost 64 B of DRAM and one row activation each"}},
            "gpu_launches": head["launches"],
            "exchange": {"what": exchange,
                         "transport": (None if world == 1 else "stores from the placement kernel into rank 0's HBM over NVLink "
                                       "(CUDA IPC mapping), epoch word per shard, device-side wait on rank 0: no collective"),
                         "collect_ms": head["collect_ms"], "rows_at_root": head["collected"],
                         "ms": head["step_ms_rank0"] - kms,
                         "note": "ms = rank 0's step time outside its own three kernels: launch gaps plus, at N > 1, waiting "
                                 "for the slowest shard and rebuilding the row numbers (collect_ms, CUDA events on the root's "
                                 "collect stream)"},
            "clocks": clocks.summary(),
        }
        line.update(extra)
        if e2e:
            line["e2e"] = e2e
        if world == 1 and not args.no_cpu_baseline:
            base = cpu_reference_run(args.cpu_sample_rows, 3, 1)
            line["cpu_baseline"] = {k: base[k] for k in ("value", "unit", "cores", "kind", "sample", "parallel_soa")}
        print(json.dumps(line))
    if run is not None:
        run.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------------
def run_extra(args):
    """The other configs of SURVEY.md 8(d) on one GPU (not the driver's headline line)."""
    import numpy as np
    import torch

    pkg = importlib.import_module("aca-dotnet-workshop_b200")
    tt = pkg.tasktable
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    peak, peak_src = peaks()
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    med = lambda xs: sorted(xs)[len(xs) // 2]

    if args.workload == "sweep10m":
        flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
        n = 10_000_000
        cfg = pkg.default_cfg(0x5EED0001, n)
        table = pkg.TaskTable(0).generate(cfg, 0, n)
        nbytes = (n + 7) // 8
        pristine = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        tt._check(tt.lib().tt_snapshot_flags(table._h, None, pristine.data_ptr()))
        cap = int(n * 0.032) + 65536
        d_rows = torch.empty(cap, dtype=torch.int32, device=dev)
        d_ids = torch.empty((cap, 16), dtype=torch.uint8, device=dev)
        wall, kms, sms = [], [], []
        # two passes: the call as a caller makes it (TT_SWEEP_NO_TIMING: no events between the kernels) on the wall
        # clock, then the same call with the library's events for the kernel times
        for timed in (False, True):
            for i in range(args.warmup + args.steps):
                tt._check(tt.lib().tt_restore_flags(table._h, None, pristine.data_ptr()))
                flush.zero_()  # L2 flush: the 42.5 MB of hot columns would otherwise stay in the 126 MB L2
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                st, _ = table.sweep_device(tt.MODE_STORE, tt.SWEEP_MARK | (0 if timed else tt.SWEEP_NO_TIMING),
                                           d_rows.data_ptr(), d_ids.data_ptr(), cap, stream=stream.cuda_stream)
                dt = time.perf_counter() - t0
                if i >= args.warmup and not timed:
                    wall.append(dt * 1e3)
                if i >= args.warmup and timed:
                    kms.append(st.kernel_ms)
                    sms.append(st.sweep_ms)
        hits = st.n_listed
        line = {"metric": METRIC, "workload": "config 2: 10M-task sweep, 1 GPU, L2 flushed between steps",
                "value": n / (med(wall) * 1e-3), "unit": UNIT, "ms_per_step_wall": med(wall), "kernel_ms": med(kms),
                "sweep_kernel_ms": med(sms), "hits": hits,
                "roofline": {"bound": "hbm", "achieved": (n * BYTES_PER_TASK + hits * 4) / (med(sms) * 1e-3) / 1e9,
                             "peak": peak, "unit": "GB/s",
                             "frac": (n * BYTES_PER_TASK + hits * 4) / (med(sms) * 1e-3) / 1e9 / peak,
                             "note": "K1 alone; 7.9 us at peak, so launch latency dominates at this size"}}
        print(json.dumps(line))
        table.close()
        return

    if args.workload == "creator":
        n, n_creators = 100_000_000, 1_000_000
        cfg = pkg.default_cfg(0x5EED0003, n, n_creators=n_creators)
        table = pkg.TaskTable(0).generate(cfg, 0, n)
        rng = np.random.default_rng(3)
        ks = rng.integers(0, n_creators, size=1000)
        names = [("user%07d@mail%d.example" % (k, k % 97)).encode() for k in ks]
        names += [b"absent%d@nowhere.example" % i for i in range(10)]
        single_scan, single_wall, hits = [], [], 0
        for i, q in enumerate(names[:200]):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = table.get_by_creator(q)
            single_wall.append((time.perf_counter() - t0) * 1e3)
            single_scan.append(table.last_op_stats().scan_ms)
            hits += r.count
            r.free()
        batch = {}
        for q in (16, 64, 256, 1000):
            scans, walls = [], []
            for rep in range(5):
                t0 = time.perf_counter()
                rs = table.get_by_creators(names[:q])
                walls.append((time.perf_counter() - t0) * 1e3)
                st = table.last_op_stats()
                scans.append(st.scan_ms)
                for r in rs:
                    r.free()
            batch[q] = {"scan_ms": med(scans), "order_ms": st.order_ms, "wall_ms": med(walls), "matches": st.matches,
                        "candidates": st.candidates, "queries_per_sec": q / (med(walls) * 1e-3),
                        "GBps": n * 8 / (med(scans) * 1e-3) / 1e9}
        sc = med(single_scan[5:])
        line = {"metric": "creator_filter_rows_per_sec", "workload": "config 4: GetTasksByCreator over 100M tasks, 1M creators",
                "value": n / (sc * 1e-3), "unit": "rows/s", "scan_ms": sc, "wall_ms_per_query": med(single_wall[5:]),
                "avg_hits_per_query": hits / 200,
                "roofline": {"bound": "hbm", "achieved": n * 8 / (sc * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": n * 8 / (sc * 1e-3) / 1e9 / peak, "kernel": "tt::creator_scan_kernel<true>",
                             "algorithmic_bytes": n * 8, "note": "8 B/row hash column; inputs 800 MB >> L2"},
                "batch": batch}
        print(json.dumps(line))
        table.close()
        return

    if args.workload == "creator-threads":
        # GetTasksByCreator from 64 request threads at once against the config-4 table (the reference's manager is a
        # singleton shared by every ASP.NET request thread, Api/Program.cs:13), beside one thread issuing the same
        # queries one after the other.  Concurrent calls are combined into shared table passes by the library.
        n, n_creators = 100_000_000, 1_000_000
        cfg = pkg.default_cfg(0x5EED0003, n, n_creators=n_creators)
        table = pkg.TaskTable(0).generate(cfg, 0, n)
        rng = np.random.default_rng(64)
        names = [("user%07d@mail%d.example" % (k, k % 97)).encode() for k in rng.integers(0, n_creators, size=64 * 50)]

        def issue(qs, lat):
            for q in qs:
                t0 = time.perf_counter()
                r = table.get_by_creator(q)
                lat.append((time.perf_counter() - t0) * 1e3)
                r.free()

        issue(names[:20], [])  # warm-up
        lat1 = []
        t0 = time.perf_counter()
        issue(names[:400], lat1)
        single = 400 / (time.perf_counter() - t0)
        p0, r0 = table.creator_combining_stats()
        lats = [[] for _ in range(64)]
        ths = [threading.Thread(target=issue, args=(names[i * 50:(i + 1) * 50], lats[i])) for i in range(64)]
        t0 = time.perf_counter()
        for th in ths:
            th.start()
        for th in ths:
            th.join()
        wall = time.perf_counter() - t0
        p1, r1 = table.creator_combining_stats()
        lat64 = sorted(x for l in lats for x in l)
        lat1.sort()
        pct = lambda xs, p: xs[min(len(xs) - 1, int(len(xs) * p))]
        cpu = None
        if not args.no_cpu_baseline:
            from oracle import oracle as orc

            sample, sample_creators = 10_000_000, 100_000
            ot = orc.OracleTable()
            ot.add_synth(orc.default_cfg(0x5EED0003, sample, n_creators=sample_creators), 0, sample)
            qs = [("user%07d@mail%d.example" % (k, k % 97)).encode() for k in rng.integers(0, sample_creators, size=5)]
            t0 = time.perf_counter()
            for q in qs:
                ot.get_by_creator(orc.MODE_FAKE, q)
            cdt = (time.perf_counter() - t0) / len(qs)
            cpu = {"value": sample / cdt, "unit": "rows/s", "cores": 1, "kind": "port", "ms_per_query": cdt * 1e3,
                   "sample": f"FakeTasksManager.GetTasksByCreator (LINQ Where + OrderByDescending, restated) over a {sample}-task "
                             f"List<TaskModel> with {sample_creators} creators, 1 thread"}
            ot.close()
        line = {"metric": "creator_queries_per_sec", "workload": "config 4 table (100M tasks, 1M creators): GetTasksByCreator from "
                "64 request threads x 50 queries, beside 1 thread x 400 queries", "value": 64 * 50 / wall, "unit": "queries/s",
                "threads": 64, "p50_ms": pct(lat64, 0.5), "p99_ms": pct(lat64, 0.99),
                "table_passes": p1 - p0, "requests": r1 - r0, "requests_per_pass": (r1 - r0) / max(1, p1 - p0),
                "single_thread": {"value": single, "unit": "queries/s", "p50_ms": pct(lat1, 0.5), "p99_ms": pct(lat1, 0.99)},
                "speedup_vs_single_thread": 64 * 50 / wall / single,
                "note": "Python request threads (ctypes releases the GIL inside the call); concurrent calls are combined "
                        "into one batch per table pass (tt_creator_combining_stats)",
                "cpu_baseline": cpu}
        print(json.dumps(line))
        table.close()
        return

    if args.workload == "json":
        # GetYesterdaysDueTasks (STORE) over the 1 B-task table, then the list as the
        # JSON text OverdueTasksController.Get sends.  CPU beside it: the oracle's System.Text.Json
        # restatement over a sample of the same list, one thread (JsonSerializer is sequential).
        n = args.rows
        cfg = pkg.default_cfg(SEED, n)
        table = pkg.TaskTable(local).generate(cfg, 0, n).generate_payload(cfg, 0)
        res = table.get_yesterdays_due(tt.MODE_STORE)
        size_ms, emit_ms, wall, nbytes = [], [], [], 0
        host = None
        for i in range(args.warmup + args.steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            j = res.json(tt.JSON_RESPONSE)
            dt = time.perf_counter() - t0
            st = table.last_json_stats()
            nbytes = j.size
            if host is None:
                host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
            if i == args.warmup + args.steps - 1:
                t1 = time.perf_counter()
                j.copy_to(host.data_ptr())
                d2h_ms = (time.perf_counter() - t1) * 1e3
            j.free()
            if i >= args.warmup:
                size_ms.append(st.size_ms)
                emit_ms.append(st.emit_ms)
                wall.append(dt * 1e3)
        text = bytes(host.numpy()[:2000])
        tail = bytes(host.numpy()[-200:])
        ok = text.startswith(b'[{"taskId":"') and tail.endswith(b'}]')
        cpu = None
        if not args.no_cpu_baseline:
            from oracle import oracle as orc

            sample_rows = 4_000_000
            ot = orc.OracleTable()
            ot.add_synth(orc.default_cfg(SEED, n), 0, sample_rows)
            listed = ot.get_yesterdays_due(orc.MODE_STORE)
            t0 = time.perf_counter()
            ref_text = ot.to_json(listed, orc.JSON_RESPONSE)
            cdt = time.perf_counter() - t0
            cpu = {"value": listed.size / cdt, "unit": "records/s", "cores": 1, "kind": "port",
                   "sample": f"{listed.size} records ({len(ref_text)} bytes) of the same list, oracle restatement of "
                             "System.Text.Json, 1 thread", "MBps": len(ref_text) / cdt / 1e6}
            ot.close()
        em = med(emit_ms)
        line = {"metric": "overdue_list_json_records_per_sec", "workload": f"GetYesterdaysDueTasks (STORE) over {n} tasks "
                "as the JSON array of OverdueTasksController.Get (MVC encoder)", "value": res.count / (med(wall) * 1e-3),
                "unit": "records/s", "records": res.count, "bytes": nbytes, "bytes_per_record": nbytes / max(res.count, 1),
                "size_ms": med(size_ms), "emit_ms": em, "wall_ms": med(wall), "d2h_ms_pinned": d2h_ms, "well_formed_ends": ok,
                "roofline": {"bound": "hbm", "achieved": nbytes / (em * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": nbytes / (em * 1e-3) / 1e9 / peak, "kernel": "tt::json_emit_kernel",
                             "algorithmic_bytes": nbytes,
                             "note": "output bytes only; each record also costs ~7 scattered 64 B DRAM reads "
                                     "(dates, refs, dictionary entries, escaped strings)"},
                "cpu_baseline": cpu}
        print(json.dumps(line))
        table.close()
        return

    if args.workload == "ingest":
        # the documents of the overdue list of the 1 B-task table (written by K6 with the
        # encoder Dapr's client uses) parsed and dictionary-encoded into a fresh table, straight from HBM.
        n = args.rows
        cfg = pkg.default_cfg(SEED, n)
        src = pkg.TaskTable(local).generate(cfg, 0, n).generate_payload(cfg, 0)
        res = src.get_yesterdays_due(tt.MODE_STORE)
        j = res.json(tt.JSON_DOCUMENT)
        off_dev = torch.empty(res.count + 1, dtype=torch.int64, device=dev)  # the offsets stay in HBM too
        j.offsets_to(off_dev.data_ptr())
        off = (off_dev.data_ptr(), res.count)
        parse, encode, wall = [], [], []
        for i in range(args.warmup + args.steps):
            dst = pkg.TaskTable(local)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            st = dst.load_documents(j.device_ptr(), off)
            dt = time.perf_counter() - t0
            if i >= args.warmup:
                parse.append(st.parse_ms)
                encode.append(st.encode_ms)
                wall.append(dt * 1e3)
            if i == args.warmup + args.steps - 1:
                back = dst.get_yesterdays_due(tt.MODE_STORE)
                same = back.count == res.count and back.digest()[:2] == res.digest()[:2]
            dst.close()
        cpu = None
        if not args.no_cpu_baseline:
            from oracle import oracle as orc

            ot = orc.OracleTable()
            ot.add_synth(orc.default_cfg(SEED, n), 0, 4_000_000)
            listed = ot.get_yesterdays_due(orc.MODE_STORE)
            text, doff = ot.to_documents(listed, orc.JSON_DOCUMENT)
            fresh = orc.OracleTable()
            t0 = time.perf_counter()
            assert fresh.add_documents(text, doff) == -1
            cdt = time.perf_counter() - t0
            cpu = {"value": listed.size / cdt, "unit": "documents/s", "cores": 1, "kind": "port",
                   "sample": f"{listed.size} documents ({len(text)} bytes) of the same list, oracle restatement of "
                             "JsonSerializer.Deserialize<TaskModel>, 1 thread", "MBps": len(text) / cdt / 1e6}
        pm = med(parse)
        line = {"metric": "document_ingest_docs_per_sec", "workload": f"{res.count} stored documents ({j.size} bytes, the "
                f"overdue list of a {n}-task table) -> SoA table, text and offsets resident in HBM", "value": res.count / (med(wall) * 1e-3),
                "unit": "documents/s", "documents": res.count, "bytes": j.size, "parse_ms": pm, "encode_ms": med(encode),
                "wall_ms": med(wall), "dictionaries": {"names": st.n_names, "creators": st.n_creators, "assignees": st.n_assignees},
                "launches": st.launches, "reloaded_table_lists_the_same_tasks": bool(same),
                "roofline": {"bound": "hbm", "achieved": j.size / (pm * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": j.size / (pm * 1e-3) / 1e9 / peak, "kernel": "tt::doc_parse_kernel",
                             "algorithmic_bytes": j.size, "note": "text bytes read by the parse kernel"},
                "cpu_baseline": cpu}
        print(json.dumps(line))
        j.free()
        src.close()
        return

    if args.workload == "mixed":
        # BASELINE.json config 5: ONE 1 B-task table (args.rows) sharded by TaskId range over the
        # GPUs; per tick: sweep + hit id gather on rank 0, then 64 creator filters as one batch on
        # every shard and a sum of the per-creator match counts.
        import torch.distributed as dist

        world = int(os.environ.get("WORLD_SIZE", "1"))
        rank = int(os.environ.get("RANK", "0"))
        if world > 1:
            dist.init_process_group("nccl", device_id=dev)
        n_total = args.rows
        plan = pkg.ShardPlan(n_total, world, rank)
        n = plan.n_rows
        cfg = pkg.default_cfg(SEED, n_total)
        table = pkg.TaskTable(local).generate(cfg, plan.first_row, n)
        nbytes = (n + 7) // 8
        pristine = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        tt._check(tt.lib().tt_snapshot_flags(table._h, None, pristine.data_ptr()))
        cap = int(n * 0.032) + 65536
        ex = pkg.HitExchange(plan, cap, dev, with_ids=False)
        d_ids = torch.empty((cap, 16), dtype=torch.uint8, device=dev)
        rng = np.random.default_rng(5)
        names = [("user%07d@mail%d.example" % (k, k % 97)).encode() for k in rng.integers(0, 1_000_000, size=64)]
        counts_dev = torch.zeros(64, dtype=torch.int64, device=dev)
        counts_host = torch.zeros(64, dtype=torch.int64).pin_memory()

        def sync_all():
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()

        wall, sweep_k, scan_k, order_k, matches = [], [], [], [], 0
        for i in range(args.warmup + args.steps):
            tt._check(tt.lib().tt_restore_flags(table._h, None, pristine.data_ptr()))
            if i == args.warmup - 1 and world > 1:
                ex.tune_send_cap(int(n * 0.0255))
            sync_all()
            t0 = time.perf_counter()
            st, _ = table.sweep_device(tt.MODE_STORE, tt.SWEEP_MARK, ex.rows_all.data_ptr(), d_ids.data_ptr(), cap,
                                       stream=stream.cuda_stream)
            if world > 1:
                ex.run_fixed(st.n_stored)
            rs = table.get_by_creators(names)
            for k, r in enumerate(rs):
                counts_host[k] = r.count
            counts_dev.copy_(counts_host, non_blocking=True)
            if world > 1:
                dist.all_reduce(counts_dev)
            sync_all()
            dt = time.perf_counter() - t0
            op = table.last_op_stats()
            for r in rs:
                r.free()
            if i >= args.warmup:
                wall.append(dt * 1e3)
                sweep_k.append(st.kernel_ms)
                scan_k.append(op.scan_ms)
                order_k.append(op.order_ms)
                matches = int(counts_dev.sum().item())
        w = torch.tensor([med(wall)], dtype=torch.float64, device=dev)
        hits = torch.tensor([st.n_listed], dtype=torch.int64, device=dev)
        if world > 1:
            dist.all_reduce(w, op=dist.ReduceOp.MAX)
            dist.all_reduce(hits)
        if rank == 0:
            line = {"metric": "mixed_cron_batch_ms", "n_gpus": world,
                    "workload": f"config 5: {n_total}-task table over {world} GPU(s), sweep + hit id gather + 64 creator "
                                "filters (one batch per shard) + match-count all-reduce",
                    "value": float(w.item()), "unit": "ms", "higher_is_better": False, "rows_per_gpu": n,
                    "sweep_kernels_ms": med(sweep_k), "creator_scan_ms": med(scan_k), "creator_order_ms": med(order_k),
                    "hits": int(hits.item()), "creator_matches": matches,
                    "roofline": {"bound": "hbm", "unit": "GB/s", "peak": peak,
                                 "creator_scan_achieved": n * 8 / (med(scan_k) * 1e-3) / 1e9,
                                 "creator_scan_frac": n * 8 / (med(scan_k) * 1e-3) / 1e9 / peak}}
            print(json.dumps(line))
        table.close()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload != "sweep":
        run_extra(a)
    else:
        run_b200(a)
