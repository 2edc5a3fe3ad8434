"""Soak test of the tick's end-of-call protocol (programmatic launches, the sequence word in mapped host memory, the
exchange's epoch / consumed words): thousands of back-to-back untimed ticks on small tables, every outcome compared
with the first one.  One GPU."""
import importlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
for n in (50_000, 3_000_000):
    cfg = pkg.default_cfg(0xABCD, n)
    table = pkg.TaskTable(0).generate(cfg, 0, n)
    first = table.get_yesterdays_due(tt.MODE_FAKE)
    want_rows, want_ids = first.rows().copy(), first.ids().copy()
    t0 = time.perf_counter()
    for k in range(rounds):
        r = table.get_yesterdays_due(tt.MODE_FAKE)      # an untimed listing tick + TaskIds
        assert r.count == len(want_rows), (n, k, r.count)
        if k % 97 == 0:
            assert (r.rows() == want_rows).all() and (r.ids() == want_ids).all(), (n, k)
        r.free()
    print(f"{n} rows: {rounds} ticks, {len(want_rows)} hits each, {(time.perf_counter() - t0) / rounds * 1e6:.1f} us per call", flush=True)
    table.close()

# three shards on one device, ticks into the same slots, collected and released every time
n, world = 600_000, 3
cfg = pkg.default_cfg(0xBEEF, n)
bounds = [0, 190_000, 410_001, n]
tables = [pkg.TaskTable(0).generate(cfg, bounds[g], bounds[g + 1] - bounds[g]) for g in range(world)]
tiles = max((len(t) + 8191) // 8192 for t in tables)
root, _ = tt.Exchange.create_root(0, world, 65536, tiles, tt.EXCHANGE_IDS)
ex = [root] + [root.attach_local(0, g) for g in range(1, world)]
want = None
t0 = time.perf_counter()
for k in range(rounds // 3):
    for g in range(world):
        ex[g].tick(tables[g], tt.MODE_FAKE, tt.SWEEP_LIST_ONLY | tt.SWEEP_ASYNC | tt.SWEEP_NO_TIMING)
    res = root.collect(tt.ORDER_ROWS, bounds[:world])
    for g in range(world):
        tables[g].sweep_wait()
    rows = res.rows()
    if want is None:
        want = (rows.copy(), res.ids().copy())
    assert len(rows) == len(want[0]) and (rows == want[0]).all(), k
    if k % 50 == 0:
        assert (res.ids() == want[1]).all(), k
    root.release()
    res.free()
print(f"3 shards: {rounds // 3} exchange ticks, {len(want[0])} hits each, {(time.perf_counter() - t0) / (rounds // 3) * 1e6:.1f} us per round", flush=True)
print("stress ok")
