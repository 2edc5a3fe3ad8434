"""Times tt_get_yesterdays_due (STORE: sweep + order by TaskCreatedOn) and tt_sweep through the API."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000_000
table = pkg.TaskTable(0).generate(pkg.default_cfg(0x5EED0002, n), 0, n)
for mode, name in ((tt.MODE_FAKE, "FAKE"), (tt.MODE_STORE, "STORE")):
    for rep in range(3):
        t0 = time.perf_counter()
        r = table.get_yesterdays_due(mode)
        dt = time.perf_counter() - t0
        print(f"get_yesterdays_due {name}: {r.count} hits, {dt*1e3:.2f} ms")
        r.free()
