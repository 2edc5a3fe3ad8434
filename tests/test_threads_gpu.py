"""One table shared by many request threads (the reference registers its manager as a singleton,
Api/Program.cs:13): concurrent reads and sweeps through the C ABI must stay consistent."""
import threading

import numpy as np
import pytest

from helpers import tt

pytestmark = pytest.mark.gpu


def test_concurrent_calls_on_one_table(oracle):
    n = 400_000
    cfg = oracle.default_cfg(123, n, n_creators=500)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    want_fake = ot.get_yesterdays_due(oracle.MODE_FAKE)
    creators = [ot.created_by(k) for k in range(40)]
    want_by = {c: ot.get_by_creator(oracle.MODE_FAKE, c) for c in creators}
    errors = []

    def reader(seed):
        rng = np.random.default_rng(seed)
        try:
            for _ in range(30):
                if rng.random() < 0.5:
                    c = creators[int(rng.integers(0, len(creators)))]
                    got = table.get_by_creator(c).rows()
                    assert got.tolist() == want_by[c].tolist()
                else:
                    got = table.get_yesterdays_due(tt.MODE_FAKE).rows()  # FAKE ignores the flags the sweeper changes
                    assert (got == want_fake).all()
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    def sweeper():
        try:
            total = 0
            for _ in range(5):
                res, marked, _ = table.sweep(tt.MODE_STORE)
                total += marked
            _, listed, _ = ot.cron_job(oracle.MODE_STORE)
            assert total == listed.size  # the first tick marks everything, the later ones nothing
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=reader, args=(s,)) for s in range(6)] + [threading.Thread(target=sweeper)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    assert (table.flags_bytes()[1] == ot.flags()[1]).all()
    table.close()


def test_concurrent_creator_requests_share_table_passes(oracle):
    """GetTasksByCreator from many request threads at once: requests that arrive while a pass over the table is
    running are answered together by the next pass.  Every thread still gets exactly its own creator's list."""
    n = 4_000_000
    cfg = oracle.default_cfg(321, n, n_creators=2000)
    ot = oracle.OracleTable()
    ot.add_synth(cfg, 0, n)
    table = tt.TaskTable(0).generate(cfg, 0, n)
    creators = [ot.created_by(k) for k in range(200)] + [b"nobody@example.org"]
    want = {c: ot.get_by_creator(oracle.MODE_STORE, c).tolist() for c in set(creators)}
    errors, per_thread = [], 40
    start = threading.Barrier(32)

    def request_thread(seed):
        rng = np.random.default_rng(seed)
        try:
            start.wait()
            for _ in range(per_thread):
                c = creators[int(rng.integers(0, len(creators)))]
                res = table.get_by_creator(c)
                assert res.rows().tolist() == want[c]
                res.free()
        except Exception as e:  # noqa: BLE001
            errors.append(repr(e))

    threads = [threading.Thread(target=request_thread, args=(s,)) for s in range(32)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors[:3]
    passes, requests = table.creator_combining_stats()
    assert requests == 32 * per_thread and 0 < passes < requests  # some passes answered several requests
    # a bad request in the queue fails alone at the door, not inside somebody else's batch
    with pytest.raises(tt.TaskTableError):
        tt._check(tt.lib().tt_get_by_creator(table._h, None, 5, None))
    table.close()
