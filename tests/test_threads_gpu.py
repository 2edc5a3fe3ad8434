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
