"""Shared helpers for the parity tests: the same rows in the oracle and in the device table."""
import importlib

import numpy as np

pkg = importlib.import_module("aca-dotnet-workshop_b200")
tt = pkg.tasktable


def be_u64(ids, lo_half):
    """big-endian u64 view of bytes 0..7 (lo_half False) or 8..15 of [n,16] ids"""
    part = ids[:, 8:] if lo_half else ids[:, :8]
    return np.ascontiguousarray(part).view(">u8").ravel().astype(np.uint64)


def xor_reduce(a):
    return int(np.bitwise_xor.reduce(a)) if a.size else 0


def oracle_columns(ot):
    """Everything tt_load_soa needs, taken from an OracleTable."""
    n = len(ot)
    ids, created, due = ot.export()
    comp, over = ot.flags()
    creators, refs = {}, np.empty(n, dtype=np.uint32)
    for i in range(n):
        refs[i] = creators.setdefault(ot.created_by(i), len(creators))
    return dict(task_id=ids, created_raw=created, due_raw=due, completed=comp, overdue=over, creator_ref=refs,
                creators=list(creators.keys()) or [b""])


def load_from_oracle(ot, device=0):
    return tt.TaskTable(device).load_soa(**oracle_columns(ot))
