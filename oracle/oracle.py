"""ctypes binding of oracle/liboracle.so (the CPU restatement in oracle.cpp).

TEST INFRASTRUCTURE ONLY — imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package.  Parity is UNPINNED:
see the header of oracle.cpp.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")

MODE_FAKE = 0
MODE_STORE = 1

TICKS_PER_DAY = 864000000000
DEFAULT_TODAY = 639295200000000000  # 2026-11-06T00:00:00
DEFAULT_RUN_AT = DEFAULT_TODAY + 5 * 60 * 10_000_000


class SynthCfg(C.Structure):
    """Mirror of tt_synth_cfg (include/tt_synth.h)."""

    _fields_ = [
        ("seed", C.c_uint64),
        ("n_total", C.c_uint64),
        ("today_local_ticks", C.c_int64),
        ("ppm_due_yesterday", C.c_uint32),
        ("ppm_adversarial", C.c_uint32),
        ("ppm_completed", C.c_uint32),
        ("ppm_overdue_past", C.c_uint32),
        ("ppm_created_dup", C.c_uint32),
        ("ppm_timed_due", C.c_uint32),
        ("n_creators", C.c_uint32),
        ("n_names", C.c_uint32),
    ]


def default_cfg(seed, n_total, **kw):
    cfg = SynthCfg(seed, n_total, DEFAULT_TODAY, 50000, 10000, 500000, 100000, 1000, 100000, 1000000, 0)
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


class StreamOut(C.Structure):
    _fields_ = [
        ("n_listed", C.c_int64),
        ("n_marked", C.c_int64),
        ("xor_id_hi", C.c_uint64),
        ("xor_id_lo", C.c_uint64),
        ("sum_rows", C.c_uint64),
        ("overdue_after", C.c_int64),
        ("completed", C.c_int64),
        ("threw", C.c_int32),
    ]


def build(force=False):
    src = os.path.join(HERE, "oracle.cpp")
    hdr = os.path.join(HERE, "..", "include", "tt_synth.h")
    if not force and os.path.exists(LIB_PATH):
        if not os.path.exists(src):
            return LIB_PATH
        if os.path.getmtime(LIB_PATH) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
            return LIB_PATH
    subprocess.check_call(
        ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", LIB_PATH, src]
    )
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(build())
    vp, i32, i64, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64
    L.orc_create.restype = vp
    L.orc_destroy.argtypes = [vp]
    L.orc_count.argtypes = [vp]
    L.orc_count.restype = i64
    L.orc_add.argtypes = [vp, vp, C.c_char_p, i32, C.c_char_p, i32, C.c_char_p, i32, u64, u64, i32, i32]
    L.orc_add.restype = i64
    L.orc_add_synth.argtypes = [vp, C.POINTER(SynthCfg), u64, u64]
    L.orc_add_synth.restype = i64
    L.orc_seed_fake.argtypes = [vp, i64, u64]
    L.orc_get_yesterdays_due.argtypes = [vp, i32, i64, i32, i32, vp, i64]
    L.orc_get_yesterdays_due.restype = i64
    L.orc_cron_job.argtypes = [vp, i32, i64, i64, i32, i32, C.POINTER(i64), C.POINTER(i64), vp, i64]
    L.orc_cron_job.restype = i32
    L.orc_backlog.argtypes = [vp, i32, i32, i64, i64, C.POINTER(i64), C.POINTER(i32), vp, i64]
    L.orc_backlog.restype = i64
    L.orc_mark_overdue.argtypes = [vp, i32, vp, i64, C.POINTER(i64)]
    L.orc_mark_overdue.restype = i32
    L.orc_get_by_creator.argtypes = [vp, i32, C.c_char_p, i32, vp, i64]
    L.orc_get_by_creator.restype = i64
    L.orc_get_flags.argtypes = [vp, vp, vp]
    L.orc_set_flags.argtypes = [vp, i64, i32, i32]
    L.orc_build_keys.argtypes = [vp]
    L.orc_live_count.argtypes = [vp]
    L.orc_live_count.restype = i64
    L.orc_get_by_id.argtypes = [vp, i32, vp]
    L.orc_get_by_id.restype = i64
    L.orc_mark_completed.argtypes = [vp, i32, vp]
    L.orc_update_task.argtypes = [vp, i32, vp, C.c_char_p, i32, C.c_char_p, i32, u64]
    L.orc_delete_task.argtypes = [vp, i32, vp]
    L.orc_set_all_flags.argtypes = [vp, vp, vp]
    L.orc_soa_reset.argtypes = [vp, C.POINTER(SynthCfg), u64, i32]
    L.orc_export.argtypes = [vp, vp, vp, vp]
    L.orc_created_by.argtypes = [vp, i64, C.c_char_p, i32]
    L.orc_created_by.restype = i32
    L.orc_format_stj.argtypes = [u64, i32, C.c_char_p, i32]
    L.orc_format_converter.argtypes = [u64, C.c_char_p, i32]
    L.orc_guid_string.argtypes = [vp, C.c_char_p]
    L.orc_serialize.argtypes = [vp, vp, i64, i32, i32, vp, i64]
    L.orc_serialize.restype = i64
    L.orc_string_field.argtypes = [vp, i64, i32, C.c_char_p, i32]
    L.orc_string_field.restype = i32
    L.orc_add_documents.argtypes = [vp, vp, vp, i64, i32]
    L.orc_add_documents.restype = i64
    L.orc_serialize_documents.argtypes = [vp, vp, i64, i32, i32, vp, i64, vp]
    L.orc_serialize_documents.restype = i64
    L.orc_stream_sweep.argtypes = [C.POINTER(SynthCfg), u64, u64, i32, i64, i64, i32, C.POINTER(StreamOut)]
    L.orc_stream_creators.argtypes = [C.POINTER(SynthCfg), u64, u64, C.c_char_p, C.c_void_p, i32, i32, C.c_void_p]
    L.orc_stream_creators.restype = None
    L.orc_soa_create.argtypes = [C.POINTER(SynthCfg), u64, i64, i32]
    L.orc_soa_create.restype = vp
    L.orc_soa_destroy.argtypes = [vp]
    L.orc_soa_sweep.argtypes = [vp, i64, i64, i32, vp, i64]
    L.orc_soa_sweep.restype = i64
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


JSON_RESPONSE = 0  # ASP.NET Core MVC output formatter (UnsafeRelaxedJsonEscaping)
JSON_DOCUMENT = 1  # Dapr client bodies (JavaScriptEncoder.Default)


class OracleTable:
    """The reference's task list / state store, restated (oracle.cpp)."""

    def __init__(self):
        self._l = lib()
        self._h = C.c_void_p(self._l.orc_create())

    def close(self):
        if self._h:
            self._l.orc_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def __len__(self):
        return self._l.orc_count(self._h)

    def add(self, task_id, name, created_by, assigned_to, created_raw, due_raw, completed=False, overdue=False):
        tid = np.frombuffer(bytes(task_id), dtype=np.uint8).copy()
        nb = name.encode() if isinstance(name, str) else bytes(name)
        cb = created_by.encode() if isinstance(created_by, str) else bytes(created_by)
        ab = assigned_to.encode() if isinstance(assigned_to, str) else bytes(assigned_to)
        return self._l.orc_add(self._h, _ptr(tid), nb, len(nb), cb, len(cb), ab, len(ab), created_raw, due_raw,
                               int(completed), int(overdue))

    def add_synth(self, cfg, first_row, n):
        return self._l.orc_add_synth(self._h, C.byref(cfg), first_row, n)

    def seed_fake(self, now_utc_ticks, rng_seed=1):
        self._l.orc_seed_fake(self._h, now_utc_ticks, rng_seed)

    def get_yesterdays_due(self, mode, today=DEFAULT_TODAY, local_offset_min=0, string_eq=False):
        out = np.empty(len(self), dtype=np.uint32)
        n = self._l.orc_get_yesterdays_due(self._h, mode, today, local_offset_min, int(string_eq), _ptr(out), len(out))
        return out[:n].copy()

    def cron_job(self, mode, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT, local_offset_min=0, string_eq=False):
        """-> (threw, listed_rows, n_overdue)"""
        out = np.empty(len(self), dtype=np.uint32)
        nl, no = C.c_int64(0), C.c_int64(0)
        st = self._l.orc_cron_job(self._h, mode, today, run_at, local_offset_min, int(string_eq), C.byref(nl),
                                  C.byref(no), _ptr(out), len(out))
        return bool(st), out[: nl.value].copy(), no.value

    def backlog(self, mode, mark=False, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT):
        """EXTENSION: tasks due yesterday or earlier [and pending, STORE]; with mark, the cron
        handler's filter + MarkOverdueTasks.  -> (rows in list order, n_overdue, threw)."""
        n = len(self)
        out = np.empty(max(n, 1), dtype=np.uint32)
        n_over, threw = C.c_int64(0), C.c_int32(0)
        k = self._l.orc_backlog(self._h, mode, 1 if mark else 0, today, run_at, C.byref(n_over), C.byref(threw), _ptr(out), n)
        return out[:k].copy(), n_over.value, bool(threw.value)

    def mark_overdue(self, mode, ids16):
        ids = np.ascontiguousarray(ids16, dtype=np.uint8).reshape(-1, 16)
        miss = C.c_int64(0)
        st = self._l.orc_mark_overdue(self._h, mode, _ptr(ids), ids.shape[0], C.byref(miss))
        return bool(st), miss.value

    def get_by_creator(self, mode, created_by):
        cb = created_by.encode() if isinstance(created_by, str) else bytes(created_by)
        out = np.empty(len(self), dtype=np.uint32)
        n = self._l.orc_get_by_creator(self._h, mode, cb, len(cb), _ptr(out), len(out))
        if n == -2:  # FakeTasksManager.cs:70 calls .Equals on a null TaskCreatedBy
            raise RuntimeError("NullReferenceException")
        return out[:n].copy()

    def flags(self):
        n = len(self)
        c = np.empty(n, dtype=np.uint8)
        o = np.empty(n, dtype=np.uint8)
        self._l.orc_get_flags(self._h, _ptr(c), _ptr(o))
        return c, o

    # -- ITasksManager point operations (return codes: see oracle.cpp) --
    def live_count(self):
        return self._l.orc_live_count(self._h)

    def get_by_id(self, mode, task_id):
        a = np.frombuffer(bytes(task_id), dtype=np.uint8).copy()
        return self._l.orc_get_by_id(self._h, mode, _ptr(a))

    def mark_completed(self, mode, task_id):
        a = np.frombuffer(bytes(task_id), dtype=np.uint8).copy()
        return self._l.orc_mark_completed(self._h, mode, _ptr(a))

    def update_task(self, mode, task_id, name, assigned_to, due_raw):
        a = np.frombuffer(bytes(task_id), dtype=np.uint8).copy()
        nb, ab = name.encode(), assigned_to.encode()
        return self._l.orc_update_task(self._h, mode, _ptr(a), nb, len(nb), ab, len(ab), due_raw)

    def delete_task(self, mode, task_id):
        a = np.frombuffer(bytes(task_id), dtype=np.uint8).copy()
        return self._l.orc_delete_task(self._h, mode, _ptr(a))

    def build_keys(self):
        self._l.orc_build_keys(self._h)

    def set_all_flags(self, completed, overdue):
        c = np.ascontiguousarray(completed, dtype=np.uint8)
        o = np.ascontiguousarray(overdue, dtype=np.uint8)
        assert c.size == len(self) and o.size == len(self)
        self._l.orc_set_all_flags(self._h, _ptr(c), _ptr(o))

    def set_flags(self, row, completed, overdue):
        self._l.orc_set_flags(self._h, row, int(completed), int(overdue))

    def export(self):
        """-> (ids[n,16] u8, created_raw u64[n], due_raw u64[n])"""
        n = len(self)
        ids = np.empty((n, 16), dtype=np.uint8)
        cr = np.empty(n, dtype=np.uint64)
        du = np.empty(n, dtype=np.uint64)
        self._l.orc_export(self._h, _ptr(ids), _ptr(cr), _ptr(du))
        return ids, cr, du

    def created_by(self, row):
        return self.string_field(row, 1)

    def string_field(self, row, which):
        """UTF-8 of TaskName (which=0), TaskCreatedBy (1) or TaskAssignedTo (2) of one row."""
        buf = C.create_string_buffer(4096)
        n = self._l.orc_string_field(self._h, row, which, buf, 4096)
        return buf.raw[:n]

    def add_documents(self, text, off, local_offset_min=0):
        """JsonSerializer.Deserialize<TaskModel> of every document text[off[i]:off[i+1]], appended in order.
        Returns -1, or the index of the first document the reference would reject (nothing is added)."""
        buf = np.frombuffer(bytes(text) + b"\0", dtype=np.uint8)
        off = np.ascontiguousarray(off, dtype=np.uint64)
        return int(self._l.orc_add_documents(self._h, _ptr(buf), _ptr(off), off.size - 1, local_offset_min))

    def to_documents(self, rows, flavor=JSON_DOCUMENT, local_offset_min=0):
        """One stored document per listed row -> (bytes, offsets uint64[n + 1])."""
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        off = np.zeros(rows.size + 1, dtype=np.uint64)
        need = self._l.orc_serialize_documents(self._h, _ptr(rows), rows.size, flavor, local_offset_min, None, 0, None)
        out = np.empty(max(need, 1), dtype=np.uint8)
        self._l.orc_serialize_documents(self._h, _ptr(rows), rows.size, flavor, local_offset_min, _ptr(out), need, _ptr(off))
        return out[:need].tobytes(), off

    def to_json(self, rows, flavor=JSON_RESPONSE, local_offset_min=0):
        """JsonSerializer.Serialize(List<TaskModel>) of the listed rows, in that order (bytes)."""
        rows = np.ascontiguousarray(rows, dtype=np.uint32)
        need = self._l.orc_serialize(self._h, _ptr(rows), rows.size, flavor, local_offset_min, None, 0)
        out = np.empty(need, dtype=np.uint8)
        self._l.orc_serialize(self._h, _ptr(rows), rows.size, flavor, local_offset_min, _ptr(out), need)
        return out.tobytes()


def format_stj(raw, local_offset_min=0):
    buf = C.create_string_buffer(64)
    lib().orc_format_stj(raw, local_offset_min, buf, 64)
    return buf.value.decode()


def format_converter(raw):
    buf = C.create_string_buffer(64)
    lib().orc_format_converter(raw, buf, 64)
    return buf.value.decode()


def guid_string(id16):
    a = np.frombuffer(bytes(id16), dtype=np.uint8).copy()
    buf = C.create_string_buffer(40)
    lib().orc_guid_string(_ptr(a), buf)
    return buf.value.decode()


def stream_sweep(cfg, first_row, n, mode, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT, threads=None):
    out = StreamOut()
    lib().orc_stream_sweep(C.byref(cfg), first_row, n, mode, today, run_at, threads or (os.cpu_count() or 1),
                           C.byref(out))
    return out


class CreatorOut(C.Structure):
    _fields_ = [("count", C.c_int64), ("xor_id_hi", C.c_uint64), ("xor_id_lo", C.c_uint64), ("sum_rows", C.c_uint64),
                ("min_row", C.c_int64), ("max_row", C.c_int64)]


def stream_creators(cfg, first_row, n, queries, threads=None):
    """GetTasksByCreator for every query over rows generated on the fly (no table is stored): per query the
    order-free summary (count, XOR of TaskIds, sum / min / max of the rows).  queries: list of bytes."""
    qs = [q.encode() if isinstance(q, str) else bytes(q) for q in queries]
    off = np.zeros(len(qs) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(q) for q in qs], dtype=np.uint64) if qs else 0
    pool = b"".join(qs) or b"\0"
    out = (CreatorOut * len(qs))()
    lib().orc_stream_creators(C.byref(cfg), first_row, n, pool, _ptr(off), len(qs), threads or (os.cpu_count() or 1), out)
    return list(out)


class SoaTable:
    """Best-effort multi-core SoA sweep (context number, not the reference's layout)."""

    def __init__(self, cfg, first_row, n, threads=None):
        self._l = lib()
        self.n = n
        self.threads = threads or (os.cpu_count() or 1)
        self._cfg, self._first = cfg, first_row
        self._h = C.c_void_p(self._l.orc_soa_create(C.byref(cfg), first_row, n, self.threads))

    def reset(self):
        self._l.orc_soa_reset(self._h, C.byref(self._cfg), self._first, self.threads)

    def sweep(self, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT, cap=None):
        cap = cap or self.n
        out = np.empty(cap, dtype=np.uint32)
        n = self._l.orc_soa_sweep(self._h, today, run_at, self.threads, _ptr(out), cap)
        return n, out[: min(n, cap)]

    def close(self):
        if self._h:
            self._l.orc_soa_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()
