"""ctypes binding of libtasktable.so — one function per entry point of include/tasktable.h.

This is the same binding surface a C# ``[DllImport]`` wrapper uses: plain
pointers and sizes.  The library is loaded from this directory (built in-tree by
``make`` / ``__graft_entry__.build()``); a missing library is an ImportError-like
``TaskTableError`` at first use, never a silent fallback.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

MODE_FAKE = 0
MODE_STORE = 1
SWEEP_LIST_ONLY = 0
SWEEP_MARK = 1
SWEEP_DUE_ON_OR_BEFORE = 2  # TT_SWEEP_DUE_ON_OR_BEFORE: the generalised predicate (OR it into the flags)
SWEEP_NO_TIMING = 4         # TT_SWEEP_NO_TIMING: no CUDA events between the kernels, timing fields stay 0
SWEEP_ASYNC = 8             # TT_SWEEP_ASYNC: queue the tick and return; sweep_wait() collects the outcome

TT_OK = 0
TT_ERR_ARG = -1
TT_ERR_CUDA = -2
TT_ERR_OOM = -3
TT_ERR_NOT_IMPLEMENTED = -4
TT_ERR_CAPACITY = -5
TT_ERR_STATE = -6

FIELD_CREATED_RAW = 1
FIELD_DUE_RAW = 2
FIELD_COMPLETED = 3
FIELD_OVERDUE = 4
FIELD_CREATOR_REF = 5

JSON_RESPONSE = 0  # TT_JSON_RESPONSE: the MVC output formatter's encoder
JSON_DOCUMENT = 1  # TT_JSON_DOCUMENT: the encoder of Dapr's client bodies

EXCHANGE_IDS = 1       # TT_EXCHANGE_IDS: the slots carry the hits' TaskIds
EXCHANGE_CREATED = 2   # TT_EXCHANGE_CREATED: ... and their TaskCreatedOn (the cross-shard OrderBy)
ORDER_ROWS = 0         # TT_ORDER_ROWS
ORDER_CREATED_ASC = 1  # TT_ORDER_CREATED_ASC
COLLECT_RELEASE = 0x100
MAX_SHARDS = 32

TICKS_PER_DAY = 864000000000
DEFAULT_TODAY = 639295200000000000  # 2026-11-06T00:00:00
DEFAULT_RUN_AT = DEFAULT_TODAY + 5 * 60 * 10_000_000
TILE_ROWS = 8192
STATIC_MAX_ROWS = 16777216  # TT_STATIC_MAX_ROWS
STATIC_MAX_HITS = 1048576   # TT_STATIC_MAX_HITS

# every symbol include/tasktable.h declares (tests check the library exports all of them)
EXPORTS = [
    "tt_create", "tt_destroy", "tt_last_error", "tt_generate", "tt_load_soa", "tt_row_count",
    "tt_get_yesterdays_due", "tt_mark_overdue", "tt_sweep", "tt_get_by_creator",
    "tt_result_count", "tt_result_rows", "tt_result_ids", "tt_result_field", "tt_result_free",
    "tt_snapshot_flags", "tt_restore_flags", "tt_count_flags", "tt_result_digest",
    "tt_sweep_device", "tt_get_device_view", "tt_sweep_static", "tt_get_by_creators", "tt_last_op_stats",
    "tt_get_by_id", "tt_insert_task", "tt_update_task", "tt_mark_completed", "tt_delete_task", "tt_live_count",
    "tt_save_snapshot", "tt_load_snapshot", "tt_pack_hits", "tt_expand_hits",
    "tt_sweep_push", "tt_peer_alloc", "tt_peer_open", "tt_peer_close", "tt_peer_free",
    "tt_load_payload", "tt_generate_payload", "tt_result_json", "tt_json_size", "tt_json_copy", "tt_json_device",
    "tt_json_free", "tt_last_json_stats", "tt_json_offsets", "tt_load_documents",
    "tt_get_due_on_or_before", "tt_sweep_backlog", "tt_tick_host", "tt_set_task_strings", "tt_upsert_task",
    "tt_sweep_wait", "tt_creator_combining_stats", "tt_last_order_info",
    "tt_exchange_create_root", "tt_exchange_attach", "tt_exchange_attach_local", "tt_exchange_destroy",
    "tt_exchange_slot_bytes", "tt_exchange_tick", "tt_exchange_epoch", "tt_exchange_collect", "tt_exchange_release",
    "tt_exchange_last_collect_ms",
    "tt_gresult_count", "tt_gresult_shard_counts", "tt_gresult_rows", "tt_gresult_ids", "tt_gresult_digest", "tt_gresult_free",
    "tt_shards_create", "tt_shards_destroy", "tt_shards_count", "tt_shards_table", "tt_shards_generate",
    "tt_shards_generate_payload", "tt_shards_load_soa", "tt_shards_row_count", "tt_shards_get_yesterdays_due",
    "tt_shards_get_due_on_or_before", "tt_shards_sweep", "tt_shards_mark_overdue", "tt_shards_get_by_creator",
    "tt_shards_get_by_id", "tt_shards_insert_task", "tt_shards_update_task", "tt_shards_mark_completed",
    "tt_shards_delete_task",
]


class TaskTableError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"tasktable error {code}: {text}")
        self.code = code


class NotImplementedInReference(TaskTableError):
    """FAKE mode: FakeTasksManager.MarkOverdueTasks throws NotImplementedException
    (Api/Services/FakeTasksManager.cs:102-105)."""


class SynthCfg(C.Structure):
    """tt_synth_cfg (include/tt_synth.h)."""

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
    """BASELINE.json config-5 mix: 5 % due yesterday, 1 % adversarial, 50 % completed."""
    cfg = SynthCfg(seed, n_total, DEFAULT_TODAY, 50000, 10000, 500000, 100000, 1000, 100000, 1000000, 0)
    for k, v in kw.items():
        setattr(cfg, k, v)
    return cfg


class SoaView(C.Structure):
    _fields_ = [
        ("n", C.c_int64),
        ("task_id", C.c_void_p),
        ("created_raw", C.c_void_p),
        ("due_raw", C.c_void_p),
        ("completed", C.c_void_p),
        ("overdue", C.c_void_p),
        ("creator_ref", C.c_void_p),
        ("n_creators", C.c_uint32),
        ("creator_off", C.c_void_p),
        ("creator_pool", C.c_void_p),
    ]


class SweepStats(C.Structure):
    _fields_ = [
        ("n_listed", C.c_int64),
        ("n_stored", C.c_int64),
        ("n_marked", C.c_int64),
        ("resume_row", C.c_int64),
        ("kernel_ms", C.c_float),
        ("launches", C.c_int32),
        ("sweep_ms", C.c_float),
        ("gather_ms", C.c_float),
    ]


class OpStats(C.Structure):
    _fields_ = [
        ("scan_ms", C.c_float),
        ("order_ms", C.c_float),
        ("candidates", C.c_int64),
        ("matches", C.c_int64),
        ("launches", C.c_int32),
        ("queries", C.c_int32),
    ]


class DeviceView(C.Structure):
    _fields_ = [
        ("n", C.c_int64),
        ("n_padded", C.c_int64),
        ("due_code", C.c_void_p),
        ("completed", C.c_void_p),
        ("overdue", C.c_void_p),
        ("creator_hash", C.c_void_p),
        ("task_id", C.c_void_p),
        ("created_raw", C.c_void_p),
        ("device", C.c_int32),
    ]


class StringsView(C.Structure):
    """tt_strings_view: a dictionary-encoded string column (host arrays)."""
    _fields_ = [("ref", C.c_void_p), ("n_strings", C.c_uint32), ("off", C.c_void_p), ("pool", C.c_void_p)]


class JsonStats(C.Structure):
    _fields_ = [("size_ms", C.c_float), ("emit_ms", C.c_float), ("bytes", C.c_int64), ("records", C.c_int64),
                ("launches", C.c_int32)]


class IngestStats(C.Structure):
    _fields_ = [("n_docs", C.c_int64), ("text_bytes", C.c_int64), ("first_bad", C.c_int64), ("first_bad_reason", C.c_int32),
                ("n_names", C.c_uint32), ("n_creators", C.c_uint32), ("n_assignees", C.c_uint32),
                ("parse_ms", C.c_float), ("encode_ms", C.c_float), ("launches", C.c_int32)]


class PushTarget(C.Structure):
    """tt_push_target: where tt_sweep_push stores the compact hit list (device addresses)."""
    _fields_ = [("off16", C.c_void_p), ("tile_count", C.c_void_p), ("count", C.c_void_p)]


class PeerHandle(C.Structure):
    _fields_ = [("bytes", C.c_uint8 * 64)]


def lib_path():
    return os.path.join(HERE, "libtasktable.so")


_lib = None


def lib():
    """Loads libtasktable.so (RTLD_LOCAL).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise TaskTableError(TT_ERR_STATE, f"{path} is missing: run `make` (or __graft_entry__.build()); "
                                           "there is no CPU fallback")
    L = C.CDLL(path)
    vp, i32, i64, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64
    P = C.POINTER
    L.tt_last_error.restype = C.c_char_p
    L.tt_last_error.argtypes = []
    L.tt_create.argtypes = [i32, P(vp)]
    L.tt_destroy.argtypes = [vp]
    L.tt_generate.argtypes = [vp, P(SynthCfg), u64, u64]
    L.tt_load_soa.argtypes = [vp, P(SoaView)]
    L.tt_row_count.argtypes = [vp, P(i64)]
    L.tt_get_yesterdays_due.argtypes = [vp, i32, i64, P(vp)]
    L.tt_mark_overdue.argtypes = [vp, i32, vp, i64, P(i64), P(i64)]
    L.tt_sweep.argtypes = [vp, i32, i64, i64, P(vp), P(i64)]
    L.tt_get_by_creator.argtypes = [vp, C.c_char_p, i32, P(vp)]
    L.tt_get_by_creators.argtypes = [vp, P(C.c_char_p), P(i32), i32, P(vp)]
    L.tt_last_op_stats.argtypes = [vp, P(OpStats)]
    L.tt_creator_combining_stats.argtypes = [vp, P(i64), P(i64)]
    L.tt_last_order_info.argtypes = [vp, P(C.c_int32), P(C.c_int32), P(i64), P(C.c_int32)]
    L.tt_get_by_id.argtypes = [vp, vp, P(vp)]
    L.tt_insert_task.argtypes = [vp, vp, u64, u64, C.c_char_p, i32, P(i64)]
    L.tt_update_task.argtypes = [vp, vp, u64, P(i32)]
    L.tt_upsert_task.argtypes = [vp, vp, u64, u64, C.c_char_p, i32, i32, i32, P(i64), P(i32)]
    L.tt_mark_completed.argtypes = [vp, vp, P(i32)]
    L.tt_delete_task.argtypes = [vp, vp, P(i32)]
    L.tt_live_count.argtypes = [vp, P(i64)]
    L.tt_pack_hits.argtypes = [vp, vp, i64, vp, vp, P(i64), vp]
    L.tt_expand_hits.argtypes = [i32, vp, i64, vp, vp, vp, vp]
    L.tt_save_snapshot.argtypes = [vp, C.c_char_p]
    L.tt_load_snapshot.argtypes = [vp, C.c_char_p]
    L.tt_result_count.argtypes = [vp, P(i64)]
    L.tt_result_rows.argtypes = [vp, vp]
    L.tt_result_ids.argtypes = [vp, vp]
    L.tt_result_field.argtypes = [vp, i32, vp]
    L.tt_result_free.argtypes = [vp]
    L.tt_snapshot_flags.argtypes = [vp, vp, vp]
    L.tt_restore_flags.argtypes = [vp, vp, vp]
    L.tt_count_flags.argtypes = [vp, P(i64), P(i64)]
    L.tt_result_digest.argtypes = [vp, P(u64), P(u64), P(u64)]
    L.tt_sweep_device.argtypes = [vp, i32, i32, i64, i64, i64, vp, vp, i64, P(SweepStats), vp]
    L.tt_sweep_wait.argtypes = [vp, P(SweepStats)]
    L.tt_get_device_view.argtypes = [vp, P(DeviceView)]
    L.tt_sweep_static.argtypes = [P(DeviceView), i32, i32, i64, i64, vp, vp, i64, P(SweepStats), vp]
    L.tt_sweep_push.argtypes = [vp, i32, i32, i64, i64, vp, vp, i64, P(PushTarget), P(SweepStats), vp]
    L.tt_peer_alloc.argtypes = [i32, u64, P(vp), P(PeerHandle)]
    L.tt_peer_open.argtypes = [i32, P(PeerHandle), P(vp)]
    L.tt_peer_close.argtypes = [i32, vp]
    L.tt_peer_free.argtypes = [i32, vp]
    L.tt_load_payload.argtypes = [vp, P(StringsView), P(StringsView)]
    L.tt_generate_payload.argtypes = [vp, P(SynthCfg), u64]
    L.tt_result_json.argtypes = [vp, i32, i32, P(vp)]
    L.tt_json_size.argtypes = [vp, P(i64)]
    L.tt_json_copy.argtypes = [vp, vp]
    L.tt_json_device.argtypes = [vp, P(vp)]
    L.tt_json_free.argtypes = [vp]
    L.tt_last_json_stats.argtypes = [vp, P(JsonStats)]
    L.tt_get_due_on_or_before.argtypes = [vp, i32, i64, P(vp)]
    L.tt_sweep_backlog.argtypes = [vp, i32, i64, i64, P(vp), P(i64)]
    L.tt_tick_host.argtypes = [vp, i32, i32, i64, i64, vp, vp, vp, vp, i64, i32, P(SweepStats)]
    L.tt_set_task_strings.argtypes = [vp, vp, C.c_char_p, i32, C.c_char_p, i32, P(i32)]
    L.tt_json_offsets.argtypes = [vp, vp]
    L.tt_load_documents.argtypes = [vp, vp, vp, i64, i32, P(IngestStats)]
    L.tt_exchange_create_root.argtypes = [i32, i32, i64, i64, i32, P(vp), P(PeerHandle)]
    L.tt_exchange_attach.argtypes = [i32, i32, i32, i64, i64, i32, P(PeerHandle), P(vp)]
    L.tt_exchange_attach_local.argtypes = [vp, i32, i32, P(vp)]
    L.tt_exchange_destroy.argtypes = [vp]
    L.tt_exchange_slot_bytes.argtypes = [vp, P(u64)]
    L.tt_exchange_tick.argtypes = [vp, vp, i32, i32, i64, i64, vp, vp, P(SweepStats), vp]
    L.tt_exchange_epoch.argtypes = [vp, P(u64)]
    L.tt_exchange_collect.argtypes = [vp, i32, P(u64), u64, P(vp)]
    L.tt_exchange_release.argtypes = [vp, u64]
    L.tt_exchange_last_collect_ms.argtypes = [vp, P(C.c_float)]
    L.tt_gresult_count.argtypes = [vp, P(i64)]
    L.tt_gresult_shard_counts.argtypes = [vp, P(i64), i32]
    L.tt_gresult_rows.argtypes = [vp, vp]
    L.tt_gresult_ids.argtypes = [vp, vp]
    L.tt_gresult_digest.argtypes = [vp, P(u64), P(u64), P(u64)]
    L.tt_gresult_free.argtypes = [vp]
    L.tt_shards_create.argtypes = [P(i32), i32, P(vp)]
    L.tt_shards_destroy.argtypes = [vp]
    L.tt_shards_count.argtypes = [vp, P(i32)]
    L.tt_shards_table.argtypes = [vp, i32, P(vp), P(u64)]
    L.tt_shards_generate.argtypes = [vp, P(SynthCfg), u64]
    L.tt_shards_generate_payload.argtypes = [vp, P(SynthCfg), u64]
    L.tt_shards_load_soa.argtypes = [vp, P(SoaView)]
    L.tt_shards_row_count.argtypes = [vp, P(i64)]
    L.tt_shards_get_yesterdays_due.argtypes = [vp, i32, i64, P(vp)]
    L.tt_shards_get_due_on_or_before.argtypes = [vp, i32, i64, P(vp)]
    L.tt_shards_sweep.argtypes = [vp, i32, i32, i64, i64, P(vp), P(i64)]
    L.tt_shards_mark_overdue.argtypes = [vp, i32, vp, i64, P(i64), P(i64)]
    L.tt_shards_get_by_creator.argtypes = [vp, C.c_char_p, i32, P(vp)]
    L.tt_shards_get_by_id.argtypes = [vp, vp, P(vp)]
    L.tt_shards_insert_task.argtypes = [vp, vp, u64, u64, C.c_char_p, i32, P(i32), P(i64)]
    L.tt_shards_update_task.argtypes = [vp, vp, u64, P(i32)]
    L.tt_shards_mark_completed.argtypes = [vp, vp, P(i32)]
    L.tt_shards_delete_task.argtypes = [vp, vp, P(i32)]
    for name in EXPORTS:
        if name != "tt_last_error":
            getattr(L, name).restype = i32
    _lib = L
    return L


def _check(rc, allow=()):
    if rc == TT_OK or rc in allow:
        return rc
    text = lib().tt_last_error().decode(errors="replace")
    if rc == TT_ERR_NOT_IMPLEMENTED:
        raise NotImplementedInReference(rc, text)
    raise TaskTableError(rc, text)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Result:
    """tt_result: the hit list of one query, resident on the device until fetched."""

    def __init__(self, handle):
        self._h = handle
        n = C.c_int64(0)
        _check(lib().tt_result_count(self._h, C.byref(n)))
        self.count = n.value

    def rows(self):
        out = np.empty(self.count, dtype=np.uint32)
        _check(lib().tt_result_rows(self._h, _ptr(out)))
        return out

    def ids(self):
        out = np.empty((self.count, 16), dtype=np.uint8)
        _check(lib().tt_result_ids(self._h, _ptr(out)))
        return out

    def field(self, field):
        dt = {FIELD_CREATED_RAW: np.uint64, FIELD_DUE_RAW: np.uint64, FIELD_COMPLETED: np.uint8,
              FIELD_OVERDUE: np.uint8, FIELD_CREATOR_REF: np.uint32}[field]
        out = np.empty(self.count, dtype=dt)
        _check(lib().tt_result_field(self._h, field, _ptr(out)))
        return out

    def json(self, flavor=JSON_RESPONSE, local_offset_min=0):
        """tt_result_json: the rows of this result as the JSON array the reference's controllers
        return.  -> JsonText (the text stays on the device until .bytes() is called)."""
        h = C.c_void_p(0)
        _check(lib().tt_result_json(self._h, flavor, local_offset_min, C.byref(h)))
        return JsonText(h)

    def to_json(self, flavor=JSON_RESPONSE, local_offset_min=0):
        j = self.json(flavor, local_offset_min)
        try:
            return j.bytes()
        finally:
            j.free()

    def digest(self):
        hi, lo, s = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        _check(lib().tt_result_digest(self._h, C.byref(hi), C.byref(lo), C.byref(s)))
        return hi.value, lo.value, s.value

    def free(self):
        if self._h:
            lib().tt_result_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class JsonText:
    """tt_json: a JSON text resident on the device."""

    def __init__(self, handle):
        self._h = handle
        n = C.c_int64(0)
        _check(lib().tt_json_size(self._h, C.byref(n)))
        self.size = n.value

    def bytes(self):
        out = np.empty(self.size, dtype=np.uint8)
        _check(lib().tt_json_copy(self._h, _ptr(out)))
        return out.tobytes()

    def copy_to(self, dst_ptr):
        """dst_ptr: address of host (e.g. pinned) or device memory with room for .size bytes."""
        _check(lib().tt_json_copy(self._h, dst_ptr))

    def device_ptr(self):
        p = C.c_void_p(0)
        _check(lib().tt_json_device(self._h, C.byref(p)))
        return p.value

    def offsets_to(self, dst_ptr):
        """tt_json_offsets into caller memory (host or device address): records + 1 uint64 values."""
        _check(lib().tt_json_offsets(self._h, dst_ptr))

    def offsets(self, records):
        """Where every record's piece of the text starts (uint64[records + 1], the last = .size)."""
        out = np.empty(records + 1, dtype=np.uint64)
        _check(lib().tt_json_offsets(self._h, _ptr(out)))
        return out

    def free(self):
        if self._h:
            lib().tt_json_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def encode_strings(values):
    """Dictionary-encodes one bytes value per row -> (ref uint32[n], off uint32[k+1], pool uint8[]), the
    arrays a tt_strings_view points at."""
    index, ref = {}, np.empty(len(values), dtype=np.uint32)
    for i, v in enumerate(values):
        ref[i] = index.setdefault(bytes(v), len(index))
    keys = list(index.keys())
    off = np.zeros(len(keys) + 1, dtype=np.uint32)
    off[1:] = np.cumsum([len(k) for k in keys], dtype=np.uint64)
    pool = np.frombuffer(b"".join(keys) or b"\0", dtype=np.uint8).copy()
    return ref, off, pool


def peer_alloc(device, nbytes):
    """-> (device address, 64-byte handle) of zeroed device memory other processes can map."""
    p, h = C.c_void_p(0), PeerHandle()
    _check(lib().tt_peer_alloc(device, nbytes, C.byref(p), C.byref(h)))
    return p.value, bytes(h.bytes)


def peer_open(device, handle_bytes):
    """Maps memory another process allocated with peer_alloc; -> device address valid on `device`."""
    h = PeerHandle()
    C.memmove(h.bytes, handle_bytes, 64)
    p = C.c_void_p(0)
    _check(lib().tt_peer_open(device, C.byref(h), C.byref(p)))
    return p.value


def peer_close(device, ptr):
    _check(lib().tt_peer_close(device, ptr))


def peer_free(device, ptr):
    _check(lib().tt_peer_free(device, ptr))


def expand_hits(device, d_tile_count, n_tiles, d_off16, d_scratch, d_rows_out, stream=None):
    """Inverse of TaskTable.pack_hits on any device (device addresses)."""
    _check(lib().tt_expand_hits(device, d_tile_count, n_tiles, d_off16, d_scratch, d_rows_out, stream))


class TaskTable:
    """tt_table: one shard of the task table, resident in the HBM of one GPU."""

    def __init__(self, device=0):
        h = C.c_void_p()
        _check(lib().tt_create(device, C.byref(h)))
        self._h = h
        self.device = device
        self._owned = True

    @classmethod
    def _borrowed(cls, handle, device):
        """A table owned by someone else (a shard of a ShardedTable): close() does not destroy it."""
        t = cls.__new__(cls)
        t._h, t.device, t._owned = handle, device, False
        return t

    # -- filling ---------------------------------------------------------------------------
    def generate(self, cfg, first_row, n_rows):
        if C.sizeof(cfg) != C.sizeof(SynthCfg):
            raise TypeError("cfg must have the layout of tt_synth_cfg")
        _check(lib().tt_generate(self._h, C.cast(C.byref(cfg), C.POINTER(SynthCfg)), first_row, n_rows))
        return self

    def load_soa(self, task_id, created_raw, due_raw, completed, overdue, creator_ref, creators):
        """creators: list of bytes (the distinct TaskCreatedBy strings, UTF-8)."""
        n = len(created_raw)
        ids = np.ascontiguousarray(task_id, dtype=np.uint8).reshape(n, 16) if n else np.zeros((0, 16), np.uint8)
        cr = np.ascontiguousarray(created_raw, dtype=np.uint64)
        du = np.ascontiguousarray(due_raw, dtype=np.uint64)
        co = np.ascontiguousarray(completed, dtype=np.uint8)
        ov = np.ascontiguousarray(overdue, dtype=np.uint8)
        ref = np.ascontiguousarray(creator_ref, dtype=np.uint32)
        off = np.zeros(len(creators) + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(c) for c in creators], dtype=np.uint64) if creators else 0
        pool = np.frombuffer(b"".join(creators) or b"\0", dtype=np.uint8).copy()
        v = SoaView(n, _ptr(ids), _ptr(cr), _ptr(du), _ptr(co), _ptr(ov), _ptr(ref), len(creators), _ptr(off),
                    _ptr(pool))
        _check(lib().tt_load_soa(self._h, C.byref(v)))
        return self

    def load_payload(self, task_names=None, assigned_to=None):
        """TaskName / TaskAssignedTo, one bytes value (UTF-8) per row; either may be None."""
        keep, views = [], []
        for values in (task_names, assigned_to):
            if values is None:
                views.append(None)
                continue
            ref, off, pool = encode_strings(values)
            keep.append((ref, off, pool))
            views.append(StringsView(_ptr(ref), len(off) - 1, _ptr(off), _ptr(pool)))
        _check(lib().tt_load_payload(self._h, C.byref(views[0]) if views[0] else None,
                                     C.byref(views[1]) if views[1] else None))
        return self

    def load_documents(self, text, doc_off, local_offset_min=0, allow_bad=False):
        """tt_load_documents: fills this (empty) table from JSON documents.  text: bytes / uint8 array, or
        a device address (int); doc_off: n + 1 offsets (array-like), or (device address, n).  -> IngestStats (with allow_bad a rejected load
        returns its stats instead of raising)."""
        if isinstance(doc_off, tuple):  # (device address of uint64[n + 1], n): offsets already on the device
            off_ptr, n_docs = doc_off
            off = None
        else:
            off = np.ascontiguousarray(doc_off, dtype=np.uint64)
            off_ptr, n_docs = _ptr(off), off.size - 1
        keep = None
        if isinstance(text, int):
            ptr = text
        else:
            keep = np.frombuffer(bytes(text), dtype=np.uint8) if not isinstance(text, np.ndarray) else text
            ptr = _ptr(keep) if keep.size else None
        st = IngestStats()
        rc = lib().tt_load_documents(self._h, ptr, off_ptr, n_docs, local_offset_min, C.byref(st))
        if not (allow_bad and rc == TT_ERR_ARG and st.first_bad >= 0):
            _check(rc)
        return st

    def generate_payload(self, cfg, first_row):
        """Synthetic TaskName / TaskAssignedTo for the rows generate() made (same cfg, same first_row)."""
        if C.sizeof(cfg) != C.sizeof(SynthCfg):
            raise TypeError("cfg must have the layout of tt_synth_cfg")
        _check(lib().tt_generate_payload(self._h, C.cast(C.byref(cfg), C.POINTER(SynthCfg)), first_row))
        return self

    def last_json_stats(self):
        st = JsonStats()
        _check(lib().tt_last_json_stats(self._h, C.byref(st)))
        return st

    def __len__(self):
        n = C.c_int64(0)
        _check(lib().tt_row_count(self._h, C.byref(n)))
        return n.value

    # -- ITasksManager operations ----------------------------------------------------------
    def get_yesterdays_due(self, mode, today=DEFAULT_TODAY):
        r = C.c_void_p()
        _check(lib().tt_get_yesterdays_due(self._h, mode, today, C.byref(r)))
        return Result(r)

    def get_due_on_or_before(self, mode, today=DEFAULT_TODAY):
        """tt_get_due_on_or_before: GetYesterdaysDueTasks with the generalised predicate (an extension)."""
        h = C.c_void_p(0)
        _check(lib().tt_get_due_on_or_before(self._h, mode, today, C.byref(h)))
        return Result(h)

    def sweep_backlog(self, mode, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT):
        """tt_sweep_backlog: the fused cron tick with the generalised predicate -> (Result, n_marked, threw)."""
        h, marked = C.c_void_p(0), C.c_int64(0)
        rc = lib().tt_sweep_backlog(self._h, mode, today, run_at, C.byref(h), C.byref(marked))
        _check(rc, allow=(TT_ERR_NOT_IMPLEMENTED,))
        return Result(h), marked.value, rc == TT_ERR_NOT_IMPLEMENTED

    def mark_overdue(self, mode, ids16):
        ids = np.ascontiguousarray(ids16, dtype=np.uint8).reshape(-1, 16)
        marked, missing = C.c_int64(0), C.c_int64(0)
        _check(lib().tt_mark_overdue(self._h, mode, _ptr(ids), ids.shape[0], C.byref(marked), C.byref(missing)))
        return marked.value, missing.value

    def sweep(self, mode, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT):
        """-> (result, n_marked, threw) ; threw mirrors FAKE mode's NotImplementedException."""
        r = C.c_void_p()
        marked = C.c_int64(0)
        rc = _check(lib().tt_sweep(self._h, mode, today, run_at, C.byref(r), C.byref(marked)),
                    allow=(TT_ERR_NOT_IMPLEMENTED,))
        return Result(r), marked.value, rc == TT_ERR_NOT_IMPLEMENTED

    def get_by_creator(self, created_by):
        b = created_by.encode() if isinstance(created_by, str) else bytes(created_by)
        r = C.c_void_p()
        _check(lib().tt_get_by_creator(self._h, b, len(b), C.byref(r)))
        return Result(r)

    def get_by_creators(self, creators):
        """One table pass for a whole batch of creators -> list of Result, one per query."""
        bs = [c.encode() if isinstance(c, str) else bytes(c) for c in creators]
        n = len(bs)
        arr = (C.c_char_p * n)(*bs)
        lens = (C.c_int32 * n)(*[len(b) for b in bs])
        out = (C.c_void_p * n)()
        _check(lib().tt_get_by_creators(self._h, arr, lens, n, out))
        return [Result(C.c_void_p(out[i])) for i in range(n)]

    def creator_combining_stats(self):
        """-> (table passes, requests they answered) of the get_by_creator calls made so far."""
        p, r = C.c_int64(0), C.c_int64(0)
        _check(lib().tt_creator_combining_stats(self._h, C.byref(p), C.byref(r)))
        return p.value, r.value

    def last_order_info(self):
        """-> (key bits, radix passes, groups put right after the prefix passes, fell back) of the last ordered list."""
        b, p, f, fb = C.c_int32(0), C.c_int32(0), C.c_int64(0), C.c_int32(0)
        _check(lib().tt_last_order_info(self._h, C.byref(b), C.byref(p), C.byref(f), C.byref(fb)))
        return b.value, p.value, f.value, bool(fb.value)

    def last_op_stats(self):
        st = OpStats()
        _check(lib().tt_last_op_stats(self._h, C.byref(st)))
        return st

    # -- single-row operations (ITasksManager.cs:8-12) -----------------------------------------
    @staticmethod
    def _id(task_id):
        a = np.frombuffer(bytes(task_id), dtype=np.uint8).copy()
        assert a.size == 16
        return a

    def get_by_id(self, task_id):
        r = C.c_void_p()
        a = self._id(task_id)
        _check(lib().tt_get_by_id(self._h, _ptr(a), C.byref(r)))
        return Result(r)

    def insert_task(self, task_id, created_raw, due_raw, created_by):
        b = created_by.encode() if isinstance(created_by, str) else bytes(created_by)
        row = C.c_int64(-1)
        a = self._id(task_id)
        _check(lib().tt_insert_task(self._h, _ptr(a), created_raw, due_raw, b, len(b), C.byref(row)))
        return row.value

    def upsert_task(self, task_id, created_raw, due_raw, created_by, is_completed, is_overdue):
        """tt_upsert_task: the state store's SaveStateAsync of a whole record -> (row, existed)."""
        b = created_by.encode() if isinstance(created_by, str) else bytes(created_by)
        row, existed = C.c_int64(-1), C.c_int32(0)
        a = self._id(task_id)
        _check(lib().tt_upsert_task(self._h, _ptr(a), created_raw, due_raw, b, len(b), int(bool(is_completed)),
                                    int(bool(is_overdue)), C.byref(row), C.byref(existed)))
        return row.value, bool(existed.value)

    def update_task(self, task_id, due_raw):
        f = C.c_int32(0)
        a = self._id(task_id)
        _check(lib().tt_update_task(self._h, _ptr(a), due_raw, C.byref(f)))
        return bool(f.value)

    def set_task_strings(self, task_id, task_name=None, assigned_to=None):
        """TaskName / TaskAssignedTo of one task (bytes, UTF-8); None leaves that member as it is."""
        f = C.c_int32(0)
        a = self._id(task_id)
        _check(lib().tt_set_task_strings(self._h, _ptr(a), task_name, -1 if task_name is None else len(task_name),
                                         assigned_to, -1 if assigned_to is None else len(assigned_to), C.byref(f)))
        return bool(f.value)

    def mark_completed(self, task_id):
        f = C.c_int32(0)
        a = self._id(task_id)
        _check(lib().tt_mark_completed(self._h, _ptr(a), C.byref(f)))
        return bool(f.value)

    def delete_task(self, task_id):
        f = C.c_int32(0)
        a = self._id(task_id)
        _check(lib().tt_delete_task(self._h, _ptr(a), C.byref(f)))
        return bool(f.value)

    def live_count(self):
        n = C.c_int64(0)
        _check(lib().tt_live_count(self._h, C.byref(n)))
        return n.value

    # -- device-resident sweep ---------------------------------------------------------------
    def sweep_device(self, mode, flags, d_rows, d_ids, cap, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT, first_row=0,
                     stream=None, allow_capacity=False):
        """d_rows / d_ids: device addresses (int).  Returns (SweepStats, rc)."""
        st = SweepStats()
        rc = lib().tt_sweep_device(self._h, mode, flags, today, run_at, first_row, d_rows, d_ids, cap, C.byref(st),
                                   stream)
        _check(rc, allow=(TT_ERR_CAPACITY,) if allow_capacity else ())
        return st, rc

    def sweep_wait(self, allow_capacity=False):
        """tt_sweep_wait: completes a tick queued with SWEEP_ASYNC -> (SweepStats, rc)."""
        st = SweepStats()
        rc = lib().tt_sweep_wait(self._h, C.byref(st))
        _check(rc, allow=(TT_ERR_CAPACITY,) if allow_capacity else ())
        return st, rc

    def sweep_push(self, mode, flags, d_rows, d_ids, cap, off16, tile_count, count, today=DEFAULT_TODAY,
                   run_at=DEFAULT_RUN_AT, stream=None, allow_capacity=False):
        """tt_sweep_push: sweep_device whose placement kernel also stores the compact hit list
        (off16 / tile_count / count: device addresses, possibly peer-mapped memory of the root GPU).
        d_rows / d_ids may be None."""
        target = PushTarget(off16, tile_count, count)
        st = SweepStats()
        rc = lib().tt_sweep_push(self._h, mode, flags, today, run_at, d_rows, d_ids, cap, C.byref(target), C.byref(st),
                                 stream)
        _check(rc, allow=(TT_ERR_CAPACITY,) if allow_capacity else ())
        return st, rc

    def tick_host(self, mode, flags, h_completed, h_overdue, h_rows, h_ids, cap, chunks=0, today=DEFAULT_TODAY,
                  run_at=DEFAULT_RUN_AT, allow_capacity=False):
        """tt_tick_host: the tick with the flag planes coming from host memory and the hits going to host
        memory, transfers pipelined with the kernels.  All four buffers are host ADDRESSES (ints; pinned
        memory recommended), h_completed / h_overdue / h_ids may be None.  -> (SweepStats, rc)."""
        st = SweepStats()
        rc = lib().tt_tick_host(self._h, mode, flags, today, run_at, h_completed, h_overdue, h_rows, h_ids, cap, chunks,
                                C.byref(st))
        _check(rc, allow=(TT_ERR_CAPACITY,) if allow_capacity else ())
        return st, rc

    def sweep_static(self, mode, flags, d_rows, d_ids, cap, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT, stream=None,
                     allow_capacity=False):
        """tt_sweep_static: the sweep for callers that cannot allocate device memory.  It runs the
        same kernels as sweep_device over this table's columns, but keeps all working memory in
        the library's fixed ~4 MB device array, so the table must have at most STATIC_MAX_ROWS
        rows and cap at most STATIC_MAX_HITS (larger sizes raise TaskTableError, TT_ERR_ARG).
        Returns (SweepStats, rc) like sweep_device; the timing fields stay 0."""
        view = self.device_view()
        st = SweepStats()
        rc = lib().tt_sweep_static(C.byref(view), mode, flags, today, run_at, d_rows, d_ids, cap, C.byref(st), stream)
        _check(rc, allow=(TT_ERR_CAPACITY,) if allow_capacity else ())
        return st, rc

    # -- state -----------------------------------------------------------------------------------
    def snapshot_flags(self):
        nb = (len(self) + 7) // 8
        c = np.zeros(nb, dtype=np.uint8)
        o = np.zeros(nb, dtype=np.uint8)
        _check(lib().tt_snapshot_flags(self._h, _ptr(c), _ptr(o)))
        return c, o

    def flags_bytes(self):
        """(completed, overdue) as one byte per row."""
        n = len(self)
        c, o = self.snapshot_flags()
        return (np.unpackbits(c, bitorder="little")[:n], np.unpackbits(o, bitorder="little")[:n])

    def restore_flags(self, completed_bits=None, overdue_bits=None):
        _check(lib().tt_restore_flags(self._h, _ptr(completed_bits), _ptr(overdue_bits)))

    def count_flags(self):
        c, o = C.c_int64(0), C.c_int64(0)
        _check(lib().tt_count_flags(self._h, C.byref(c), C.byref(o)))
        return c.value, o.value

    def pack_hits(self, d_rows, n, d_off16, d_tile_count, stream=None):
        """Compact form of the hit list of the last sweep_device call: 2-byte in-tile offsets + per-tile
        counts (device addresses).  Returns the number of tiles."""
        nt = C.c_int64(0)
        _check(lib().tt_pack_hits(self._h, d_rows, n, d_off16, d_tile_count, C.byref(nt), stream))
        return nt.value

    def save_snapshot(self, path):
        _check(lib().tt_save_snapshot(self._h, os.fsencode(path)))

    def load_snapshot(self, path):
        _check(lib().tt_load_snapshot(self._h, os.fsencode(path)))
        return self

    def device_view(self):
        v = DeviceView()
        _check(lib().tt_get_device_view(self._h, C.byref(v)))
        return v

    def close(self):
        if self._h:
            if self._owned:
                lib().tt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GResult:
    """tt_gresult: a list over the whole sharded table — global row numbers and TaskIds."""

    def __init__(self, handle):
        self._h = handle
        n = C.c_int64(0)
        _check(lib().tt_gresult_count(self._h, C.byref(n)))
        self.count = n.value

    def shard_counts(self):
        out = (C.c_int64 * MAX_SHARDS)()
        _check(lib().tt_gresult_shard_counts(self._h, out, MAX_SHARDS))
        return list(out)

    def rows(self):
        out = np.empty(self.count, dtype=np.uint64)
        _check(lib().tt_gresult_rows(self._h, _ptr(out)))
        return out

    def ids(self):
        out = np.empty((self.count, 16), dtype=np.uint8)
        _check(lib().tt_gresult_ids(self._h, _ptr(out)))
        return out

    def digest(self):
        hi, lo, s = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
        _check(lib().tt_gresult_digest(self._h, C.byref(hi), C.byref(lo), C.byref(s)))
        return hi.value, lo.value, s.value

    def free(self):
        if self._h:
            lib().tt_gresult_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Exchange:
    """tt_exchange: one shard's end of the cross-shard exchange (include/tasktable.h).  The root's owns the
    receive buffer; the others map it — from another process through the 64-byte handle, inside one
    process with attach_local."""

    def __init__(self, handle, device, rank, world, root):
        self._h, self.device, self.rank, self.world, self.root = handle, device, rank, world, root

    @classmethod
    def create_root(cls, device, world, cap_per_shard, max_tiles_per_shard, flags=0):
        """-> (Exchange, handle bytes for the other processes)."""
        h, ph = C.c_void_p(0), PeerHandle()
        _check(lib().tt_exchange_create_root(device, world, cap_per_shard, max_tiles_per_shard, flags, C.byref(h), C.byref(ph)))
        return cls(h, device, 0, world, True), bytes(ph.bytes)

    @classmethod
    def attach(cls, device, rank, world, cap_per_shard, max_tiles_per_shard, flags, handle_bytes):
        ph = PeerHandle()
        C.memmove(ph.bytes, handle_bytes, 64)
        h = C.c_void_p(0)
        _check(lib().tt_exchange_attach(device, rank, world, cap_per_shard, max_tiles_per_shard, flags, C.byref(ph), C.byref(h)))
        return cls(h, device, rank, world, False)

    def attach_local(self, device, rank):
        h = C.c_void_p(0)
        _check(lib().tt_exchange_attach_local(self._h, device, rank, C.byref(h)))
        return Exchange(h, device, rank, self.world, False)

    @property
    def epoch(self):
        e = C.c_uint64(0)
        _check(lib().tt_exchange_epoch(self._h, C.byref(e)))
        return e.value

    @property
    def slot_bytes(self):
        b = C.c_uint64(0)
        _check(lib().tt_exchange_slot_bytes(self._h, C.byref(b)))
        return b.value

    def tick(self, table, mode, flags, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT, stream=None, allow_capacity=False,
             d_rows=None, d_ids=None):
        """This shard's cron tick; its result lands in its slot of the root's buffer.  -> (SweepStats, rc).
        With SWEEP_ASYNC in flags the stats come from table.sweep_wait().  d_rows / d_ids: device addresses
        for the shard's own copy of its list (optional)."""
        st = SweepStats()
        rc = lib().tt_exchange_tick(self._h, table._h, mode, flags, today, run_at, d_rows, d_ids, C.byref(st), stream)
        _check(rc, allow=(TT_ERR_CAPACITY,) if allow_capacity else ())
        return st, rc

    def collect(self, order, first_rows, expected_epoch=None):
        """root: waits (on the device) until every shard has published `expected_epoch` (default: this
        shard's own epoch), -> GResult."""
        fr = (C.c_uint64 * self.world)(*[int(x) for x in first_rows])
        h = C.c_void_p(0)
        _check(lib().tt_exchange_collect(self._h, order, fr, self.epoch if expected_epoch is None else expected_epoch, C.byref(h)))
        return GResult(h)

    def release(self, epoch=None):
        _check(lib().tt_exchange_release(self._h, self.epoch if epoch is None else epoch))

    def last_collect_ms(self):
        ms = C.c_float(0)
        _check(lib().tt_exchange_last_collect_ms(self._h, C.byref(ms)))
        return ms.value

    def close(self):
        if self._h:
            lib().tt_exchange_destroy(self._h)
            self._h = None


class ShardedTable:
    """tt_shards: the task table split by TaskId range over several GPUs, driven from one process, behind
    the same operations as TaskTable (the reference registers one ITasksManager, Api/Program.cs:13).
    devices: one CUDA ordinal per shard (a device may repeat)."""

    def __init__(self, devices):
        arr = (C.c_int32 * len(devices))(*devices)
        h = C.c_void_p(0)
        _check(lib().tt_shards_create(arr, len(devices), C.byref(h)))
        self._h, self.devices = h, list(devices)

    def generate(self, cfg, n_total):
        _check(lib().tt_shards_generate(self._h, C.cast(C.byref(cfg), C.POINTER(SynthCfg)), n_total))
        return self

    def load_soa(self, task_id, created_raw, due_raw, completed, overdue, creator_ref, creators):
        n = len(created_raw)
        ids = np.ascontiguousarray(task_id, dtype=np.uint8).reshape(n, 16) if n else np.zeros((0, 16), np.uint8)
        cr = np.ascontiguousarray(created_raw, dtype=np.uint64)
        du = np.ascontiguousarray(due_raw, dtype=np.uint64)
        co = np.ascontiguousarray(completed, dtype=np.uint8)
        ov = np.ascontiguousarray(overdue, dtype=np.uint8)
        ref = np.ascontiguousarray(creator_ref, dtype=np.uint32)
        off = np.zeros(len(creators) + 1, dtype=np.uint32)
        off[1:] = np.cumsum([len(c) for c in creators], dtype=np.uint64) if creators else 0
        pool = np.frombuffer(b"".join(creators) or b"\0", dtype=np.uint8).copy()
        v = SoaView(n, _ptr(ids), _ptr(cr), _ptr(du), _ptr(co), _ptr(ov), _ptr(ref), len(creators), _ptr(off), _ptr(pool))
        _check(lib().tt_shards_load_soa(self._h, C.byref(v)))
        return self

    def __len__(self):
        n = C.c_int64(0)
        _check(lib().tt_shards_row_count(self._h, C.byref(n)))
        return n.value

    def shard(self, g):
        """-> (TaskTable of shard g, still owned by this object; global row of its first row)."""
        h, first = C.c_void_p(0), C.c_uint64(0)
        _check(lib().tt_shards_table(self._h, g, C.byref(h), C.byref(first)))
        return TaskTable._borrowed(h, self.devices[g]), first.value

    def get_yesterdays_due(self, mode, today=DEFAULT_TODAY):
        h = C.c_void_p(0)
        _check(lib().tt_shards_get_yesterdays_due(self._h, mode, today, C.byref(h)))
        return GResult(h)

    def get_due_on_or_before(self, mode, today=DEFAULT_TODAY):
        h = C.c_void_p(0)
        _check(lib().tt_shards_get_due_on_or_before(self._h, mode, today, C.byref(h)))
        return GResult(h)

    def sweep(self, mode, flags=0, today=DEFAULT_TODAY, run_at=DEFAULT_RUN_AT):
        """-> (GResult in TaskId-range order, n_marked, threw)."""
        h, marked = C.c_void_p(0), C.c_int64(0)
        rc = _check(lib().tt_shards_sweep(self._h, mode, flags, today, run_at, C.byref(h), C.byref(marked)),
                    allow=(TT_ERR_NOT_IMPLEMENTED,))
        return GResult(h), marked.value, rc == TT_ERR_NOT_IMPLEMENTED

    def mark_overdue(self, mode, ids16):
        ids = np.ascontiguousarray(ids16, dtype=np.uint8).reshape(-1, 16)
        marked, missing = C.c_int64(0), C.c_int64(0)
        _check(lib().tt_shards_mark_overdue(self._h, mode, _ptr(ids), ids.shape[0], C.byref(marked), C.byref(missing)))
        return marked.value, missing.value

    def get_by_creator(self, created_by):
        b = created_by.encode() if isinstance(created_by, str) else bytes(created_by)
        h = C.c_void_p(0)
        _check(lib().tt_shards_get_by_creator(self._h, b, len(b), C.byref(h)))
        return GResult(h)

    def get_by_id(self, task_id):
        a = TaskTable._id(task_id)
        h = C.c_void_p(0)
        _check(lib().tt_shards_get_by_id(self._h, _ptr(a), C.byref(h)))
        return GResult(h)

    def insert_task(self, task_id, created_raw, due_raw, created_by):
        """-> (shard, row inside that shard)."""
        b = created_by.encode() if isinstance(created_by, str) else bytes(created_by)
        a = TaskTable._id(task_id)
        g, row = C.c_int32(-1), C.c_int64(-1)
        _check(lib().tt_shards_insert_task(self._h, _ptr(a), created_raw, due_raw, b, len(b), C.byref(g), C.byref(row)))
        return g.value, row.value

    def _point(self, fn, task_id, *args):
        f = C.c_int32(0)
        a = TaskTable._id(task_id)
        _check(fn(self._h, _ptr(a), *args, C.byref(f)))
        return bool(f.value)

    def update_task(self, task_id, due_raw):
        return self._point(lib().tt_shards_update_task, task_id, due_raw)

    def mark_completed(self, task_id):
        return self._point(lib().tt_shards_mark_completed, task_id)

    def delete_task(self, task_id):
        return self._point(lib().tt_shards_delete_task, task_id)

    def close(self):
        if self._h:
            lib().tt_shards_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
