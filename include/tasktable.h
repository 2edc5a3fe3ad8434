/* tasktable.h — C ABI of the B200-native task table (libtasktable.so).
 *
 * This is the drop-in boundary for the one data-parallel path of
 * chsakell/aca-dotnet-workshop: the cron-driven overdue-task sweep and the creator filter
 * behind ITasksManager (Api/Services/ITasksManager.cs:5-15; "Api" =
 * TasksTracker.TasksManager.Backend.Api, "Svc" = TasksTracker.Processor.Backend.Svc under
 * /root/reference).  A C# ITasksManager implementation P/Invokes these entry points
 * (INTEGRATION.md shows the [DllImport] stubs); every signature uses plain pointers and
 * sizes only.  All functions return an int32 status (TT_OK = 0, negative = error) and never
 * throw; tt_last_error() gives the text for the calling thread.
 *
 * The table lives in HBM as a structure of arrays:
 *   hot  : due_code u32/row (day number + "time of day != 0" bit + DateTimeKind),
 *          completed and overdue bit-planes (1 bit/row each), creator_hash u64/row
 *   cold : task_id 16 B/row, created_raw and due_raw u64/row (.NET dateData), creator_ref
 *          u32/row into a creator dictionary (offset table + UTF-8 pool)
 * Rows keep the caller's order (the reference's list order); row numbers are shard-local.
 *
 * Thread safety: one tt_table may be shared by all request threads (the reference registers
 * its manager as a singleton, Api/Program.cs:13); calls on the same table are serialised by
 * an internal lock, except that concurrent tt_get_by_creator calls are COMBINED into shared table
 * passes (tt_creator_combining_stats).  tt_result objects are independent of the lock once returned.
 *
 * There is no CPU fallback: every operation runs as CUDA kernels on the table's device and
 * fails with TT_ERR_CUDA if no device is usable.
 */
#ifndef TASKTABLE_H
#define TASKTABLE_H

#include <stdint.h>

#include "tt_synth.h"

#ifdef __cplusplus
extern "C" {
#endif

#define TT_OK 0
#define TT_ERR_ARG (-1)             /* bad argument (null, negative size, today not a midnight, ...) */
#define TT_ERR_CUDA (-2)            /* CUDA runtime / launch failure, text in tt_last_error()        */
#define TT_ERR_OOM (-3)             /* device or pinned-host allocation failed                       */
#define TT_ERR_NOT_IMPLEMENTED (-4) /* FAKE-mode MarkOverdueTasks: Api/Services/FakeTasksManager.cs:102-105 throws */
#define TT_ERR_CAPACITY (-5)        /* caller-provided output buffer too small; call is resumable    */
#define TT_ERR_STATE (-6)           /* table empty / already loaded / result of another table        */

/* Which reference implementation's semantics to follow (they differ, SURVEY.md §0 finding 1):
 * FAKE  = Api/Services/FakeTasksManager.cs:107-112  tick equality only, source order, Mark throws
 * STORE = Api/Services/TasksStoreManager.cs:104-149 exact "yyyy-MM-ddT00:00:00" match, pending
 *         rows only, ordered by TaskCreatedOn, Mark sets IsOverDue                              */
#define TT_MODE_FAKE 0
#define TT_MODE_STORE 1

typedef struct tt_table tt_table;
typedef struct tt_result tt_result;

/* Host-side structure-of-arrays view handed to tt_load_soa.  One entry per TaskModel
 * (Api/Models/TaskModel.cs:3-13); TaskName / TaskAssignedTo are not touched by any kernel and
 * stay with the host wrapper. */
typedef struct tt_soa_view {
  int64_t n;                   /* rows                                                          */
  const uint8_t* task_id;      /* 16*n bytes, canonical big-endian order of Guid.ToString("D")  */
  const uint64_t* created_raw; /* TaskCreatedOn as .NET dateData (ticks | kind<<62)             */
  const uint64_t* due_raw;     /* TaskDueDate  as .NET dateData                                 */
  const uint8_t* completed;    /* IsCompleted, one byte per row (0/1)                           */
  const uint8_t* overdue;      /* IsOverDue,   one byte per row (0/1)                           */
  const uint32_t* creator_ref; /* index of the row's TaskCreatedBy in the dictionary below      */
  uint32_t n_creators;         /* dictionary entries                                            */
  const uint32_t* creator_off; /* n_creators+1 byte offsets into creator_pool                   */
  const uint8_t* creator_pool; /* UTF-8 bytes of the distinct TaskCreatedBy strings             */
} tt_soa_view;

/* Fields tt_result_field can gather for the rows of a result. */
#define TT_FIELD_CREATED_RAW 1 /* uint64 per row */
#define TT_FIELD_DUE_RAW 2     /* uint64 per row */
#define TT_FIELD_COMPLETED 3   /* uint8  per row */
#define TT_FIELD_OVERDUE 4     /* uint8  per row (state when the result was produced is NOT kept: current state) */
#define TT_FIELD_CREATOR_REF 5 /* uint32 per row */

/* ---- lifetime -------------------------------------------------------------------------- */

/* Replaces the reference's `new FakeTasksManager()` / DI construction (Api/Program.cs:13).
 * `device` is the CUDA ordinal that will hold this table (one table per GPU shard). */
int32_t tt_create(int32_t device, tt_table** out);
int32_t tt_destroy(tt_table* t);

/* Text of the last error raised on the calling thread (never NULL). */
const char* tt_last_error(void);

/* Fills the table on the device with rows [first_row, first_row + n_rows) of the synthetic
 * table described by cfg (include/tt_synth.h) — the stand-in for
 * FakeTasksManager.GenerateRandomTasks (Api/Services/FakeTasksManager.cs:10-25) at benchmark
 * sizes.  first_row/n_rows select this shard's TaskId range. */
int32_t tt_generate(tt_table* t, const tt_synth_cfg* cfg, uint64_t first_row, uint64_t n_rows);

/* Loads caller rows (the reference's _tasksList / state-store documents) into HBM. */
int32_t tt_load_soa(tt_table* t, const tt_soa_view* v);

int32_t tt_row_count(tt_table* t, int64_t* n);

/* ---- the ITasksManager operations on the path ------------------------------------------- */

/* ITasksManager.GetYesterdaysDueTasks (Api/Services/ITasksManager.cs:14);
 * FAKE: FakeTasksManager.cs:107-112 (source order), STORE: TasksStoreManager.cs:104-139
 * (pending rows, ascending TaskCreatedOn, ties in row order).  `today_local_ticks` is
 * DateTime.Today (local midnight) in ticks.  Does not modify the table. */
int32_t tt_get_yesterdays_due(tt_table* t, int32_t mode, int64_t today_local_ticks, tt_result** out);

/* ITasksManager.MarkOverdueTasks (ITasksManager.cs:13; TasksStoreManager.cs:141-149): sets
 * IsOverDue for every listed TaskId that is in the table.  ids16 = 16*n bytes in canonical
 * order.  FAKE mode returns TT_ERR_NOT_IMPLEMENTED like the reference.  *n_missing counts ids
 * with no live row: the reference's SaveStateAsync would create a document for them, which needs
 * the whole record — call tt_upsert_task for those (the host mirrors do).  The table need not be
 * in TaskId order: the TaskId index is rebuilt on the device when inserts have invalidated it. */
int32_t tt_mark_overdue(tt_table* t, int32_t mode, const uint8_t* ids16, int64_t n, int64_t* n_marked,
                        int64_t* n_missing);

/* The whole cron tick in one device pass — replaces the three-process round trip of
 * Svc/Controllers/ScheduledTasksManagerController.cs:19-46: GetYesterdaysDueTasks, keep
 * rows with runAt.Date > TaskDueDate.Date (:34), MarkOverdueTasks if any survive (:40-45).
 * *out lists what GetYesterdaysDueTasks returned (row order); *n_marked = rows whose
 * IsOverDue was set.  FAKE mode with survivors returns TT_ERR_NOT_IMPLEMENTED (the result is
 * still produced, the table is unchanged), as the reference's cron handler would observe. */
int32_t tt_sweep(tt_table* t, int32_t mode, int64_t today_local_ticks, int64_t run_at_utc_ticks, tt_result** out,
                 int64_t* n_marked);

/* The two calls above with the generalised predicate TT_SWEEP_DUE_ON_OR_BEFORE (see below): what
 * GetYesterdaysDueTasks / the cron tick would be if they asked for "due yesterday or earlier and
 * still pending".  No reference counterpart; same modes, order and errors. */
int32_t tt_get_due_on_or_before(tt_table* t, int32_t mode, int64_t today_local_ticks, tt_result** out);
int32_t tt_sweep_backlog(tt_table* t, int32_t mode, int64_t today_local_ticks, int64_t run_at_utc_ticks, tt_result** out,
                         int64_t* n_marked);

/* ITasksManager.GetTasksByCreator (ITasksManager.cs:7; FakeTasksManager.cs:68-72,
 * TasksStoreManager.cs:54-69): ordinal byte-exact match of TaskCreatedBy (64-bit hash scan,
 * then the candidate's string is compared), descending TaskCreatedOn, ties in row order. */
int32_t tt_get_by_creator(tt_table* t, const char* utf8, int32_t len, tt_result** out);
/* The reference's manager is a singleton shared by every ASP.NET request thread (Api/Program.cs:13), so
 * tt_get_by_creator calls arrive concurrently.  They do not queue for one table pass each: calls that
 * arrive while a pass is running are combined, and the next pass answers all of them at once (the batch
 * kernel of tt_get_by_creators reads the hash column once for up to 1024 creators).  These counters say
 * how many passes answered how many requests.  No reference counterpart. */
int32_t tt_creator_combining_stats(tt_table* t, int64_t* passes, int64_t* requests);
/* How the last OrderBy(TaskCreatedOn) of this table (Api/Services/TasksStoreManager.cs:136; the ordered
 * tt_get_yesterdays_due / tt_get_due_on_or_before) was sorted: the bits in which the creation times of the
 * listed tasks differ, the radix passes run (over a prefix of those bits only when they are many: the top 32 for
 * 25 M tasks), the groups of equal prefixes that stood out of order afterwards and were put right in place, and
 * whether such groups were too long or too many, so that ordinary passes over every digit finished the job.
 * Any of the pointers may be null.  No reference counterpart. */
int32_t tt_last_order_info(tt_table* t, int32_t* key_bits, int32_t* passes, int64_t* groups_fixed, int32_t* fell_back);

/* The same for a batch of creators in ONE pass over the table (the hash column is read once for
 * all of them): out[i] receives the result of query i, nq <= 1024.  No reference counterpart as
 * a single call — it is GetTasksByCreator (ITasksManager.cs:7) for nq concurrent requests. */
int32_t tt_get_by_creators(tt_table* t, const char* const* utf8, const int32_t* lens, int32_t nq, tt_result** out);

/* Device timings of the last tt_get_by_creator(s) call on this table. */
typedef struct tt_op_stats {
  float scan_ms;      /* K4 hash-column scan + exact confirm                                   */
  float order_ms;     /* K5 ordering + TaskId gather of the matches                             */
  int64_t candidates; /* rows whose 64-bit hash matched a query                                 */
  int64_t matches;    /* rows whose string matched (all queries together)                       */
  int32_t launches;
  int32_t queries;
} tt_op_stats;
int32_t tt_last_op_stats(tt_table* t, tt_op_stats* out);

/* ---- the single-row operations of ITasksManager ------------------------------------------
 * Not data-parallel and kept simple (a lookup kernel + a one-thread field write); they exist so
 * the table can stand behind the whole interface.  TaskName / TaskAssignedTo stay with the host
 * wrapper, and so does generating the Guid and reading the clock.  *found = 1 if a task with
 * that id exists (the reference's `true` / non-null result), 0 otherwise. */

/* ITasksManager.GetTaskById (ITasksManager.cs:8; FakeTasksManager.cs:62-66,
 * TasksStoreManager.cs:47-52): a result with one row, or an empty one. */
int32_t tt_get_by_id(tt_table* t, const uint8_t* id16, tt_result** out);
/* ITasksManager.CreateNewTask (ITasksManager.cs:9; FakeTasksManager.cs:32-47,
 * TasksStoreManager.cs:22-38): appends a pending task at the end of the list; *row = its row. */
int32_t tt_insert_task(tt_table* t, const uint8_t* id16, uint64_t created_raw, uint64_t due_raw, const char* created_by_utf8,
                       int32_t len, int64_t* row);
/* The state store's upsert, SaveStateAsync(key = TaskId.ToString(), whole record): what
 * TasksStoreManager.MarkOverdueTasks does for every listed task (TasksStoreManager.cs:146-147) — a task
 * with this id has every kernel-visible field overwritten (*existed = 1), otherwise the task is appended
 * (*existed = 0); *row = its row.  The host mirrors use it for the ids tt_mark_overdue reports missing. */
int32_t tt_upsert_task(tt_table* t, const uint8_t* id16, uint64_t created_raw, uint64_t due_raw, const char* created_by_utf8,
                       int32_t len, int32_t is_completed, int32_t is_overdue, int64_t* row, int32_t* existed);
/* ITasksManager.UpdateTask (ITasksManager.cs:10; FakeTasksManager.cs:87-100,
 * TasksStoreManager.cs:84-102): the due date is the only updated field a kernel reads. */
int32_t tt_update_task(tt_table* t, const uint8_t* id16, uint64_t due_raw, int32_t* found);
/* ITasksManager.MarkTaskCompleted (ITasksManager.cs:11; FakeTasksManager.cs:74-85,
 * TasksStoreManager.cs:71-82). */
int32_t tt_mark_completed(tt_table* t, const uint8_t* id16, int32_t* found);
/* ITasksManager.DeleteTask (ITasksManager.cs:12; FakeTasksManager.cs:49-60,
 * TasksStoreManager.cs:40-45).  The row becomes a tombstone; row numbers of other tasks and
 * their relative order do not change. */
int32_t tt_delete_task(tt_table* t, const uint8_t* id16, int32_t* found);
/* Tasks in the table, not counting deleted ones (tt_row_count counts rows, tombstones included). */
int32_t tt_live_count(tt_table* t, int64_t* n);

/* ---- results --------------------------------------------------------------------------- */

int32_t tt_result_count(const tt_result* r, int64_t* n);
/* shard-local row numbers, uint32 per hit */
int32_t tt_result_rows(tt_result* r, uint32_t* out);
/* TaskIds, 16 bytes per hit */
int32_t tt_result_ids(tt_result* r, uint8_t* out16);
/* any other column for the hits (TT_FIELD_*) */
int32_t tt_result_field(tt_result* r, int32_t field, void* out);
int32_t tt_result_free(tt_result* r);

/* ---- state access used by the parity tests and by the benchmark's per-step reset --------- */

/* Bit-planes, (n+7)/8 bytes each, bit (r & 7) of byte (r >> 3) = row r.  Either may be NULL.
 * The buffers may be host (pageable or pinned) or device memory (copied with cudaMemcpyDefault). */
int32_t tt_snapshot_flags(tt_table* t, uint8_t* completed_bits, uint8_t* overdue_bits);
int32_t tt_restore_flags(tt_table* t, const uint8_t* completed_bits, const uint8_t* overdue_bits);
/* Population counts of the two planes (computed on the device). */
int32_t tt_count_flags(tt_table* t, int64_t* n_completed, int64_t* n_overdue);
/* Order-free digest of a result computed on the device: XOR of the TaskIds (hi = bytes 0..7,
 * lo = bytes 8..15, big-endian) and the sum of the row numbers.  Lets the 1 B-row parity test
 * compare with the streaming oracle without moving 25 M ids to the host. */
int32_t tt_result_digest(tt_result* r, uint64_t* xor_id_hi, uint64_t* xor_id_lo, uint64_t* sum_rows);

/* ---- device-resident form of the sweep (no host copy of the hits) ------------------------
 * Same kernels as tt_sweep, writing into caller-owned DEVICE buffers on a caller stream
 * (cudaStream_t passed as void*; NULL = the table's own stream).  d_rows: uint32[cap],
 * d_ids: 16*cap bytes (may be NULL to skip the TaskId gather); d_rows == NULL counts only.
 * h_stats is host memory, filled when the call returns (the call synchronises the stream).
 * The table keeps an internal scratch list of `cap` row numbers; it is (re)allocated only when
 * a call asks for a larger cap than any call before it.
 * If the hits do not fit in cap the call returns TT_ERR_CAPACITY with n_listed = the number
 * of hits there are and n_stored = how many were written (a row-ordered subset).  A marking
 * sweep has marked exactly the stored ones: call again (they no longer match) to get the
 * rest.  A listing sweep changed nothing: call again with cap >= n_listed. */
typedef struct tt_sweep_stats {
  int64_t n_listed;   /* hits in [first_row, n): what GetYesterdaysDueTasks returns            */
  int64_t n_stored;   /* hits actually written to d_rows / d_ids                               */
  int64_t n_marked;   /* rows whose IsOverDue was set                                          */
  int64_t resume_row; /* row count when complete, else first_row                               */
  float kernel_ms;    /* CUDA-event time of this call's kernels on their stream (sweep + gather) */
  int32_t launches;   /* kernels launched by this call                                          */
  float sweep_ms;     /* K1: predicate + compaction + flag write                                */
  float gather_ms;    /* K2 + K3: row order and TaskId gather for the listed rows               */
} tt_sweep_stats;

#define TT_SWEEP_MARK 1     /* apply the cron predicate and set IsOverDue (the fused tick)      */
#define TT_SWEEP_LIST_ONLY 0 /* GetYesterdaysDueTasks only                                        */
#define TT_SWEEP_DUE_ON_OR_BEFORE 2 /* OR it in: list (and mark) every task whose due DATE is yesterday or
                                       earlier, whatever its time of day or Kind — the backlog the exact
                                       "due yesterday at 00:00:00" match leaves behind when a cron run was
                                       missed or a due date carries a time (the edge cases
                                       docs/aca/07-aca-cron-bindings/index.md:106 mentions).  An extension:
                                       the reference has no such query.  Flag tests of
                                       STORE mode still apply; with TT_SWEEP_MARK, run_at must lie after
                                       yesterday. */

#define TT_SWEEP_NO_TIMING 4 /* OR it in: record no CUDA events between the kernels; the three timing fields of
                                the stats stay 0 (the events cost a few microseconds of a tick)                 */

#define TT_SWEEP_ASYNC 8     /* OR it in (tt_sweep_device, tt_sweep_push): queue the tick's kernels and return at
                                once; h_stats is filled by tt_sweep_wait, which also reports TT_ERR_CAPACITY.  Until
                                then no other call may be made on the table (another sweep is refused with
                                TT_ERR_STATE).  The kernels report their outcome into mapped pinned memory, so the
                                tick itself contains no copy and no host round trip.                            */

int32_t tt_sweep_device(tt_table* t, int32_t mode, int32_t flags, int64_t today_local_ticks,
                        int64_t run_at_utc_ticks, int64_t first_row, uint32_t* d_rows, void* d_ids, int64_t cap,
                        tt_sweep_stats* h_stats, void* stream);
/* Completes a tick queued with TT_SWEEP_ASYNC: waits for its stream, fills h_stats. */
int32_t tt_sweep_wait(tt_table* t, tt_sweep_stats* h_stats);

/* ---- the tick for a caller whose task state and results live in HOST memory -------------------
 * The whole of Svc/Controllers/ScheduledTasksManagerController.cs:19-46 as tt_sweep does it, but with
 * the two flag planes arriving from host memory and the listed rows + TaskIds delivered into host
 * memory, the three stages pipelined over `chunks` row ranges of the table (0 = 8): while chunk c is
 * swept on the table's stream, chunk c+1's planes are uploaded on a second stream and chunk c-1's
 * hits are downloaded on a third, so the PCIe link works in both directions under the kernels.
 * h_completed_bits / h_overdue_bits: the layout of tt_snapshot_flags, either may be NULL (that plane
 * stays as it is on the device); h_rows: uint32[cap]; h_ids16: 16*cap bytes or NULL.  Use pinned
 * memory: pageable buffers make the copies synchronous and the stages stop overlapping.
 * The rows arrive in ascending row order (tt_sweep's order).  stats->n_listed / n_stored / n_marked as
 * in tt_sweep_device; the timing fields are sums over the chunks.  TT_ERR_CAPACITY if the hits do not
 * fit in cap (the first cap are delivered; a marking tick has marked all it listed — the device's overdue
 * plane is then ahead of the caller's: read it back with tt_snapshot_flags before the next tick uploads
 * the host copy again).  On every return, error returns included, no transfer to or from the caller's
 * buffers is still running. */
int32_t tt_tick_host(tt_table* t, int32_t mode, int32_t flags, int64_t today_local_ticks, int64_t run_at_utc_ticks,
                     const uint8_t* h_completed_bits, const uint8_t* h_overdue_bits, uint32_t* h_rows, uint8_t* h_ids16,
                     int64_t cap, int32_t chunks, tt_sweep_stats* stats);

/* Device addresses of the table's columns, for callers that exchange shard results over
 * NCCL or reset state between benchmark steps.  Any out pointer may be NULL. */
typedef struct tt_device_view {
  int64_t n;           /* rows                      */
  int64_t n_padded;    /* rows allocated (tile multiple) */
  void* due_code;      /* uint32[n_padded]          */
  void* completed;     /* uint32[n_padded/32]       */
  void* overdue;       /* uint32[n_padded/32]       */
  void* creator_hash;  /* uint64[n_padded]          */
  void* task_id;       /* 16 B * n_padded           */
  void* created_raw;   /* uint64[n_padded]          */
  int32_t device;
} tt_device_view;
int32_t tt_get_device_view(tt_table* t, tt_device_view* out);

/* ---- compact hit lists (what shards exchange) ---------------------------------
 * A row-ordered hit list is equivalent to "hits per 8192-row tile" + a 13-bit offset per hit:
 * 2 bytes per hit instead of 4.  No reference counterpart (the reference ships whole TaskModel
 * objects as JSON between its processes, Svc/Controllers/ScheduledTasksManagerController.cs:28,44).
 * tt_pack_hits   : after a complete tt_sweep_device call on this table, writes d_off16[i] = offset of
 *                  d_rows[i] in its tile (i < n) and copies the per-tile hit counts of that sweep
 *                  (uint32 per tile, *n_tiles of them = n_padded / 8192) into d_tile_count.
 * tt_expand_hits : the inverse, on any device: rebuilds the shard-local row numbers.  d_scratch needs
 *                  n_tiles + n_tiles/1024 + 1 uint32 of device memory. */
int32_t tt_pack_hits(tt_table* t, const uint32_t* d_rows, int64_t n, uint16_t* d_off16, uint32_t* d_tile_count,
                     int64_t* n_tiles, void* stream);
int32_t tt_expand_hits(int32_t device, const uint32_t* d_tile_count, int64_t n_tiles, const uint16_t* d_off16,
                       uint32_t* d_scratch, uint32_t* d_rows_out, void* stream);

/* ---- the sweep fused with the cross-shard exchange -----------------------------
 * tt_sweep_push is tt_sweep_device over the whole table whose last kernel, while it puts the hits
 * in row order and gathers their TaskIds, also stores the COMPACT form of the result (the layout of
 * tt_pack_hits: off16[i] for the n_stored hits, one count per 8192-row tile, and the hit count)
 * into `target`.  The target buffers may belong to another GPU: the root shard allocates them with
 * tt_peer_alloc, every other process maps them with tt_peer_open (CUDA IPC; stores travel over
 * NVLink) and passes the mapped addresses of ITS slot.  The exchange then needs no pass of its own
 * and no collective: once every shard's call has returned and the processes have met at a barrier,
 * the root reads all slots (tt_expand_hits rebuilds row numbers).  d_rows / d_ids (this shard's own
 * copy of the result) may both be NULL.  Replaces, for one box, the JSON hop of the hit list between
 * the reference's processes (Svc/Controllers/ScheduledTasksManagerController.cs:28,44).
 * target->off16 must be 16-byte aligned and hold `cap` entries. */
typedef struct tt_push_target {
  void* off16;      /* uint16[cap]                          */
  void* tile_count; /* uint32[n_padded / 8192]               */
  void* count;      /* uint64[1]: hits stored (= n_stored)   */
} tt_push_target;
int32_t tt_sweep_push(tt_table* t, int32_t mode, int32_t flags, int64_t today_local_ticks, int64_t run_at_utc_ticks,
                      uint32_t* d_rows, void* d_ids, int64_t cap, const tt_push_target* target,
                      tt_sweep_stats* h_stats, void* stream);

/* Device memory that other processes on the box can map.  tt_peer_alloc: cudaMalloc (zeroed) on
 * `device` + an exportable handle (64 opaque bytes, send them to the other processes any way you
 * like); tt_peer_open: maps it for kernels running on `device` of the calling process (peer access
 * is enabled as needed); close what you opened, free what you allocated.  No reference counterpart. */
typedef struct tt_peer_handle {
  uint8_t bytes[64];
} tt_peer_handle;
int32_t tt_peer_alloc(int32_t device, uint64_t bytes, void** d_ptr, tt_peer_handle* handle);
int32_t tt_peer_open(int32_t device, const tt_peer_handle* handle, void** d_ptr);
int32_t tt_peer_close(int32_t device, void* d_ptr);
int32_t tt_peer_free(int32_t device, void* d_ptr);

/* ---- payload strings and the JSON form of a result ---------------------------
 * TaskName and TaskAssignedTo (Api/Models/TaskModel.cs:6,10) are never read by the sweep or the
 * filter; they are kept as dictionary-encoded string columns only so that a result can leave the
 * GPU as the text the reference's controllers send.  Without a payload both print as "" (the
 * reference's default, `= string.Empty`); rows added later by tt_insert_task print "" until
 * tt_set_task_strings gives them their strings. */
typedef struct tt_strings_view {
  const uint32_t* ref;  /* one entry per row: index into the dictionary, 0xFFFFFFFF = empty string, 0xFFFFFFFE = null */
  uint32_t n_strings;   /* dictionary entries                                                      */
  const uint32_t* off;  /* n_strings + 1 byte offsets into pool                                    */
  const uint8_t* pool;  /* UTF-8 (must be well formed; checked when the first JSON is produced)    */
} tt_strings_view;
/* Either view may be NULL (that column is left as it is).  Call after tt_load_soa / tt_generate. */
int32_t tt_load_payload(tt_table* t, const tt_strings_view* task_name, const tt_strings_view* assigned_to);
/* The synthetic payload of include/tt_synth.h for the rows tt_generate made (same cfg, same first_row):
 * "Task number: {i}" names (Api/Services/FakeTasksManager.cs:16; one in eight with characters JSON must
 * escape) and "assignee{0..49}@mail.com" (FakeTasksManager.cs:21). */
int32_t tt_generate_payload(tt_table* t, const tt_synth_cfg* cfg, uint64_t first_row);

/* TaskName / TaskAssignedTo of ONE task, for the single-row methods that set them: CreateNewTask
 * (ITasksManager.cs:9; after tt_insert_task) and UpdateTask (ITasksManager.cs:10; FakeTasksManager.cs:92-94,
 * TasksStoreManager.cs:91-93).  A length of -1 leaves that member as it is, -2 sets it to null (a C# string may
 * hold null; it is then written as the JSON literal null).  New strings are appended to the
 * column's dictionary (the column is created if the table had no payload).  *found as in tt_update_task. */
int32_t tt_set_task_strings(tt_table* t, const uint8_t* id16, const char* task_name, int32_t name_len,
                            const char* assigned_to, int32_t assigned_len, int32_t* found);

/* tt_result_json writes the rows of a result, in result order, as the JSON array ASP.NET Core
 * produces for the List<TaskModel> the controllers return (Api/Controllers/OverdueTasksController.cs:22-26,
 * Api/Controllers/TasksController.cs:21-25): camelCase names in declaration order, no whitespace,
 * Guid "D" form, DateTime as System.Text.Json prints it (docs/aca/05-aca-dapr-pubsubapi/index.md:60-69),
 * isCompleted / isOverDue as they are in the table NOW.  `flavor` selects the string encoder:
 * TT_JSON_RESPONSE = the MVC output formatter (UnsafeRelaxedJsonEscaping), TT_JSON_DOCUMENT = the
 * bodies Dapr's client writes (SaveStateAsync / PublishEventAsync, Api/Services/TasksStoreManager.cs:35,155;
 * JavaScriptEncoder.Default).  local_offset_min is the UTC offset printed for Kind = Local values.
 * The text stays on the device until tt_json_copy (dst: host or device memory, tt_json_size bytes,
 * no terminator). */
#define TT_JSON_RESPONSE 0
#define TT_JSON_DOCUMENT 1
typedef struct tt_json tt_json;
int32_t tt_result_json(tt_result* r, int32_t flavor, int32_t local_offset_min, tt_json** out);
int32_t tt_json_size(const tt_json* j, int64_t* bytes);
int32_t tt_json_copy(tt_json* j, void* dst);
int32_t tt_json_device(tt_json* j, const void** d_text);
int32_t tt_json_free(tt_json* j);
typedef struct tt_json_stats {
  float size_ms;   /* K6a: record sizes + offsets            */
  float emit_ms;   /* K6b: formatting + write                */
  int64_t bytes;   /* length of the text                     */
  int64_t records;
  int32_t launches;
} tt_json_stats;
int32_t tt_last_json_stats(tt_table* t, tt_json_stats* out);

/* ---- bulk load from the state store's documents -----------------------------
 * The reference stores every task as one JSON document under the key TaskId.ToString()
 * (Api/Services/TasksStoreManager.cs:35) and reads them back through System.Text.Json
 * (GetStateAsync<TaskModel> :50, QueryStateAsync<TaskModel> :61,130).  tt_load_documents parses n
 * such documents on the device — one thread per document, members in any order, whitespace,
 * unknown members, every string escape, the date forms System.Text.Json accepts, names matched
 * ignoring ASCII case — and fills an EMPTY table with them in document order: the fixed columns,
 * and TaskName / TaskCreatedBy / TaskAssignedTo dictionary-encoded on the device (hash sort +
 * exact confirm).  Document i is docs[doc_off[i] .. doc_off[i+1]); it may carry one leading '['
 * or ',' and one trailing ',' or ']' (a JSON array cut at its element boundaries, which is what
 * tt_json_offsets returns for a tt_result_json text).  docs and doc_off may each be host or device
 * memory (offsets that run backwards or outside [doc_off[0], doc_off[n]] make that document bad).
 * local_offset_min: the machine's UTC offset, used for dates that carry an
 * offset (System.Text.Json turns them into local time, Kind = Local).
 * If any document is malformed the call fails with TT_ERR_ARG, stats->first_bad / first_bad_reason
 * say which and why, and the table stays empty.  A string member that is null in its document stays
 * null (a missing one is "", TaskModel.cs:6,7,10): it prints as null again, and a task whose
 * TaskCreatedBy is null matches no creator (the store's EQ filter; TT_FIELD_CREATOR_REF gives
 * 0xFFFFFFFE for it). */
typedef struct tt_ingest_stats {
  int64_t n_docs;
  int64_t text_bytes;
  int64_t first_bad;         /* -1 when every document parsed */
  int32_t first_bad_reason;  /* 1 shape, 2 Guid, 3 date, 4 string, 5 value type */
  uint32_t n_names, n_creators, n_assignees; /* dictionary sizes */
  float parse_ms;            /* K7a */
  float encode_ms;           /* column fill + the three dictionary encodings */
  int32_t launches;
} tt_ingest_stats;
int32_t tt_load_documents(tt_table* t, const void* docs, const uint64_t* doc_off, int64_t n, int32_t local_offset_min,
                          tt_ingest_stats* stats);
/* Byte offsets of the records of a tt_result_json text: off[i] = where record i's piece starts (its
 * leading '[' or ','), off[n] = the length of the text; n + 1 values copied to dst (host or device). */
int32_t tt_json_offsets(tt_json* j, uint64_t* dst);

/* ---- snapshot ---------------------------------------------------------------------------------
 * Saves / restores the whole table (all columns, flags, creator dictionary, tombstones) as one
 * file in the device layout — the counterpart of the reference keeping its tasks in a state store
 * across restarts (Api/Services/TasksStoreManager.cs:35 SaveStateAsync).  tt_load_snapshot needs
 * an empty table on the same kind of build (same tile size).  The payload columns (TaskName /
 * TaskAssignedTo), when the table has them, are part of the file. */
int32_t tt_save_snapshot(tt_table* t, const char* path);
int32_t tt_load_snapshot(tt_table* t, const char* path);

/* ---- the table sharded by TaskId range over several GPUs ------------------------
 * The reference registers ONE manager (Api/Program.cs:13) and its methods return ONE ordered list
 * (Api/Services/ITasksManager.cs:7,14); these entry points give the same contract over G shards.
 *
 * Layer 1, tt_exchange_*: one exchange object per shard.  The root shard's exchange owns a receive
 * buffer in its GPU's HBM with one slot per shard; a shard in another process maps it through the
 * 64-byte handle (CUDA IPC), a shard in the same process attaches to the root's object.
 * tt_exchange_tick is the shard's cron tick (Svc/Controllers/ScheduledTasksManagerController.cs:19-46)
 * whose placement kernel stores the result straight into the shard's slot — per tile a hit count and
 * a list offset, per hit a 13-bit in-tile offset and, if the exchange was made for them, the TaskId
 * (TT_EXCHANGE_IDS) and TaskCreatedOn (TT_EXCHANGE_CREATED) — over NVLink when the slot is on another
 * GPU; its last kernel publishes the tick's epoch in the slot.  No collective is involved: on the root,
 * tt_exchange_collect waits for the epoch words ON THE DEVICE (bounded by TT_SHARD_TIMEOUT_MS,
 * default 10 s: a shard that never ticks gives TT_ERR_STATE, not a hung GPU), lays the slots end to
 * end and rebuilds the row numbers.  Every shard must tick once per collect; expected_epoch is the
 * number of ticks made so far (tt_exchange_epoch).  A shard does not overwrite its slot before the
 * root has released the previous epoch (tt_exchange_release, or TT_COLLECT_RELEASE).
 *
 * order = TT_ORDER_ROWS: the shards' row-ordered lists end to end — shards are TaskId ranges, so this
 * is the order of FakeTasksManager.GetYesterdaysDueTasks (Api/Services/FakeTasksManager.cs:107-112)
 * over a table kept in TaskId order.  The result is a view of the exchange's buffers until it is first
 * read; read or free it before the next tick.  order = TT_ORDER_CREATED_ASC: OrderBy(TaskCreatedOn)
 * over the hits of all shards (Api/Services/TasksStoreManager.cs:136), stable, ties in TaskId-range
 * order: one radix sort of the pushed ticks on the root.  first_rows[r] = global row of shard r's
 * first row (what the result's row numbers are relative to). */
typedef struct tt_exchange tt_exchange;
typedef struct tt_gresult tt_gresult;
#define TT_EXCHANGE_IDS 1
#define TT_EXCHANGE_CREATED 2
#define TT_ORDER_ROWS 0
#define TT_ORDER_CREATED_ASC 1
#define TT_COLLECT_RELEASE 0x100 /* OR into order (TT_ORDER_ROWS only): release the epoch inside the collect; the
                                    result's TaskIds must then be read before any shard's next tick reaches K3g  */
int32_t tt_exchange_create_root(int32_t device, int32_t world, int64_t cap_per_shard, int64_t max_tiles_per_shard,
                                int32_t flags, tt_exchange** out, tt_peer_handle* handle /* may be NULL */);
int32_t tt_exchange_attach(int32_t device, int32_t rank, int32_t world, int64_t cap_per_shard, int64_t max_tiles_per_shard,
                           int32_t flags, const tt_peer_handle* handle, tt_exchange** out);
int32_t tt_exchange_attach_local(tt_exchange* root, int32_t device, int32_t rank, tt_exchange** out);
/* Destroy the shards' exchanges before the root's, and read or free every result of tt_exchange_collect first (an
 * unread TT_ORDER_ROWS result still points into the root's buffers). */
int32_t tt_exchange_destroy(tt_exchange* ex);
int32_t tt_exchange_slot_bytes(const tt_exchange* ex, uint64_t* bytes);
/* flags: TT_SWEEP_* as for tt_sweep_device (TT_SWEEP_ASYNC: finish with tt_sweep_wait on the table).
 * d_rows / d_ids: the shard's own copy of its list (uint32[cap], 16 B x cap; device memory), either
 * may be NULL.  TT_ERR_CAPACITY if the shard's hits do not fit in cap_per_shard (the slot then holds a
 * subset). */
int32_t tt_exchange_tick(tt_exchange* ex, tt_table* t, int32_t mode, int32_t flags, int64_t today_local_ticks,
                         int64_t run_at_utc_ticks, uint32_t* d_rows, void* d_ids, tt_sweep_stats* h_stats, void* stream);
int32_t tt_exchange_epoch(const tt_exchange* ex, uint64_t* epoch);
int32_t tt_exchange_collect(tt_exchange* root, int32_t order, const uint64_t* first_rows, uint64_t expected_epoch,
                            tt_gresult** out);
int32_t tt_exchange_release(tt_exchange* root, uint64_t epoch);
/* Device time of the last collect on this root (wait for the slowest shard + rebuilding the rows). */
int32_t tt_exchange_last_collect_ms(const tt_exchange* root, float* ms);

int32_t tt_gresult_count(const tt_gresult* r, int64_t* n);
int32_t tt_gresult_shard_counts(const tt_gresult* r, int64_t* counts, int32_t n_shards);
int32_t tt_gresult_rows(tt_gresult* r, uint64_t* out);   /* global row numbers, result order            */
int32_t tt_gresult_ids(tt_gresult* r, uint8_t* out16);   /* TaskIds, 16 bytes each, result order        */
int32_t tt_gresult_digest(tt_gresult* r, uint64_t* xor_id_hi, uint64_t* xor_id_lo, uint64_t* sum_rows);
int32_t tt_gresult_free(tt_gresult* r);

/* Layer 2, tt_shards_*: the whole ITasksManager (Api/Services/ITasksManager.cs:5-15) over G shards
 * driven from ONE process — what a C# host registers as its singleton.  devices[g] holds shard g (a
 * device may appear more than once).  Shard g owns the g-th TaskId range: equal-width ranges of the
 * 128-bit id space for tt_shards_load_soa and for tables filled by inserts; for tt_shards_generate
 * the g-th equal share of the synthetic rows (whose ids grow with the row).  Global row numbers count
 * the shards' rows end to end.  The list calls size the exchange with a count-only pass (K1 alone)
 * and grow it on demand.  Lists are ordered as the reference orders them: GetYesterdaysDueTasks by
 * TaskCreatedOn in STORE mode (TasksStoreManager.cs:136) and in TaskId-range order in FAKE mode;
 * GetTasksByCreator by TaskCreatedOn descending (FakeTasksManager.cs:70, TasksStoreManager.cs:66);
 * ties keep TaskId-range order.  The single-row methods go to the shard that owns the id's range. */
typedef struct tt_shards tt_shards;
int32_t tt_shards_create(const int32_t* devices, int32_t n_shards, tt_shards** out);
int32_t tt_shards_destroy(tt_shards* s);
int32_t tt_shards_count(const tt_shards* s, int32_t* n_shards);
/* Shard `shard`'s table (still owned by s; for state access such as tt_snapshot_flags) and the global
 * row number of its first row.  Either out pointer may be NULL. */
int32_t tt_shards_table(tt_shards* s, int32_t shard, tt_table** out, uint64_t* first_row);
int32_t tt_shards_generate(tt_shards* s, const tt_synth_cfg* cfg, uint64_t n_total);
int32_t tt_shards_generate_payload(tt_shards* s, const tt_synth_cfg* cfg, uint64_t n_total);
int32_t tt_shards_load_soa(tt_shards* s, const tt_soa_view* v);
int32_t tt_shards_row_count(tt_shards* s, int64_t* n);
int32_t tt_shards_get_yesterdays_due(tt_shards* s, int32_t mode, int64_t today_local_ticks, tt_gresult** out);
int32_t tt_shards_get_due_on_or_before(tt_shards* s, int32_t mode, int64_t today_local_ticks, tt_gresult** out);
/* The cron tick over all shards (tt_sweep; flags = 0 or TT_SWEEP_DUE_ON_OR_BEFORE): *out lists what
 * GetYesterdaysDueTasks returned, in TaskId-range order; FAKE mode with survivors returns
 * TT_ERR_NOT_IMPLEMENTED with the list produced and nothing changed. */
int32_t tt_shards_sweep(tt_shards* s, int32_t mode, int32_t flags, int64_t today_local_ticks, int64_t run_at_utc_ticks,
                        tt_gresult** out, int64_t* n_marked);
int32_t tt_shards_mark_overdue(tt_shards* s, int32_t mode, const uint8_t* ids16, int64_t n, int64_t* n_marked,
                               int64_t* n_missing);
int32_t tt_shards_get_by_creator(tt_shards* s, const char* utf8, int32_t len, tt_gresult** out);
int32_t tt_shards_get_by_id(tt_shards* s, const uint8_t* id16, tt_gresult** out);
int32_t tt_shards_insert_task(tt_shards* s, const uint8_t* id16, uint64_t created_raw, uint64_t due_raw,
                              const char* created_by_utf8, int32_t len, int32_t* shard, int64_t* local_row);
int32_t tt_shards_update_task(tt_shards* s, const uint8_t* id16, uint64_t due_raw, int32_t* found);
int32_t tt_shards_mark_completed(tt_shards* s, const uint8_t* id16, int32_t* found);
int32_t tt_shards_delete_task(tt_shards* s, const uint8_t* id16, int32_t* found);

/* ---- sweep without any allocation ----------------------------------------------------------
 * This entry point has no counterpart in the reference.  It exists for hosts that cannot let a
 * library allocate device memory while they are serving: a process that handed the whole GPU to
 * its own memory pool at start-up (cudaMalloc would simply fail), or a request path that must not
 * run into the device-wide synchronisation and the unpredictable latency of cudaMalloc/cudaFree.
 * tt_sweep_device keeps its working memory in the table and grows it on demand; this variant
 * runs the same kernels with all of that working memory (control block, per-tile words and the
 * intermediate hit list) in one fixed array that is part of the library image — about 4 MB that
 * every GPU which loads the library holds until the process exits.  It calls no cudaMalloc,
 * cudaFree, event/stream creation or any other allocating routine.
 *
 * The price is a size limit: cols->n_padded <= TT_STATIC_MAX_ROWS and cap <= TT_STATIC_MAX_HITS,
 * anything larger is refused with TT_ERR_ARG before a kernel is launched.
 *
 * cols   : the table's columns as device pointers (from tt_get_device_view, or the caller's own
 *          arrays in the same layout: padded to a multiple of 8192 rows, padding rows with
 *          due_code 0xFFFFFFFF and zero flag bits).  cols->device must own the memory.
 * d_rows : uint32[cap] device memory, d_ids: 16*cap bytes or NULL; d_rows == NULL counts only.
 * Results, error codes, the flags (TT_SWEEP_MARK, TT_SWEEP_DUE_ON_OR_BEFORE) and the meaning of
 * TT_ERR_CAPACITY are those of tt_sweep_device; the timing fields of h_stats stay 0 (timing would
 * need events).  The call does not lock the table the columns came from: fetch the view again
 * after anything that can move the columns (tt_insert_task growing the table, a reload).
 *
 * The first call on each device looks up the device address of the fixed array; make that first
 * call before any stream capture is started.  All calls share the array and it keeps whatever
 * the previous call left in it; each call re-initialises the words that must start at zero, and
 * calls from several threads are serialised inside the library.  Two calls that nevertheless
 * overlap on the GPU (for example from two processes' worth of streams in one context created
 * behind the library's back) would race on it. */
#define TT_STATIC_MAX_ROWS 16777216 /* 2048 tiles of 8192 rows */
#define TT_STATIC_MAX_HITS 1048576
int32_t tt_sweep_static(const tt_device_view* cols, int32_t mode, int32_t flags, int64_t today_local_ticks,
                        int64_t run_at_utc_ticks, uint32_t* d_rows, void* d_ids, int64_t cap,
                        tt_sweep_stats* h_stats, void* stream);

#ifdef __cplusplus
}
#endif

#endif /* TASKTABLE_H */
