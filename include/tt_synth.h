/* tt_synth.h — deterministic synthetic TaskModel rows, shared by host and device.
 *
 * A row is a pure function of (cfg, global_row): the CPU oracle, the host loader and every
 * GPU shard materialise identical rows without shipping the table over PCIe.  This header
 * defines INPUTS only (what a synthetic task looks like); it contains none of the sweep /
 * filter logic, so the oracle including it does not make the oracle depend on the product.
 *
 * Field domains follow the reference's producers:
 *   TaskId          Guid.NewGuid() = RFC 4122 v4   (Api/Services/FakeTasksManager.cs:36)
 *   TaskCreatedOn   DateTime.UtcNow, Kind = Utc    (Api/Services/FakeTasksManager.cs:39)
 *   TaskDueDate     <input type="date"> midnight, Kind = Unspecified
 *                   (Ui/Pages/Tasks/Create.cshtml:22) or a queue message with a time of
 *                   day (Svc/Controllers/ExternalTasksProcessorController.cs:29-33)
 *   TaskCreatedBy   an e-mail-like string          (Api/Services/FakeTasksManager.cs:18)
 *   TaskName        "Task number: {i}"             (Api/Services/FakeTasksManager.cs:16); one name
 *                   in eight carries characters JSON has to escape
 *   TaskAssignedTo  "assignee{rnd.Next(50)}@mail.com" (Api/Services/FakeTasksManager.cs:21)
 *
 * DateTime values use .NET's internal 64-bit "dateData": ticks (100 ns since
 * 0001-01-01T00:00:00) in the low 62 bits, DateTimeKind in the top 2 bits.
 */
#ifndef TT_SYNTH_H
#define TT_SYNTH_H

#include <stdint.h>

#if defined(__CUDACC__)
#define TT_HD __host__ __device__ __forceinline__
#else
#define TT_HD static inline
#endif

#define TT_TICKS_PER_DAY 864000000000LL
#define TT_TICKS_MASK 0x3FFFFFFFFFFFFFFFULL
#define TT_KIND_UNSPECIFIED 0u
#define TT_KIND_UTC 1u
#define TT_KIND_LOCAL 2u

/* 2026-11-06T00:00:00 (day 739925 since 0001-01-01) and the cron's 00:05 run time. */
#define TT_DEFAULT_TODAY_TICKS 639295200000000000LL
#define TT_DEFAULT_RUN_AT_TICKS (TT_DEFAULT_TODAY_TICKS + 5LL * 60 * 10000000LL)

typedef struct tt_synth_cfg {
  uint64_t seed;
  uint64_t n_total;            /* rows of the whole table, all shards together          */
  int64_t today_local_ticks;   /* local midnight "DateTime.Today" the table is built around */
  uint32_t ppm_due_yesterday;  /* due == yesterday 00:00:00, Kind Unspecified (exact hit)   */
  uint32_t ppm_adversarial;    /* yesterday's date but time-of-day != 0 or Kind != Unspecified */
  uint32_t ppm_completed;      /* IsCompleted, independent                                  */
  uint32_t ppm_overdue_past;   /* IsOverDue already set, among rows due before yesterday    */
  uint32_t ppm_created_dup;    /* TaskCreatedOn copied from the previous row (sort ties)    */
  uint32_t ppm_timed_due;      /* among other rows: due date carries a time of day, Kind Utc */
  uint32_t n_creators;         /* distinct TaskCreatedBy values                             */
  uint32_t n_names;            /* distinct TaskName values (0 = 1 000 000); row r has name r % n_names */
} tt_synth_cfg;

typedef struct tt_synth_row {
  uint64_t id_hi, id_lo; /* TaskId in canonical big-endian ("D" string) order, hi = bytes 0..7 */
  uint64_t created_raw;  /* dateData of TaskCreatedOn */
  uint64_t due_raw;      /* dateData of TaskDueDate   */
  uint32_t creator;      /* index k of the creator string tt_synth_creator(k) */
  uint8_t completed, overdue;
} tt_synth_row;

/* BASELINE.json config 5 mix: 5 % due yesterday, 50 % completed, 1 % adversarial. */
TT_HD tt_synth_cfg tt_synth_default(uint64_t seed, uint64_t n_total) {
  tt_synth_cfg c;
  c.seed = seed;
  c.n_total = n_total;
  c.today_local_ticks = TT_DEFAULT_TODAY_TICKS;
  c.ppm_due_yesterday = 50000;
  c.ppm_adversarial = 10000;
  c.ppm_completed = 500000;
  c.ppm_overdue_past = 100000;
  c.ppm_created_dup = 1000;
  c.ppm_timed_due = 100000;
  c.n_creators = 1000000;
  c.n_names = 0;
  return c;
}

TT_HD uint64_t tt_mix64(uint64_t z) { /* splitmix64 finaliser */
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

/* k-th independent 64-bit draw for a row */
TT_HD uint64_t tt_draw(uint64_t seed, uint64_t row, uint32_t k) {
  return tt_mix64(tt_mix64(seed ^ (row * 0xD1342543DE82EF95ULL)) + (uint64_t)k * 0xA0761D6478BD642FULL);
}

/* multiply-shift range reduction of the top 32 bits: uniform in [0, n) */
TT_HD uint32_t tt_range(uint64_t h, uint32_t n) { return (uint32_t)(((h >> 32) * (uint64_t)n) >> 32); }

TT_HD uint64_t tt_synth_created_base(const tt_synth_cfg* c, uint64_t row) {
  /* uniform in the two years before `today`, 100 ns resolution */
  uint64_t span = 730ULL * (uint64_t)TT_TICKS_PER_DAY;
  uint64_t h = tt_draw(c->seed, row, 3);
  return (uint64_t)c->today_local_ticks - 1 - (h % span);
}

TT_HD tt_synth_row tt_synth_make(const tt_synth_cfg* c, uint64_t row) {
  tt_synth_row r;
  const int64_t today_day = c->today_local_ticks / TT_TICKS_PER_DAY;
  const int64_t yday = today_day - 1;

  /* --- TaskDueDate --- */
  uint64_t h0 = tt_draw(c->seed, row, 0);
  uint64_t h1 = tt_draw(c->seed, row, 1);
  uint32_t u = tt_range(h0, 1000000u);
  int64_t day;
  uint64_t tod = 0;
  uint32_t kind = TT_KIND_UNSPECIFIED;
  if (u < c->ppm_due_yesterday) {
    day = yday;
  } else if (u < c->ppm_due_yesterday + c->ppm_adversarial) {
    day = yday;
    uint32_t variant = (uint32_t)(h1 & 3u);
    uint64_t t = 1 + (h1 >> 8) % (uint64_t)(TT_TICKS_PER_DAY - 1);
    if (variant == 0) { tod = t; }
    else if (variant == 1) { kind = TT_KIND_UTC; }
    else if (variant == 2) { kind = TT_KIND_LOCAL; }
    else { tod = t; kind = TT_KIND_UTC; }
  } else {
    /* uniform over [today-365, today+365] \ {yesterday}: 730 candidate days */
    day = today_day - 365 + (int64_t)tt_range(h1, 730u);
    if (day >= yday) day += 1;
    if (tt_range(h0 << 32, 1000000u) < c->ppm_timed_due) {
      tod = 1 + (h1 & 0xFFFFFFFFULL) * 201ULL % (uint64_t)(TT_TICKS_PER_DAY - 1);
      kind = TT_KIND_UTC;
    }
  }
  r.due_raw = ((uint64_t)day * (uint64_t)TT_TICKS_PER_DAY + tod) | ((uint64_t)kind << 62);

  /* --- flags --- */
  uint64_t h2 = tt_draw(c->seed, row, 2);
  r.completed = tt_range(h2, 1000000u) < c->ppm_completed;
  r.overdue = (day < yday) && (tt_range(h2 << 32, 1000000u) < c->ppm_overdue_past);

  /* --- TaskCreatedOn (Kind Utc), with forced duplicates of the previous row --- */
  uint64_t h4 = tt_draw(c->seed, row, 4);
  uint64_t src = (row > 0 && tt_range(h4, 1000000u) < c->ppm_created_dup) ? row - 1 : row;
  r.created_raw = tt_synth_created_base(c, src) | ((uint64_t)TT_KIND_UTC << 62);

  /* --- TaskCreatedBy --- */
  r.creator = tt_range(h4 << 32, c->n_creators ? c->n_creators : 1u);

  /* --- TaskId: v4 GUID whose top 48 bits grow with the row, so the table is sorted by
   *     TaskId and an equal-width TaskId range is an equal share of rows --- */
  uint64_t h5 = tt_draw(c->seed, row, 5);
  uint64_t h6 = tt_draw(c->seed, row, 6);
  uint64_t n = c->n_total ? c->n_total : 1;
  uint64_t stride = (1ULL << 48) / n;
  if (stride == 0) stride = 1;
  uint64_t top48 = row * stride + (h5 % stride);
  r.id_hi = (top48 << 16) | 0x4000ULL | ((h5 >> 48) & 0x0FFFULL);
  r.id_lo = (h6 & 0x3FFFFFFFFFFFFFFFULL) | 0x8000000000000000ULL;
  return r;
}

/* Creator string k: "user{k:07d}@mail{k%97}.example" (SURVEY.md §8d config 4).
 * Writes at most 40 bytes (no terminator); returns the length. */
TT_HD int tt_synth_creator(uint32_t k, char* out) {
  int p = 0;
  out[p++] = 'u'; out[p++] = 's'; out[p++] = 'e'; out[p++] = 'r';
  char tmp[10];
  int nd = 0;
  uint32_t v = k;
  do { tmp[nd++] = (char)('0' + v % 10); v /= 10; } while (v);
  for (int i = nd; i < 7; ++i) out[p++] = '0';
  while (nd) out[p++] = tmp[--nd];
  out[p++] = '@'; out[p++] = 'm'; out[p++] = 'a'; out[p++] = 'i'; out[p++] = 'l';
  v = k % 97;
  if (v >= 10) out[p++] = (char)('0' + v / 10);
  out[p++] = (char)('0' + v % 10);
  const char* tail = ".example";
  for (int i = 0; i < 8; ++i) out[p++] = tail[i];
  return p;
}

/* TaskName / TaskAssignedTo of a row, as indices into two small dictionaries (the strings never
 * reach the sweep or the filter; they only matter when a result is written out as JSON). */
#define TT_SYNTH_ASSIGNEES 50u
TT_HD uint32_t tt_synth_n_names(const tt_synth_cfg* c) { return c->n_names ? c->n_names : 1000000u; }
TT_HD uint32_t tt_synth_name_ref(const tt_synth_cfg* c, uint64_t row) { return (uint32_t)(row % tt_synth_n_names(c)); }
TT_HD uint32_t tt_synth_assignee_ref(const tt_synth_cfg* c, uint64_t row) {
  return tt_range(tt_draw(c->seed, row, 7), TT_SYNTH_ASSIGNEES);
}

TT_HD int tt_synth_put_uint(uint32_t v, char* out) {
  char tmp[10];
  int nd = 0, p = 0;
  do { tmp[nd++] = (char)('0' + v % 10); v /= 10; } while (v);
  while (nd) out[p++] = tmp[--nd];
  return p;
}

/* Name k: "Task number: {k}"; when k % 8 == 5: `Task "{k}" <urgent> & caf\u00e9 \\ {k}` + TAB, i.e. a
 * quote, HTML-sensitive characters, a two-byte UTF-8 character, a backslash and a control
 * character.  Writes at most 64 bytes (UTF-8, no terminator); returns the length. */
TT_HD int tt_synth_name(uint32_t k, char* out) {
  int p = 0;
  if (k % 8 != 5) {
    const char* head = "Task number: ";
    for (int i = 0; i < 13; ++i) out[p++] = head[i];
    p += tt_synth_put_uint(k, out + p);
    return p;
  }
  const char* a = "Task \"";
  for (int i = 0; i < 6; ++i) out[p++] = a[i];
  p += tt_synth_put_uint(k, out + p);
  const char* b = "\" <urgent> & caf";
  for (int i = 0; i < 16; ++i) out[p++] = b[i];
  out[p++] = (char)0xC3; out[p++] = (char)0xA9; /* U+00E9 */
  out[p++] = ' '; out[p++] = '\\'; out[p++] = ' ';
  p += tt_synth_put_uint(k, out + p);
  out[p++] = '\t';
  return p;
}

/* Assignee k: "assignee{k}@mail.com".  At most 24 bytes; returns the length. */
TT_HD int tt_synth_assignee(uint32_t k, char* out) {
  int p = 0;
  const char* a = "assignee";
  for (int i = 0; i < 8; ++i) out[p++] = a[i];
  p += tt_synth_put_uint(k, out + p);
  const char* b = "@mail.com";
  for (int i = 0; i < 9; ++i) out[p++] = b[i];
  return p;
}

/* 64-bit hash of the UTF-8 bytes of TaskCreatedBy: 8-byte little-endian chunks (zero padded
 * tail) folded through the splitmix64 finaliser, seeded with the length.  Fixed and seedless:
 * the value is part of the stored table format. */
TT_HD uint64_t tt_hash64(const uint8_t* p, uint32_t len) {
  uint64_t h = tt_mix64(0x7461736B73ULL ^ (uint64_t)len);
  uint32_t i = 0;
  while (i < len) {
    uint64_t c = 0;
    uint32_t m = len - i < 8 ? len - i : 8;
    for (uint32_t b = 0; b < m; ++b) c |= (uint64_t)p[i + b] << (8 * b);
    h = tt_mix64(h ^ c);
    i += 8;
  }
  return h;
}

#endif /* TT_SYNTH_H */
