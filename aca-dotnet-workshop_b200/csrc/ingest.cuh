// ingest.cuh — K7: bulk load of the state store's documents into the SoA table.
//
// The step right before the path.  The reference keeps every task as one JSON
// document under the key TaskId.ToString() (Api/Services/TasksStoreManager.cs:35 SaveStateAsync) and
// gets them back as TaskModel objects through System.Text.Json (GetStateAsync<TaskModel> :50,
// QueryStateAsync<TaskModel> :61,130; the queue handler receives the same shape,
// Svc/Controllers/ExternalTasksProcessorController.cs:29-33).  K7a restates that deserialisation
// for TaskModel (Api/Models/TaskModel.cs:5-12) with one thread per document:
//   * members in any order, any JSON whitespace between tokens, unknown members skipped (nested
//     objects / arrays included), a repeated member keeps its last value, a missing one keeps the
//     default (Guid.Empty, DateTime.MinValue, false, "");
//   * member names compared ignoring ASCII case (JsonSerializerDefaults.Web), escapes allowed;
//   * taskId: the "D" form, hex in either case;  isCompleted / isOverDue: true | false;
//   * taskCreatedOn / taskDueDate: ISO 8601 extended as System.Text.Json reads it — date only, or
//     date 'T' hh:mm[:ss[.f…]] with an optional 'Z' (Kind Utc) or ±hh[:mm] offset (converted to the
//     machine's local time, Kind Local); no suffix = Kind Unspecified; fractions beyond 7 digits
//     are truncated;
//   * strings: every JSON escape including \uXXXX surrogate pairs, unescaped IN PLACE in the
//     device copy of the text (unescaping only shrinks); null stays null (the ref becomes kRefNull).
// A document may carry one leading '[' or ',' and one trailing ',' or ']' so that a JSON array
// split at its element boundaries (the offsets K6 hands out) loads as it is.
// Anything else marks the document bad: the load reports the first bad document and changes
// nothing.
//
// Two kernels share the work: doc_parse_fast_kernel takes the documents that are in exactly the shape the
// reference's serializer writes (and judges no other), doc_parse_kernel — everything described above — takes
// the rest.
//
// The three string members are then dictionary-encoded on the device through a hash table keyed by the
// 64-bit hash of the (unescaped) string — the smallest document number of a string is its representative,
// every document's bytes are compared with its representative's (exact confirm, like the creator filter),
// a scan numbers the representatives in order of first appearance; two strings with one hash send the
// member to the round-1 encoder (sort by hash, compare neighbours).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tt_synth.h"
#include "json.cuh"
#include "radix.cuh"
#include "tt_bits.h"

namespace tt {

constexpr uint32_t kDocOk = 0;
constexpr uint32_t kDocSyntax = 1;  // not the JSON object shape described above
constexpr uint32_t kDocGuid = 2;    // taskId is not a "D" Guid
constexpr uint32_t kDocDate = 3;    // a date System.Text.Json would reject
constexpr uint32_t kDocString = 4;  // bad escape, raw control character or malformed UTF-8 in a string
constexpr uint32_t kDocType = 5;    // a known member with a value of the wrong type
constexpr uint32_t kDocDeferred = 100;  // internal: not in the serializer's own shape, left to the general parser

struct DocCols {
  uint4* id;
  uint64_t* created;
  uint64_t* due;
  uint8_t* flags;  // bit 0 IsCompleted, bit 1 IsOverDue
  uint64_t* str_start[3];  // 0 TaskName, 1 TaskCreatedBy, 2 TaskAssignedTo: offset of the unescaped bytes in the text
  uint32_t* str_len[3];
  uint64_t* str_hash[3];   // tt_hash64 of the unescaped bytes
  uint32_t* status;
};

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }

__device__ __forceinline__ const uint8_t* skip_ws(const uint8_t* p, const uint8_t* end) {
  while (p < end && is_ws(*p)) ++p;
  return p;
}

// ---- eight bytes at a time -----------------------------------------------------------------------
// The parser is one thread per document, and a byte-at-a-time state machine is a chain of dependent
// one-byte loads (an L1 round trip each: ncu's source page put most stall samples on the compare after
// LDG.E.U8).  The hot loops therefore look at eight bytes per step.
//
// load8: the eight bytes starting at p, any alignment, little-endian (two aligned 64-bit loads and a
// funnel shift).  It may read up to 15 bytes past p + 8 rounded up — the device copy of the text is
// allocated with 16 bytes of slack, and bytes beyond the caller's `end` are masked off by the callers.
__device__ __forceinline__ uint64_t load8(const uint8_t* p) {
  const uint64_t* q = reinterpret_cast<const uint64_t*>(reinterpret_cast<uintptr_t>(p) & ~(uintptr_t)7);
  const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(p) & 7u) * 8u;
  const uint64_t lo = q[0], hi = q[1];
  return sh ? (lo >> sh) | (hi << (64u - sh)) : lo;
}

constexpr uint64_t kOnes = 0x0101010101010101ull, kHighs = 0x8080808080808080ull;

// High bit set in every byte of x that equals c (exact for the LOWEST such byte; bytes above a match may be
// flagged by the borrow, which is harmless where only the first flagged byte is used).
__device__ __forceinline__ uint64_t bytes_eq(uint64_t x, uint32_t c) {
  const uint64_t y = x ^ (kOnes * c);
  return (y - kOnes) & ~y & kHighs;
}

// High bit set in the first byte of x that a plain string run must stop at: '"', '\\', a control
// character (< 0x20) or a non-ASCII byte (>= 0x80); possibly in later bytes too.
__device__ __forceinline__ uint64_t string_special(uint64_t x) {
  return (x & kHighs) | ((x - kOnes * 0x20u) & ~x & kHighs) | bytes_eq(x, '"') | bytes_eq(x, '\\');
}

// Number of leading bytes of [r, end) that are plain string bytes (none of the above).
__device__ __forceinline__ uint32_t plain_run(const uint8_t* r, const uint8_t* end) {
  uint32_t n = 0;
  while (r + n < end) {
    const uint32_t left = (uint32_t)min((ptrdiff_t)8, end - (r + n));
    const uint64_t m = string_special(load8(r + n));
    if (m) {
      const uint32_t k = (uint32_t)(__ffsll((long long)m) - 1) >> 3;
      return n + (k < left ? k : left);
    }
    n += left;
    if (left < 8) break;
  }
  return n;
}

// tt_hash64 (include/tt_synth.h) of [p, p + len), reading eight bytes per step: the same value.
__device__ __forceinline__ uint64_t hash_text(const uint8_t* p, uint32_t len) {
  uint64_t h = tt_mix64(0x7461736B73ULL ^ (uint64_t)len);
  for (uint32_t i = 0; i < len; i += 8) {
    uint64_t c = load8(p + i);
    const uint32_t m = len - i;
    if (m < 8) c &= (1ull << (8 * m)) - 1;
    h = tt_mix64(h ^ c);
  }
  return h;
}

__device__ __forceinline__ int hex_val(uint32_t c) {
  if (c >= '0' && c <= '9') return (int)c - '0';
  c |= 0x20;
  if (c >= 'a' && c <= 'f') return (int)c - 'a' + 10;
  return -1;
}

// \uXXXX at r (after the 'u'); returns the code unit or -1
__device__ __forceinline__ int read_u4(const uint8_t* r, const uint8_t* end) {
  if (end - r < 4) return -1;
  int v = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int h = hex_val(r[k]);
    if (h < 0) return -1;
    v = (v << 4) | h;
  }
  return v;
}

// One (possibly escaped) scalar of a JSON string starting at r (not the closing quote).  Writes its
// UTF-8 bytes to out[0..*nout) and returns the pointer after it, or nullptr (bad string).
__device__ const uint8_t* read_string_scalar(const uint8_t* r, const uint8_t* end, uint8_t* out, int* nout) {
  uint32_t c = *r++;
  if (c < 0x20) return nullptr;
  if (c != '\\') {
    int extra = 0;
    if (c >= 0x80) {
      if ((c >> 5) == 6) extra = 1;
      else if ((c >> 4) == 14) extra = 2;
      else if ((c >> 3) == 30) extra = 3;
      else return nullptr;
      if (end - r < extra) return nullptr;
    }
    out[0] = (uint8_t)c;
    for (int j = 0; j < extra; ++j) {
      const uint32_t d = r[j];
      if ((d >> 6) != 2) return nullptr;
      out[1 + j] = (uint8_t)d;
    }
    *nout = 1 + extra;
    return r + extra;
  }
  if (r >= end) return nullptr;
  c = *r++;
  uint32_t cp;
  switch (c) {
    case '"': cp = '"'; break;
    case '\\': cp = '\\'; break;
    case '/': cp = '/'; break;
    case 'b': cp = 8; break;
    case 'f': cp = 12; break;
    case 'n': cp = 10; break;
    case 'r': cp = 13; break;
    case 't': cp = 9; break;
    case 'u': {
      int u = read_u4(r, end);
      if (u < 0) return nullptr;
      r += 4;
      cp = (uint32_t)u;
      if (cp >= 0xD800 && cp <= 0xDBFF) {  // high surrogate: the low one must follow as \uXXXX
        if (end - r < 6 || r[0] != '\\' || r[1] != 'u') return nullptr;
        const int lo = read_u4(r + 2, end);
        if (lo < 0xDC00 || lo > 0xDFFF) return nullptr;
        r += 6;
        cp = 0x10000 + ((cp - 0xD800) << 10) + ((uint32_t)lo - 0xDC00);
      } else if (cp >= 0xDC00 && cp <= 0xDFFF) {
        return nullptr;
      }
      break;
    }
    default: return nullptr;
  }
  if (cp < 0x80) { out[0] = (uint8_t)cp; *nout = 1; }
  else if (cp < 0x800) { out[0] = (uint8_t)(0xC0 | (cp >> 6)); out[1] = (uint8_t)(0x80 | (cp & 0x3F)); *nout = 2; }
  else if (cp < 0x10000) {
    out[0] = (uint8_t)(0xE0 | (cp >> 12)); out[1] = (uint8_t)(0x80 | ((cp >> 6) & 0x3F)); out[2] = (uint8_t)(0x80 | (cp & 0x3F));
    *nout = 3;
  } else {
    out[0] = (uint8_t)(0xF0 | (cp >> 18)); out[1] = (uint8_t)(0x80 | ((cp >> 12) & 0x3F));
    out[2] = (uint8_t)(0x80 | ((cp >> 6) & 0x3F)); out[3] = (uint8_t)(0x80 | (cp & 0x3F));
    *nout = 4;
  }
  return r;
}

// A string value, unescaped in place.  p = first byte after the opening quote.  Returns the
// pointer after the closing quote (nullptr = bad string); [*start, *start + *len) holds the bytes.
__device__ const uint8_t* read_string_inplace(uint8_t* p, const uint8_t* end, uint8_t** start, uint32_t* len) {
  const uint8_t* r = p;
  uint8_t* w = p;
  while (true) {
    // fast path: a run of plain ASCII (the common case) moves one byte per step, or not at all
    // while nothing has been unescaped yet
    uint32_t c = 0;
    if (w == r) {  // nothing has been unescaped yet: the run stays where it is, eight bytes per step
      const uint32_t run = plain_run(r, end);
      r += run;
      w += run;
      if (r < end) c = *r;
    } else {
      // behind an escape the run moves down (w < r): its length from word-wise tests, its bytes eight per load
      const uint32_t run = plain_run(r, end);
      for (uint32_t k = 0; k < run; k += 8) {
        const uint64_t v = load8(r + k);
        const uint32_t left = run - k < 8u ? run - k : 8u;
        for (uint32_t j = 0; j < left; ++j) w[k + j] = (uint8_t)(v >> (8 * j));
      }
      r += run;
      w += run;
      if (r < end) c = *r;
    }
    if (r >= end) return nullptr;
    if (c == '"') break;
    uint8_t tmp[4];
    int n = 0;
    const uint8_t* nr = read_string_scalar(r, end, tmp, &n);
    if (!nr) return nullptr;
    if (w != r || c == '\\')  // nothing moves until the first escape has been seen
      for (int j = 0; j < n; ++j) w[j] = tmp[j];
    w += n;
    r = nr;
  }
  *start = p;
  *len = (uint32_t)(w - p);
  return r + 1;
}

// A short string (member name, Guid, date) unescaped into a local buffer.  *len = cap + 1 when it
// does not fit.  Returns the pointer after the closing quote, nullptr = bad string.
template <int CAP>
__device__ const uint8_t* read_string_local(const uint8_t* p, const uint8_t* end, uint8_t (&buf)[CAP], int* len) {
  const uint8_t* r = p;
  int n = 0;
  while (true) {
    if (r >= end) return nullptr;
    if (*r == '"') break;
    uint8_t tmp[4];
    int k = 0;
    r = read_string_scalar(r, end, tmp, &k);
    if (!r) return nullptr;
    for (int j = 0; j < k; ++j) {
      if (n < CAP) buf[n] = tmp[j];
      ++n;
    }
  }
  *len = n > CAP ? CAP + 1 : n;
  return r + 1;
}

// Skips any JSON value structurally (strings with their escapes, nesting by depth).
__device__ const uint8_t* skip_value(const uint8_t* p, const uint8_t* end) {
  int depth = 0;
  while (p < end) {
    const uint32_t c = *p;
    if (c == '"') {
      ++p;
      while (true) {
        if (p >= end) return nullptr;
        const uint32_t d = *p++;
        if (d == '"') break;
        if (d == '\\') {
          if (p >= end) return nullptr;
          ++p;
        }
      }
      if (depth == 0) return p;
    } else if (c == '{' || c == '[') {
      ++depth;
      ++p;
    } else if (c == '}' || c == ']') {
      if (depth == 0) return nullptr;
      ++p;
      if (--depth == 0) return p;
    } else if (depth == 0) {
      // number or literal: runs to the next delimiter
      const uint8_t* q = p;
      while (q < end && !is_ws(*q) && *q != ',' && *q != '}' && *q != ']') ++q;
      return q == p ? nullptr : q;
    } else {
      ++p;
    }
  }
  return nullptr;
}

__device__ __forceinline__ bool lit_eq(const uint8_t* p, const uint8_t* end, const char* lit, int n) {
  if (end - p < n) return false;
  for (int j = 0; j < n; ++j)
    if (p[j] != (uint8_t)lit[j]) return false;
  return true;
}

// 0 taskId, 1 taskName, 2 taskCreatedBy, 3 taskCreatedOn, 4 taskDueDate, 5 taskAssignedTo,
// 6 isCompleted, 7 isOverDue, -1 unknown.  ASCII case-insensitive.
__device__ int member_of(const uint8_t* name, int len) {
  const char* names[8] = {"taskid", "taskname", "taskcreatedby", "taskcreatedon", "taskduedate", "taskassignedto",
                          "iscompleted", "isoverdue"};
  const int lens[8] = {6, 8, 13, 13, 11, 14, 11, 9};
  for (int m = 0; m < 8; ++m) {
    if (lens[m] != len) continue;
    bool same = true;
    for (int j = 0; j < len && same; ++j) {
      uint32_t c = name[j];
      if (c >= 'A' && c <= 'Z') c |= 0x20;
      same = c == (uint8_t)names[m][j];
    }
    if (same) return m;
  }
  return -1;
}

// The member name whose opening quote is at p - 1, matched straight from the text when it holds
// no escape (the usual case): its length and one discriminating character pick the only
// candidate, then every character is compared (ASCII case folded).  *after = past the closing
// quote.  Returns the member number, -1 unknown, -2 bad string, -3 "has an escape, take the slow path".
__constant__ char kMemberNames[8][16] = {"taskid", "taskname", "taskcreatedby", "taskcreatedon", "taskduedate",
                                         "taskassignedto", "iscompleted", "isoverdue"};
__constant__ int kMemberLens[8] = {6, 8, 13, 13, 11, 14, 11, 9};
// the names as two little-endian words each (bytes 0..7 and 8..15, zero padded)
constexpr uint64_t name_word(const char* s, int from) {
  uint64_t w = 0;
  int n = 0;
  while (s[n]) ++n;
  for (int k = 0; k < 8; ++k)
    if (from + k < n) w |= (uint64_t)(uint8_t)s[from + k] << (8 * k);
  return w;
}
__constant__ uint64_t kMemberWord0[8] = {name_word("taskid", 0),       name_word("taskname", 0),    name_word("taskcreatedby", 0),
                                         name_word("taskcreatedon", 0), name_word("taskduedate", 0), name_word("taskassignedto", 0),
                                         name_word("iscompleted", 0),   name_word("isoverdue", 0)};
__constant__ uint64_t kMemberWord1[8] = {name_word("taskid", 8),       name_word("taskname", 8),    name_word("taskcreatedby", 8),
                                         name_word("taskcreatedon", 8), name_word("taskduedate", 8), name_word("taskassignedto", 8),
                                         name_word("iscompleted", 8),   name_word("isoverdue", 8)};
__device__ int member_fast(const uint8_t* p, const uint8_t* end, const uint8_t** after) {
  // One pass for the names the reference writes: "task?..." / "is?..." — the character after the
  // common prefix names the only candidate and so the length; the closing quote must sit right
  // behind it and every character must match.  Anything else takes the general route below.
  if (end - p >= 16) {
    int m = -1;
    const uint64_t r0 = load8(p), r1 = load8(p + 8);  // the whole name and its closing quote: no byte loads below
    const uint64_t f0 = r0 | (kOnes * 0x20u), f1 = r1 | (kOnes * 0x20u);
    const uint32_t c0 = (uint32_t)f0 & 0xFF;
    if (c0 == 't') {
      switch ((uint32_t)(f0 >> 32) & 0xFF) {
        case 'i': m = 0; break;
        case 'n': m = 1; break;
        case 'c': m = ((uint32_t)(f1 >> 32) & 0xFF) == 'y' ? 2 : 3; break;
        case 'd': m = 4; break;
        case 'a': m = 5; break;
        default: break;
      }
    } else if (c0 == 'i') {
      m = ((uint32_t)(f0 >> 16) & 0xFF) == 'c' ? 6 : 7;
    }
    if (m >= 0) {
      // folded to lower case by OR 0x20 (exact for a comparison with letters: only a letter can fold into a
      // letter) and compared a word at a time
      const int len = kMemberLens[m];
      const uint64_t w0 = kMemberWord0[m], w1 = kMemberWord1[m];
      const uint64_t m0 = len >= 8 ? ~0ull : (1ull << (8 * len)) - 1;
      const uint64_t m1 = len > 8 ? (1ull << (8 * (len - 8))) - 1 : 0ull;
      const uint32_t closing = (uint32_t)((len < 8 ? r0 >> (8 * len) : r1 >> (8 * (len - 8))) & 0xFF);
      if (closing == '"' && ((f0 ^ w0) & m0) == 0 && ((f1 ^ w1) & m1) == 0) {
        *after = p + len + 1;
        return m;
      }
    }
  }
  // general route: find the closing quote, then compare against the candidate of that length
  const uint8_t* q = p;
  while (q < end) {
    const uint32_t c = *q;
    if (c == '"') break;
    if (c == '\\') return -3;
    if (c < 0x20) return -2;
    ++q;
  }
  if (q >= end) return -2;
  *after = q + 1;
  const int len = (int)(q - p);
  int m;
  switch (len) {
    case 6: m = 0; break;
    case 8: m = 1; break;
    case 9: m = 7; break;
    case 11: m = (p[0] | 0x20) == 't' ? 4 : 6; break;
    case 13: m = (p[12] | 0x20) == 'y' ? 2 : 3; break;
    case 14: m = 5; break;
    default: return -1;
  }
  for (int j = 0; j < len; ++j) {
    uint32_t c = p[j];
    if (c >= 'A' && c <= 'Z') c |= 0x20;
    if (c != (uint8_t)kMemberNames[m][j]) return -1;
  }
  return m;
}

// The closing quote of a short string value that holds no escape: returns its length, or -1 when
// a backslash or control character comes first or it is longer than `limit`.
__device__ __forceinline__ int plain_string_len(const uint8_t* p, const uint8_t* end, int limit) {
  // a non-ASCII byte also ends the run: such a value takes the general route, which refuses it as a Guid / date
  const uint8_t* stop = end - p > limit + 1 ? p + limit + 1 : end;
  const uint32_t run = plain_run(p, stop);
  return p + run < stop && p[run] == '"' ? (int)run : -1;
}

// The first 48 bytes of a short string value (a Guid, a date) held in six registers: the parsers below index it
// like an array; an index known at compile time costs a shift, a computed one five selects — no memory access.
struct Chars48 {
  uint64_t w[6];
  __device__ __forceinline__ void load(const uint8_t* p, const uint8_t* end) {
#pragma unroll
    for (int k = 0; k < 6; ++k) w[k] = p + 8 * k < end ? load8(p + 8 * k) : 0ull;
  }
  __device__ __forceinline__ uint32_t operator[](int i) const {
    const uint64_t x = i < 8 ? w[0] : i < 16 ? w[1] : i < 24 ? w[2] : i < 32 ? w[3] : i < 40 ? w[4] : w[5];
    return (uint32_t)(x >> ((i & 7) * 8)) & 0xFFu;
  }
  // index of the first byte a plain run stops at ('"', '\\', control, non-ASCII), 48 if there is none
  __device__ __forceinline__ int first_special() const {
    int at = 48;
#pragma unroll
    for (int k = 5; k >= 0; --k) {
      const uint64_t m = string_special(w[k]);
      if (m) at = 8 * k + ((__ffsll((long long)m) - 1) >> 3);
    }
    return at;
  }
};

// the same interface over bytes in memory (the slow path's local buffer)
struct CharsMem {
  const uint8_t* p;
  __device__ __forceinline__ uint32_t operator[](int i) const { return p[i]; }
};

// "xxxxxxxx-xxxx-xxxx-xxxx-xxxxxxxxxxxx" -> the 16 canonical bytes packed in memory order
template <typename S>
__device__ __forceinline__ bool parse_guid(const S& s, int len, uint4* out) {
  if (len != 36) return false;
  uint32_t w[4] = {0, 0, 0, 0};
  int b = 0;
  for (int i = 0; i < 36;) {
    if (i == 8 || i == 13 || i == 18 || i == 23) {
      if (s[i] != '-') return false;
      ++i;
      continue;
    }
    const int hi = hex_val(s[i]), lo = hex_val(s[i + 1]);
    if (hi < 0 || lo < 0) return false;
    w[b >> 2] |= (uint32_t)((hi << 4) | lo) << (8 * (b & 3));
    ++b;
    i += 2;
  }
  *out = make_uint4(w[0], w[1], w[2], w[3]);
  return true;
}

// Eight ASCII hex digits (either case) -> their values as four bytes, the first pair in the lowest byte; false when a
// byte is not a hex digit.  Per byte, without carries between bytes (every addend is below 0x80):
// d = x ^ '0' is 0..9 for a digit, l = (x | 0x20) ^ 0x60 is 1..6 for a letter a-f / A-F.
__device__ __forceinline__ bool hex8(uint64_t x, uint32_t* out) {
  constexpr uint64_t k7F = kOnes * 0x7Fu;
  const uint64_t d = x ^ (kOnes * 0x30u);
  const uint64_t l = (x | (kOnes * 0x20u)) ^ (kOnes * 0x60u);
  const uint64_t dig = ~(((d & k7F) + kOnes * 0x76u) | d) & kHighs;                                  // d <= 9
  const uint64_t let = ~(((l & k7F) + kOnes * 0x79u) | l) & ((l & k7F) + k7F) & kHighs;              // 1 <= l <= 6
  if ((dig | let) != kHighs) return false;
  const uint64_t nib = (x & (kOnes * 0x0Fu)) + (let >> 7) * 9u;
  const uint64_t y = (nib << 4) | (nib >> 8);  // even bytes: first digit of a pair << 4 | second
  *out = __byte_perm((uint32_t)y, (uint32_t)(y >> 32), 0x6420);
  return true;
}

// The usual Guid — 36 characters in registers: four words of eight hex digits each, cut out around the dashes.
__device__ __forceinline__ bool parse_guid36(const Chars48& s, uint4* out) {
  const uint64_t w0 = s.w[0], w1 = s.w[1], w2 = s.w[2], w3 = s.w[3], w4 = s.w[4];
  const bool dashes = ((uint32_t)w1 & 0xFFu) == '-' && ((uint32_t)(w1 >> 40) & 0xFFu) == '-' &&
                      ((uint32_t)(w2 >> 16) & 0xFFu) == '-' && (uint32_t)(w2 >> 56) == '-';
  if (!dashes) return false;
  const uint64_t h1 = ((w1 >> 8) & 0xFFFFFFFFull) | ((w1 >> 48) << 32) | (w2 << 48);  // characters 9-12, 14-15, 16-17
  const uint64_t h2 = ((w2 >> 24) & 0xFFFFFFFFull) | (w3 << 32);                       // 19-22, 24-27
  const uint64_t h3 = (w3 >> 32) | (w4 << 32);                                         // 28-35
  uint4 id;
  if (!hex8(w0, &id.x) || !hex8(h1, &id.y) || !hex8(h2, &id.z) || !hex8(h3, &id.w)) return false;
  *out = id;
  return true;
}

template <typename S>
__device__ __forceinline__ int two_digits(const S& s, int at) {
  const int a = (int)s[at] - '0', b = (int)s[at + 1] - '0';
  if (a < 0 || a > 9 || b < 0 || b > 9) return -1;
  return a * 10 + b;
}

// System.Text.Json's DateTime reader (ISO 8601-1 extended profile), see the header comment.
template <typename S>
__device__ bool parse_datetime(const S& s, int len, int32_t local_offset_min, uint64_t* raw) {
  if (len < 10 || len > 48) return false;
  const int yh = two_digits(s, 0), yl = two_digits(s, 2), mo = two_digits(s, 5), d = two_digits(s, 8);
  if (yh < 0 || yl < 0 || mo < 0 || d < 0 || s[4] != '-' || s[7] != '-') return false;
  const int y = yh * 100 + yl;
  if (y < 1 || mo < 1 || mo > 12 || d < 1) return false;
  const bool leap = (y % 4 == 0 && y % 100 != 0) || y % 400 == 0;
  const int dim = mo == 2 ? (leap ? 29 : 28) : (mo == 4 || mo == 6 || mo == 9 || mo == 11) ? 30 : 31;
  if (d > dim) return false;
  // days since 0001-01-01 (proleptic Gregorian)
  const int yy = mo <= 2 ? y - 1 : y;
  const int era = yy / 400;
  const int yoe = yy - era * 400;
  const int doy = (153 * (mo + (mo > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const int doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  const int64_t days = (int64_t)era * 146097 + doe - 306;
  int64_t ticks = days * 864000000000ll;
  uint32_t kind = 0;
  int i = 10;
  if (i < len) {
    if (s[i] != 'T' || len < 16) return false;
    const int hh = two_digits(s, 11), mi = two_digits(s, 14);
    if (hh < 0 || hh > 23 || mi < 0 || mi > 59 || s[13] != ':') return false;
    int ss = 0;
    int64_t frac = 0;
    i = 16;
    if (i < len && s[i] == ':') {
      if (len < 19) return false;
      ss = two_digits(s, 17);
      if (ss < 0 || ss > 59) return false;
      i = 19;
      if (i < len && (s[i] == '.' || s[i] == ',')) {
        ++i;
        int nd = 0;
        while (i < len && s[i] >= '0' && s[i] <= '9') {
          if (nd < 7) frac = frac * 10 + (s[i] - '0');
          ++nd;
          ++i;
        }
        if (nd == 0 || nd > 16) return false;
        for (; nd < 7; ++nd) frac *= 10;
      }
    }
    ticks += ((int64_t)hh * 3600 + mi * 60 + ss) * 10000000ll + frac;
    if (i < len) {
      if (s[i] == 'Z') {
        kind = 1;
        ++i;
      } else if (s[i] == '+' || s[i] == '-') {
        const bool neg = s[i] == '-';
        if (len - i < 3) return false;
        const int oh = two_digits(s, i + 1);
        int om = 0;
        i += 3;
        if (i < len) {
          if (s[i] != ':' || len - i < 3) return false;
          om = two_digits(s, i + 1);
          i += 3;
        }
        if (oh < 0 || oh > 14 || om < 0 || om > 59) return false;
        const int off = (oh * 60 + om) * (neg ? -1 : 1);
        // the instant, shown on the machine's local clock (DateTimeOffset.LocalDateTime), Kind Local
        ticks += (int64_t)(local_offset_min - off) * 600000000ll;
        kind = 2;
      }
      if (i != len) return false;
    }
  }
  if (ticks < 0 || ticks > 3155378975999999999ll) return false;
  *raw = (uint64_t)ticks | ((uint64_t)kind << 62);
  return true;
}

// K7a: one thread per document.
// `text` holds the caller's bytes [window_first, window_first + bytes); doc_off are the caller's offsets.  A document
// whose range is not inside that window, or runs backwards, is a bad document.
//
// Where the bytes come from.  Thread i reads document i, so the lanes of a warp read addresses ~300 bytes apart and
// every load instruction is 32 separate L1 accesses: at ~100 loads per document the kernel was bound by the load/store
// unit, not by instructions or DRAM (shorter instruction streams did not make it faster).  A warp's 32 documents are
// one contiguous piece of the text (~9.4 KB), so the warp first copies that piece into its own part of shared memory
// with coalesced 16-byte loads (same 16-byte alignment as in global memory), and every lane parses — and unescapes in
// place — its document THERE; strings that were unescaped are copied back by their lane (few and short), offsets are
// reported relative to the text as before.  A warp whose piece does not fit (stage_bytes per warp; very long
// documents) or whose offsets are out of order parses in global memory as before.
template <int MIN_CTAS>
__global__ void __launch_bounds__(128, MIN_CTAS) doc_parse_kernel(uint8_t* text, const uint64_t* doc_off, uint64_t window_first,
                                                        uint64_t bytes, int64_t n, int32_t local_offset_min,
                                                        DocCols out, uint32_t stage_bytes, const uint32_t* todo,
                                                        int64_t n_todo) {
  extern __shared__ __align__(16) uint8_t s_text[];
  // todo != null: thread j parses document todo[j] (the ones the fast kernel left)
  const int64_t j0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t i = todo ? (j0 < n_todo ? (int64_t)todo[j0] : n) : j0;
  const uint32_t lane = threadIdx.x & 31u;
  const bool live = i < n;
  const uint64_t a = live ? doc_off[i] : 0, b = live ? doc_off[i + 1] : 0;
  const bool in_window = live && a >= window_first && b >= a && b - window_first <= bytes;
  // the warp's piece of the text: [lo, hi) relative to the window, copied from the 16-byte boundary below lo
  uint64_t lo = in_window ? a - window_first : ~0ull, hi = in_window ? b - window_first : 0ull;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const uint64_t l2 = __shfl_xor_sync(0xFFFFFFFFu, lo, d), h2 = __shfl_xor_sync(0xFFFFFFFFu, hi, d);
    lo = l2 < lo ? l2 : lo;
    hi = h2 > hi ? h2 : hi;
  }
  const uint64_t span_start = lo & ~15ull;
  const uint64_t copy_len = hi > span_start ? ((hi - span_start + 15) & ~15ull) : 0;
  uint8_t* const s_warp = s_text + (size_t)(threadIdx.x >> 5) * stage_bytes;
  bool staged;
  if (todo == nullptr) {
    staged = stage_bytes != 0 && __all_sync(0xFFFFFFFFu, !live || in_window) && lo != ~0ull && copy_len + 16 <= stage_bytes;
    if (staged) {
      const uint4* g = reinterpret_cast<const uint4*>(text + span_start);
      uint4* d = reinterpret_cast<uint4*>(s_warp);
      for (uint32_t o = lane; o < (uint32_t)(copy_len >> 4); o += 32) d[o] = g[o];
      __syncwarp();
    }
  } else {
    // The documents of a list lie anywhere and are parsed in global memory.  Staging each on its own was measured twice
    // (one after the other, and four documents' loads in flight): 10.8 ms for the whole parse against 6.8 — the documents
    // that come here are the ones with escapes, whose unescaping wants many resident warps more than near bytes.
    staged = false;
  }
  if (!live) return;
  // `base`: the address offset 0 of the text has in the space this lane parses in
  uint8_t* const base = staged ? s_warp - span_start : text;
  const uint8_t* p = base + (in_window ? a - window_first : 0);
  const uint8_t* const end = base + (in_window ? b - window_first : 0);
  uint4 id = make_uint4(0, 0, 0, 0);
  uint64_t created = 0, due = 0;
  uint32_t flags = 0;
  uint8_t* sstart[3] = {base, base, base};
  uint32_t slen[3] = {0, 0, 0};
  uint32_t status = kDocOk;

  // `peek`: the byte at p, or 0 at the end, whitespace skipped first — ONE load per token in the usual document (which has
  // no whitespace); the value is kept in a register, because the text is also written through (unescaping in place) and
  // the compiler would otherwise read *p again for every test.
  auto peek = [&](const uint8_t*& q) -> uint32_t {
    uint32_t c = q < end ? *q : 0u;
    if (c <= ' ' && c != 0u) {
      q = skip_ws(q, end);
      c = q < end ? *q : 0u;
    }
    return c;
  };
  uint32_t c = peek(p);
  if (c == '[' || c == ',') {
    ++p;
    c = peek(p);
  }
  if (!in_window || c != '{') status = kDocSyntax;
  else ++p;
  bool first = true;
  while (status == kDocOk) {
    c = peek(p);
    if (c == 0u) { status = kDocSyntax; break; }
    if (c == '}') { ++p; break; }
    if (!first) {
      if (c != ',') { status = kDocSyntax; break; }
      ++p;
      c = peek(p);
    }
    first = false;
    if (c != '"') { status = kDocSyntax; break; }
    const uint8_t* after = nullptr;
    int m = member_fast(p + 1, end, &after);
    if (m == -3) {  // an escaped name: unescape it into a local buffer first
      uint8_t name[16];
      int nlen = 0;
      after = read_string_local(p + 1, end, name, &nlen);
      m = !after ? -2 : nlen <= 16 ? member_of(name, nlen) : -1;
    }
    if (m == -2) { status = kDocString; break; }
    p = after;
    if (peek(p) != ':') { status = kDocSyntax; break; }
    ++p;
    c = peek(p);
    if (c == 0u) { status = kDocSyntax; break; }
    if (m < 0) {
      p = skip_value(p, end);
      if (!p) status = kDocSyntax;
      continue;
    }
    if (m == 6 || m == 7) {
      const uint32_t bit = m == 6 ? 1u : 2u;
      if (c == 't' && lit_eq(p, end, "true", 4)) { flags |= bit; p += 4; }
      else if (c == 'f' && lit_eq(p, end, "false", 5)) { flags &= ~bit; p += 5; }
      else status = kDocType;
      continue;
    }
    if (m == 1 || m == 2 || m == 5) {
      const int col = m == 1 ? 0 : m == 2 ? 1 : 2;
      // null stays null (flag bits 4..6 of the document's flag byte); a later repeat of the member overwrites it
      if (c == 'n' && lit_eq(p, end, "null", 4)) { slen[col] = 0; sstart[col] = base; flags |= 16u << col; p += 4; continue; }
      flags &= ~(16u << col);
      if (c != '"') { status = kDocType; continue; }
      p = read_string_inplace(const_cast<uint8_t*>(p) + 1, end, &sstart[col], &slen[col]);
      if (!p) status = kDocString;
      else if (staged && (uint32_t)(p - 1 - sstart[col]) != slen[col]) {
        // unescaped (it shrank): the text in global memory gets the bytes later kernels read at this offset
        uint8_t* g = text + (sstart[col] - base);
        for (uint32_t j = 0; j < slen[col]; ++j) g[j] = sstart[col][j];
      }
      continue;
    }
    // taskId and the two dates: short strings parsed from a local copy
    if (c != '"') { status = kDocType; continue; }
    Chars48 cs;  // the value's first 48 bytes in registers: Guid and dates are parsed without another load
    cs.load(p + 1, end);
    if (m == 0 && end - p > 37 && cs[36] == '"' && parse_guid36(cs, &id)) {  // the usual 36 plain characters
      p += 38;
      continue;
    }
    int plain = cs.first_special();
    if (plain >= 48 || p + 1 + plain >= end || cs[plain] != '"') plain = plain_string_len(p + 1, end, 48);
    if (plain >= 0) {  // no escape inside: parsed where it lies
      if (m == 0) {
        if (plain != 36 || !parse_guid36(cs, &id)) status = kDocGuid;
      } else {
        uint64_t raw = 0;
        if (!parse_datetime(cs, plain, local_offset_min, &raw)) status = kDocDate;
        else if (m == 3) created = raw;
        else due = raw;
      }
      p += plain + 2;
      continue;
    }
    uint8_t buf[48];
    int blen = 0;
    p = read_string_local(p + 1, end, buf, &blen);
    if (!p) { status = kDocString; continue; }
    const CharsMem bs{buf};
    if (m == 0) {
      if (!parse_guid(bs, blen, &id)) status = kDocGuid;
    } else {
      uint64_t raw = 0;
      if (blen > 48 || !parse_datetime(bs, blen, local_offset_min, &raw)) status = kDocDate;
      else if (m == 3) created = raw;
      else due = raw;
    }
  }
  if (status == kDocOk) {
    c = peek(p);
    if (c == ',' || c == ']') {
      ++p;
      peek(p);
    }
    if (p != end) status = kDocSyntax;
  }
  out.status[i] = status;
  out.id[i] = id;
  out.created[i] = created;
  out.due[i] = due;
  out.flags[i] = (uint8_t)flags;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    out.str_start[c][i] = (uint64_t)(sstart[c] - base);
    out.str_len[c][i] = slen[c];
    out.str_hash[c][i] = status == kDocOk ? hash_text(sstart[c], slen[c]) : 0ull;
  }
}

// The date form the serializer itself writes — yyyy-MM-ddTHH:mm:ss, then nothing, or .f to .fffffff, then nothing, Z or
// ±hh:mm — parsed from the register window with fixed positions: separators and digits are checked a word at a time,
// the two-digit numbers come out of one multiplication per word.  Same range checks and the same arithmetic as
// parse_datetime; returns false for every other form (and for values out of range), which the caller then hands to
// parse_datetime.
__device__ __forceinline__ bool parse_datetime_fixed(const Chars48& s, int len, int32_t local_offset_min, uint64_t* raw) {
  if (len < 19) return false;
  constexpr uint64_t k7F = kOnes * 0x7Fu;
  // pattern words: '0' where a digit stands, the separator itself elsewhere ("0000-00-", "00T00:00", ":00" + don't care)
  constexpr uint64_t p0 = 0x2D30302D30303030ull, p1 = 0x30303A3030543030ull, p2 = 0x30303Aull;
  const uint64_t x0 = s.w[0] ^ p0, x1 = s.w[1] ^ p1, x2 = (s.w[2] & 0xFFFFFFull) ^ p2;
  // every byte must be 0..9 (separators: 0)
  const uint64_t bad = ((((x0 & k7F) + kOnes * 0x76u) | x0) | (((x1 & k7F) + kOnes * 0x76u) | x1) | (((x2 & k7F) + kOnes * 0x76u) | x2)) & kHighs;
  const uint64_t sep = (x0 & 0xFF0000FF00000000ull) | (x1 & 0x0000FF0000FF0000ull) | (x2 & 0xFFull);
  if (bad | sep) return false;
  const uint64_t t0 = x0 * 10u + (x0 >> 8), t1 = x1 * 10u + (x1 >> 8), t2 = x2 * 10u + (x2 >> 8);  // byte k = 10 d[k] + d[k+1]
  const int yh = (int)(t0 & 0xFF), yl = (int)((t0 >> 16) & 0xFF), mo = (int)((t0 >> 40) & 0xFF);
  const int d = (int)(t1 & 0xFF), hh = (int)((t1 >> 24) & 0xFF), mi = (int)((t1 >> 48) & 0xFF), ss = (int)((t2 >> 8) & 0xFF);
  const int y = yh * 100 + yl;
  if (y < 1 || mo < 1 || mo > 12 || d < 1) return false;
  const bool leap = (y % 4 == 0 && y % 100 != 0) || y % 400 == 0;
  const int dim = mo == 2 ? (leap ? 29 : 28) : (mo == 4 || mo == 6 || mo == 9 || mo == 11) ? 30 : 31;
  if (d > dim || hh > 23 || mi > 59 || ss > 59) return false;
  const int yy = mo <= 2 ? y - 1 : y;
  const int era = yy / 400;
  const int yoe = yy - era * 400;
  const int doy = (153 * (mo + (mo > 2 ? -3 : 9)) + 2) / 5 + d - 1;
  const int doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  const int64_t days = (int64_t)era * 146097 + doe - 306;
  int64_t ticks = days * 864000000000ll + ((int64_t)hh * 3600 + mi * 60 + ss) * 10000000ll;
  int at = 19;
  if (len > 19 && ((uint32_t)(s.w[2] >> 24) & 0xFFu) == '.') {
    // up to seven fraction digits at characters 20..26; an eighth digit means a longer fraction: not this form
    const uint64_t f = ((s.w[2] >> 32) | (s.w[3] << 32)) ^ (kOnes * 0x30u);             // characters 20..27 as digit values
    const uint64_t nondigit = ((((f & k7F) + kOnes * 0x76u) | f) & kHighs);
    const int nd = nondigit ? (__ffsll((long long)nondigit) - 1) >> 3 : 8;
    if (nd == 0 || nd > 7 || 20 + nd > len) return false;
    int64_t frac = 0;
#pragma unroll
    for (int j = 0; j < 7; ++j) frac = j < nd ? frac * 10 + (int64_t)((f >> (8 * j)) & 0xFF) : frac * 10;
    ticks += frac;
    at = 20 + nd;
  }
  uint32_t kind = 0;
  if (at < len) {
    // the suffix: the (at most six) characters from `at` on
    const int wi = at >> 3, sh = (at & 7) * 8;
    const uint64_t lo = wi == 2 ? s.w[2] : wi == 3 ? s.w[3] : s.w[4], hi = wi == 2 ? s.w[3] : wi == 3 ? s.w[4] : s.w[5];
    const uint64_t z = sh ? (lo >> sh) | (hi << (64 - sh)) : lo;
    const uint32_t c0 = (uint32_t)z & 0xFFu;
    if (len - at == 1 && c0 == 'Z') {
      kind = 1;
    } else if (len - at == 6 && (c0 == '+' || c0 == '-') && ((uint32_t)(z >> 24) & 0xFFu) == ':') {
      const uint32_t a1 = ((uint32_t)(z >> 8) & 0xFFu) - '0', a2 = ((uint32_t)(z >> 16) & 0xFFu) - '0';
      const uint32_t b1 = ((uint32_t)(z >> 32) & 0xFFu) - '0', b2 = ((uint32_t)(z >> 40) & 0xFFu) - '0';
      if (a1 > 9 || a2 > 9 || b1 > 9 || b2 > 9) return false;
      const int oh = (int)(a1 * 10 + a2), om = (int)(b1 * 10 + b2);
      if (oh > 14 || om > 59) return false;
      const int off = (oh * 60 + om) * (c0 == '-' ? -1 : 1);
      ticks += (int64_t)(local_offset_min - off) * 600000000ll;
      kind = 2;
    } else {
      return false;
    }
  }
  if (ticks < 0 || ticks > 3155378975999999999ll) return false;
  *raw = (uint64_t)ticks | ((uint64_t)kind << 62);
  return true;
}

// parse_datetime out of line: the fast kernel calls it only for dates in another form
__device__ __noinline__ bool parse_datetime_other(const Chars48* s, int len, int32_t local_offset_min, uint64_t* raw) {
  return parse_datetime(*s, len, local_offset_min, raw);
}

// ---- K7a, the fast kernel: documents in the serializer's own shape -------------------------------------------------
// What the state store holds was written by the reference's own serializer (JsonSerializer.Serialize(TaskModel), camelCase
// names in declaration order, no whitespace).  For exactly that shape nothing has to be searched: every literal between the
// values is compared as whole words against compile-time constants, the values are cut out by their closing quotes, and the
// Guid, the dates and the hashes are computed by the same routines as in the general kernel.  All lanes of a warp run the
// same straight-line code (a document's bytes are in shared memory, staged as in the general kernel).  A document that
// deviates in ANY way — a backslash, a non-ASCII or control byte in a string, null, another order, whitespace, a value the
// routines refuse — is not judged here: it gets kDocDeferred and the general kernel parses it (doc_deferred_list_kernel
// collects those documents).  The general kernel costs ~7 000 warp instructions per 32 documents, this one a fraction.
template <int N>
struct LitWords8 {
  uint64_t w[(N + 6) / 8 + 1];
};
template <int N>
__host__ __device__ constexpr LitWords8<N> lit_words8(const char (&lit)[N]) {
  LitWords8<N> r = {};
  for (int k = 0; k < N - 1; ++k) r.w[k / 8] |= (uint64_t)(uint8_t)lit[k] << (8 * (k % 8));
  return r;
}
// [p, p + N - 1) equals the literal (p + N - 1 <= end is the caller's business; load8 may read a few bytes further)
template <int N>
__device__ __forceinline__ bool lit_match(const uint8_t* p, const LitWords8<N> lw) {
  constexpr int len = N - 1;
  uint64_t diff = 0;
#pragma unroll
  for (int j = 0; j * 8 < len; ++j) {
    uint64_t x = load8(p + 8 * j) ^ lw.w[j];
    if (len - 8 * j < 8) x &= (1ull << (8 * (len - 8 * j))) - 1;
    diff |= x;
  }
  return diff == 0;
}
#define TT_LIT_MATCH(P, LIT) ([&] { constexpr auto lw_ = lit_words8(LIT); return lit_match(P, lw_); }())

template <int MIN_CTAS>
__global__ void __launch_bounds__(128, MIN_CTAS) doc_parse_fast_kernel(uint8_t* text, const uint64_t* doc_off,
                                                                       uint64_t window_first, uint64_t bytes, int64_t n,
                                                                       int32_t local_offset_min, DocCols out,
                                                                       uint32_t stage_bytes, uint32_t* todo,
                                                                       unsigned long long* todo_count) {
  extern __shared__ __align__(16) uint8_t s_text[];
  __shared__ uint32_t s_defer[2];  // deferred documents of this CTA, and where its piece of the list starts
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31u;
  const bool live = i < n;
  if (threadIdx.x == 0) s_defer[0] = 0;
  const uint64_t a = live ? doc_off[i] : 0, b = live ? doc_off[i + 1] : 0;
  const bool in_window = live && a >= window_first && b >= a && b - window_first <= bytes;
  uint64_t lo = in_window ? a - window_first : ~0ull, hi = in_window ? b - window_first : 0ull;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const uint64_t l2 = __shfl_xor_sync(0xFFFFFFFFu, lo, d), h2 = __shfl_xor_sync(0xFFFFFFFFu, hi, d);
    lo = l2 < lo ? l2 : lo;
    hi = h2 > hi ? h2 : hi;
  }
  const uint64_t span_start = lo & ~15ull;
  const uint64_t copy_len = hi > span_start ? ((hi - span_start + 15) & ~15ull) : 0;
  const bool staged = __all_sync(0xFFFFFFFFu, !live || in_window) && lo != ~0ull && copy_len + 64 <= stage_bytes;
  uint8_t* const s_warp = s_text + (size_t)(threadIdx.x >> 5) * stage_bytes;
  if (staged) {
    const uint4* g = reinterpret_cast<const uint4*>(text + span_start);
    uint4* d = reinterpret_cast<uint4*>(s_warp);
    for (uint32_t o = lane; o < (uint32_t)(copy_len >> 4); o += 32) d[o] = g[o];
    __syncwarp();
  }
  // (a warp whose documents do not fit its stage, or lie outside the window, leaves them all to the general kernel)
  uint8_t* const base = s_warp - span_start;
  const uint8_t* p = base + (live && staged ? a - window_first : span_start);
  const uint8_t* const end = base + (live && staged ? b - window_first : span_start);
  // the shortest document of this shape: 11 + 36 + 14 + 19 + 19 + 10 + 17 + 10 + 20 + 16 + 4 + 13 + 4 + 1 bytes
  bool ok = live && staged && end - p >= 194;
  uint32_t c = ok ? *p : 0u;
  if (c == '[' || c == ',') ++p;
  ok = ok && TT_LIT_MATCH(p, "{\"taskId\":\"");
  p += 11;
  uint4 id = make_uint4(0, 0, 0, 0);
  uint64_t created = 0, due = 0;
  const uint8_t* sstart[3] = {base, base, base};
  uint32_t slen[3] = {0, 0, 0};
  uint32_t flags = 0;
  Chars48 cs;
  if (ok) {
    cs.load(p, end);
    ok = cs[36] == '"' && parse_guid36(cs, &id);
    p += 36;
  }
  // a string member: [p, p + run) up to the closing quote; anything but a quote at the end of the plain run defers.
  // (Unescaping here was measured: the few lanes with escapes hold their warp for 2 400 instructions, 7.7 ms for 25 M
  // documents against 4.5 + 2.4 ms when those documents go to the general kernel.)
  auto string_value = [&](int col) {
    const uint32_t run = plain_run(p, end);
    ok = ok && p + run < end && p[run] == '"';
    sstart[col] = p;
    slen[col] = run;
    p += run;
  };
  auto date_value = [&](uint64_t* raw) {
    cs.load(p, end);
    const int plain = cs.first_special();
    ok = ok && plain < 48 && p + plain < end && cs[plain] == '"' &&
         (parse_datetime_fixed(cs, plain, local_offset_min, raw) || parse_datetime_other(&cs, plain, local_offset_min, raw));
    p += plain < 48 ? plain : 0;
  };
  auto bool_value = [&](uint32_t bit) {
    const uint64_t w = load8(p);
    if ((uint32_t)w == 0x65757274u) {  // "true"
      flags |= bit;
      p += 4;
    } else if ((w & 0xFFFFFFFFFFull) == 0x65736C6166ull) {  // "false"
      p += 5;
    } else {
      ok = false;
    }
  };
  if (ok) ok = end - p >= 14 && TT_LIT_MATCH(p, "\",\"taskName\":\"");
  if (ok) { p += 14; string_value(0); }
  if (ok) ok = end - p >= 19 && TT_LIT_MATCH(p, "\",\"taskCreatedBy\":\"");
  if (ok) { p += 19; string_value(1); }
  if (ok) ok = end - p >= 19 && TT_LIT_MATCH(p, "\",\"taskCreatedOn\":\"");
  if (ok) { p += 19; date_value(&created); }
  if (ok) ok = end - p >= 17 && TT_LIT_MATCH(p, "\",\"taskDueDate\":\"");
  if (ok) { p += 17; date_value(&due); }
  if (ok) ok = end - p >= 20 && TT_LIT_MATCH(p, "\",\"taskAssignedTo\":\"");
  if (ok) { p += 20; string_value(2); }
  if (ok) ok = end - p >= 16 + 4 && TT_LIT_MATCH(p, "\",\"isCompleted\":");
  if (ok) { p += 16; bool_value(1u); }
  if (ok) ok = end - p >= 13 + 4 && TT_LIT_MATCH(p, ",\"isOverDue\":");
  if (ok) { p += 13; bool_value(2u); }
  if (ok) {
    ok = p < end && *p == '}';
    ++p;
    if (ok && p < end && (*p == ',' || *p == ']')) ++p;
    ok = ok && p == end;
  }
  if (ok) {
    out.status[i] = kDocOk;
    out.id[i] = id;
    out.created[i] = created;
    out.due[i] = due;
    out.flags[i] = (uint8_t)flags;
#pragma unroll
    for (int col = 0; col < 3; ++col) {
      out.str_start[col][i] = (uint64_t)(sstart[col] - base);
      out.str_len[col][i] = slen[col];
      out.str_hash[col][i] = hash_text(sstart[col], slen[col]);
    }
  } else if (live) {
    out.status[i] = kDocDeferred;
  }
  // the list of deferred documents: one atomic per CTA on the global counter (any order)
  const bool mine = live && !ok;
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, mine);
  __syncthreads();  // s_defer[0] = 0 is visible
  uint32_t woff = 0;
  if (m && lane == 0) woff = atomicAdd(&s_defer[0], (uint32_t)__popc(m));
  woff = __shfl_sync(0xFFFFFFFFu, woff, 0);
  __syncthreads();
  if (threadIdx.x == 0 && s_defer[0]) s_defer[1] = (uint32_t)atomicAdd(todo_count, (unsigned long long)s_defer[0]);
  __syncthreads();
  if (mine) todo[s_defer[1] + woff + __popc(m & ((1u << lane) - 1u))] = (uint32_t)i;
}

// first bad document: err[0] = min index (init ~0), err[1] = its reason (written by the owner of the min in a second pass)
__global__ void doc_first_bad_kernel(const uint32_t* status, int64_t n, unsigned long long* err) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && status[i] != kDocOk) atomicMin(err, (unsigned long long)i);
}

// The fixed-width columns of the table from the parsed documents (a warp covers one flag word).
__global__ void doc_fill_kernel(DocCols d, int64_t n, uint32_t* due_code, uint32_t* completed, uint32_t* overdue,
                                uint64_t* creator_hash, uint4* task_id, uint64_t* created_raw, uint64_t* due_raw) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = i < n;
  uint32_t f = 0;
  if (live) {
    f = d.flags[i];
    const uint64_t du = d.due[i];
    due_code[i] = tt_pack_due(du);
    due_raw[i] = du;
    created_raw[i] = d.created[i];
    task_id[i] = d.id[i];
    creator_hash[i] = d.str_hash[1][i];
  }
  const uint32_t cw = __ballot_sync(0xFFFFFFFFu, f & 1u), ow = __ballot_sync(0xFFFFFFFFu, f & 2u);
  if ((threadIdx.x & 31) == 0 && i < ((n + 31) & ~31ll)) {
    completed[i >> 5] = cw;
    overdue[i >> 5] = ow;
  }
}

// A string member that was null in the document: its ref becomes the null sentinel (json.cuh kRefNull).
__global__ void doc_null_refs_kernel(const uint8_t* flags, int member, int64_t n, uint32_t* ref) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && (flags[i] & (16u << member))) ref[i] = 0xFFFFFFFEu;
}

// rows whose TaskId is not above the previous row's (0 everywhere = the table is in TaskId order)
__global__ void doc_id_order_kernel(const uint4* id, int64_t n, unsigned long long* out_of_order) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < 1 || i >= n) return;
  const uint4 a = id[i - 1], b = id[i];
  const uint32_t x[4] = {a.x, a.y, a.z, a.w}, y[4] = {b.x, b.y, b.z, b.w};
  int cmp = 0;
  for (int k = 0; k < 4 && cmp == 0; ++k) {
    const uint32_t u = __byte_perm(x[k], 0, 0x0123), v = __byte_perm(y[k], 0, 0x0123);  // canonical order is big-endian
    cmp = u < v ? -1 : u > v ? 1 : 0;
  }
  if (cmp >= 0) atomicAdd(out_of_order, 1ull);
}

// Sort keys for the TaskId index: the low (half = 0) or high 8 canonical bytes of rows[i]'s id as a
// big-endian number.  Two stable 64-bit radix sorts, low half first, order the rows by TaskId.
__global__ void id_half_keys_kernel(const uint4* id, const uint32_t* rows, int64_t n, int half, unsigned long long* keys,
                                    uint32_t* vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = rows ? rows[i] : (uint32_t)i;
  const uint4 v = id[r];
  const uint32_t a = half ? v.x : v.z, b = half ? v.y : v.w;
  keys[i] = ((unsigned long long)__byte_perm(a, 0, 0x0123) << 32) | __byte_perm(b, 0, 0x0123);
  vals[i] = r;
}

// positions of a sorted key array whose key equals its left neighbour's (0 = all keys distinct)
__global__ void adjacent_equal_kernel(const unsigned long long* keys, int64_t n, unsigned long long* count) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const bool same = i >= 1 && i < n && keys[i] == keys[i - 1];
  const uint32_t m = __ballot_sync(0xFFFFFFFFu, same);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(count, (unsigned long long)__popc(m));
}

// ---- dictionary encoding of one string member ------------------------------------------------
__global__ void dict_keys_kernel(const uint64_t* hash, int64_t n, unsigned long long* keys, uint32_t* vals) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    keys[i] = hash[i];
    vals[i] = (uint32_t)i;
  }
}

// Byte equality of two strings of the text, eight bytes per step (both words of a step are requested together).
__device__ __forceinline__ bool same_string(const uint8_t* text, uint64_t a, uint32_t la, uint64_t b, uint32_t lb) {
  if (la != lb) return false;
  for (uint32_t j = 0; j < la; j += 8) {
    uint64_t x = load8(text + a + j) ^ load8(text + b + j);
    if (la - j < 8) x &= (1ull << (8 * (la - j))) - 1;
    if (x) return false;
  }
  return true;
}

// position `pos` of the hash-sorted order starts a new dictionary entry unless its string equals
// its left neighbour's (hash first, then length and bytes)
__device__ __forceinline__ uint32_t dict_is_head(const unsigned long long* keys, const uint32_t* docs, const uint8_t* text,
                                                 const uint64_t* start, const uint32_t* len, int64_t pos) {
  if (pos == 0) return 1u;
  if (keys[pos] != keys[pos - 1]) return 1u;
  const uint32_t a = docs[pos - 1], b = docs[pos];
  return same_string(text, start[a], len[a], start[b], len[b]) ? 0u : 1u;
}

// The head test is the expensive part (two scattered strings per position): it is done once here,
// kept as one byte per position, and only read again by dict_assign_kernel.
__global__ void __launch_bounds__(kScanBlock) dict_head_sum_kernel(const unsigned long long* keys, const uint32_t* docs,
                                                                   const uint8_t* text, const uint64_t* start,
                                                                   const uint32_t* len, int64_t n, uint8_t* head,
                                                                   uint64_t* block_sum) {
  __shared__ uint32_t s_warp[32];
  const int64_t pos = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t h = pos < n ? dict_is_head(keys, docs, text, start, len, pos) : 0u;
  if (pos < n) head[pos] = (uint8_t)h;
  uint32_t tot;
  block_exclusive_scan(h, s_warp, &tot);
  if (threadIdx.x == 0) block_sum[blockIdx.x] = tot;
}

// entry number of every position; ref[doc] = entry; the head of an entry records its document
__global__ void __launch_bounds__(kScanBlock) dict_assign_kernel(const uint32_t* docs, const uint8_t* head, int64_t n,
                                                                 const uint64_t* block_base, uint32_t* ref,
                                                                 uint32_t* entry_doc) {
  __shared__ uint32_t s_warp[32];
  const int64_t pos = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t h = pos < n ? head[pos] : 0u;
  const uint32_t before = block_exclusive_scan(h, s_warp, nullptr);
  if (pos >= n) return;
  const uint32_t entry = (uint32_t)block_base[blockIdx.x] + before + h - 1u;
  const uint32_t doc = docs[pos];
  ref[doc] = entry;
  if (h) entry_doc[entry] = doc;
}

__global__ void __launch_bounds__(kScanBlock) dict_len_sum_kernel(const uint32_t* entry_doc, const uint32_t* len, uint32_t k,
                                                                  uint64_t* block_sum) {
  __shared__ uint32_t s_warp[32];
  const uint32_t e = blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t l = e < k ? len[entry_doc[e]] : 0u;
  uint32_t tot;
  block_exclusive_scan(l, s_warp, &tot);
  if (threadIdx.x == 0) block_sum[blockIdx.x] = tot;
}

// off[e] for every entry (off[k] = total is written by the host) and the bytes of the pool
__global__ void __launch_bounds__(kScanBlock) dict_build_kernel(const uint32_t* entry_doc, const uint64_t* start,
                                                                const uint32_t* len, const uint8_t* text, uint32_t k,
                                                                const uint64_t* block_base, uint32_t* off, uint8_t* pool) {
  __shared__ uint32_t s_warp[32];
  const uint32_t e = blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t doc = e < k ? entry_doc[e] : 0u;
  const uint32_t l = e < k ? len[doc] : 0u;
  const uint32_t before = block_exclusive_scan(l, s_warp, nullptr);
  if (e >= k) return;
  const uint32_t o = (uint32_t)block_base[blockIdx.x] + before;
  off[e] = o;
  const uint8_t* src = text + start[doc];
  for (uint32_t j = 0; j < l; ++j) pool[o + j] = src[j];
}

// ---- dictionary encoding through a hash table (the usual path; the sorted one above is the fallback) ----
// An open-addressed table of {hash, first document} with at most half its slots in use.  Every document
// claims or finds the slot of its string's hash and leaves the SMALLEST document number there, so the
// representative of a string, and with it the numbering of the dictionary (entries in order of first
// appearance), does not depend on the order the threads ran in.  Two different strings with one 64-bit
// hash would share a slot: the resolve pass compares every document's bytes with its representative's and
// raises `clash`, and the caller then encodes the member by sorting instead.  `keep` masks the hash (all ones
// in production; a test narrows it to make such clashes happen).
struct DictSlot {
  unsigned long long key;  // all ones = empty
  uint32_t doc;            // all ones = none yet
  uint32_t pad;
};
constexpr unsigned long long kDictEmpty = ~0ull;

__device__ __forceinline__ unsigned long long dict_key(uint64_t h) { return h == kDictEmpty ? h - 1 : h; }
__device__ __forceinline__ uint32_t dict_home(unsigned long long k, int bits) {
  return (uint32_t)((k * 0x9E3779B97F4A7C15ull) >> (64 - bits));
}

__global__ void dict_insert_kernel(const uint64_t* hash, uint64_t keep, int64_t n, DictSlot* tab, int bits) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long k = dict_key(hash[i] & keep);
  const uint32_t mask = (1u << bits) - 1u;
  uint32_t pos = dict_home(k, bits);
  for (;;) {
    // Plain (L1-cached) reads before the atomics: a key never changes once set and a slot's document number only
    // falls, so a stale value costs at most an atomic that was not needed — while 25 M documents that share 50
    // strings would otherwise queue up on 50 addresses in L2.
    unsigned long long cur = tab[pos].key;
    if (cur == kDictEmpty) {
      cur = atomicCAS(&tab[pos].key, kDictEmpty, k);
      if (cur == kDictEmpty) cur = k;
    }
    if (cur == k) {
      if (tab[pos].doc > (uint32_t)i) atomicMin(&tab[pos].doc, (uint32_t)i);
      return;
    }
    pos = (pos + 1) & mask;
  }
}

// rep[doc] = the first document with the same string; a document that is its own representative heads an
// entry.  Per-CTA head counts go to block_sum (scanned by scan_block_sums_kernel).
__global__ void __launch_bounds__(kScanBlock) dict_resolve_kernel(const uint64_t* hash, uint64_t keep, int64_t n, const DictSlot* tab, int bits,
                                                                  const uint8_t* text, const uint64_t* start,
                                                                  const uint32_t* len, uint32_t* rep, uint64_t* block_sum,
                                                                  unsigned long long* clash) {
  __shared__ uint32_t s_warp[32];
  const int64_t i = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  uint32_t h = 0;
  if (i < n) {
    const unsigned long long k = dict_key(hash[i] & keep);
    const uint32_t mask = (1u << bits) - 1u;
    uint32_t pos = dict_home(k, bits);
    while (tab[pos].key != k) pos = (pos + 1) & mask;
    const uint32_t r = tab[pos].doc;
    if (r != (uint32_t)i && !same_string(text, start[r], len[r], start[i], len[i])) atomicAdd(clash, 1ull);
    rep[i] = r;
    h = r == (uint32_t)i;
  }
  uint32_t tot;
  block_exclusive_scan(h, s_warp, &tot);
  if (threadIdx.x == 0) block_sum[blockIdx.x] = tot;
}

// heads take their entry numbers (in document order); ref[] of a head is final after this kernel
__global__ void __launch_bounds__(kScanBlock) dict_number_kernel(const uint32_t* rep, int64_t n, const uint64_t* block_base,
                                                                 uint32_t* ref, uint32_t* entry_doc) {
  __shared__ uint32_t s_warp[32];
  const int64_t i = (int64_t)blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t h = i < n && rep[i] == (uint32_t)i;
  const uint32_t before = block_exclusive_scan(h, s_warp, nullptr);
  if (!h) return;
  const uint32_t entry = (uint32_t)block_base[blockIdx.x] + before;
  ref[i] = entry;
  entry_doc[entry] = (uint32_t)i;
}

// every other document copies its representative's entry number
__global__ void dict_follow_kernel(const uint32_t* rep, int64_t n, uint32_t* ref) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = rep[i];
  if (r != (uint32_t)i) ref[i] = ref[r];
}

}  // namespace tt
