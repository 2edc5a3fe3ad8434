// json.cuh — K6: a result list written as the JSON the reference's controllers return.
//
// The step right after the sweep / the creator filter: the reference's
// OverdueTasksController.Get (Api/Controllers/OverdueTasksController.cs:22-26) and
// TasksController.Get (Api/Controllers/TasksController.cs:21-25) return List<TaskModel>, which
// ASP.NET Core serialises with System.Text.Json: camelCase property names in declaration order
// (Api/Models/TaskModel.cs:5-12), no whitespace, Guid in "D" form, DateTime in ISO 8601 extended
// form (sample: docs/aca/05-aca-dapr-pubsubapi/index.md:60-69).  Two encoders occur in the
// reference:
//   kJsonResponse  the MVC output formatter (JavaScriptEncoder.UnsafeRelaxedJsonEscaping)
//   kJsonDocument  Dapr's SaveStateAsync / PublishEventAsync / InvokeMethodAsync bodies
//                  (Api/Services/TasksStoreManager.cs:35,155; JavaScriptEncoder.Default)
//
// Strings are dictionary-encoded on the device, so escaping is done ONCE per dictionary entry
// (esc_len / esc_write kernels, cached in the table) and the per-hit work is pure copying:
// K6a json_len_kernel sizes every record (dictionary lengths + date lengths), a two-level scan
// turns the sizes into byte offsets, and K6b json_emit_kernel lets each thread assemble one
// record in shared memory a 32-bit word at a time (JsonPacker: literals as compile-time words, TaskId
// hex and date digits four characters at a time, pre-escaped strings as words); the CTA then streams
// its contiguous piece of the text to HBM with 16-byte stores.  All byte work, HBM-bound: ≈270 B written per hit, ≈7 scattered reads.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tt_bmp_defined.h"
#include "sweep.cuh"

namespace tt {

constexpr int kJsonResponse = 0;
constexpr int kJsonDocument = 1;

// ---- L2 residency hints -----------------------------------------------------------------------
// The escaped dictionaries (a few tens of MB) are re-read by every record while gigabytes of
// output stream through the same 126 MB L2: dictionary loads ask to be evicted last, the output
// stores and the one-shot loads to be evicted first.
__device__ __forceinline__ uint64_t l2_evict_last_policy() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_evict_first_policy() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint32_t ld_hint_u32(const uint32_t* p, uint64_t pol) {
  uint32_t v;
  asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ uint4 ld_hint_v4(const uint4* p, uint64_t pol) {
  uint4 v;
  asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ uint64_t ld_hint_u64(const uint64_t* p, uint64_t pol) {
  uint64_t v;
  asm volatile("ld.global.nc.L2::cache_hint.u64 %0, [%1], %2;" : "=l"(v) : "l"(p), "l"(pol));
  return v;
}
__device__ __forceinline__ void st_hint_v4(uint4* p, uint4 v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1, %2, %3, %4}, %5;"
               :
               : "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void st_hint_u64(uint64_t* p, uint64_t v, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.u64 [%0], %1, %2;" : : "l"(p), "l"(v), "l"(pol) : "memory");
}

// BMP code points System.Text.Encodings.Web treats as defined (ForbidUndefinedCharacters): generated from an
// independent copy of the Unicode Character Database, include/tt_bmp_defined.h.
TT_BMP_TABLE kBmpDefined[2048] = {TT_BMP_DEFINED_WORDS};

// ---- escaping one Unicode scalar -----------------------------------------------------------
// Returns how a scalar is written: 0 = as is, 1 = backslash + *short_ch, 2 = \uXXXX,
// 3 = \uXXXX\uXXXX (surrogate pair).  System.Text.Json's published rules (assumptions of record in
// DESIGN.md §7): both encoders escape controls, '"' and '\\' and every supplementary-plane
// scalar; Default additionally escapes the HTML-sensitive ASCII characters and everything
// outside Basic Latin; UnsafeRelaxed keeps a BMP code point only if the encoder's table calls it
// defined — a letter, mark, number, punctuation or symbol, or U+0020: controls, format characters,
// the other separators, surrogates, private use and unassigned code points are escaped.
__device__ __forceinline__ int json_escape_kind(uint32_t cp, int flavor, uint8_t* short_ch) {
  if (cp < 0x20) {
    switch (cp) {
      case 8: *short_ch = 'b'; return 1;
      case 9: *short_ch = 't'; return 1;
      case 10: *short_ch = 'n'; return 1;
      case 12: *short_ch = 'f'; return 1;
      case 13: *short_ch = 'r'; return 1;
      default: return 2;
    }
  }
  if (cp == '\\') { *short_ch = '\\'; return 1; }
  if (cp == '"') {
    if (flavor == kJsonResponse) { *short_ch = '"'; return 1; }
    return 2;
  }
  if (cp < 0x7F) {
    if (flavor == kJsonDocument && (cp == '&' || cp == '\'' || cp == '+' || cp == '<' || cp == '>' || cp == '`')) return 2;
    return 0;
  }
  if (cp >= 0x10000) return 3;
  if (flavor == kJsonDocument) return 2;
  return (kBmpDefined[cp >> 5] >> (cp & 31)) & 1u ? 0 : 2;
}

__device__ __forceinline__ uint8_t hex_upper(uint32_t x) { return (uint8_t)(x < 10 ? '0' + x : 'A' + x - 10); }
__device__ __forceinline__ void put_u4(uint8_t* o, uint32_t unit) {
  o[0] = '\\';
  o[1] = 'u';
  o[2] = hex_upper((unit >> 12) & 15);
  o[3] = hex_upper((unit >> 8) & 15);
  o[4] = hex_upper((unit >> 4) & 15);
  o[5] = hex_upper(unit & 15);
}

// Escapes s[0, len) (UTF-8).  WRITE = false only measures.  Returns the escaped length, or ~0u
// if the bytes are not well-formed UTF-8 (a lead byte without its continuation bytes).
template <bool WRITE>
__device__ uint32_t json_escape(const uint8_t* s, uint32_t len, int flavor, uint8_t* out) {
  uint32_t i = 0, o = 0;
  while (i < len) {
    const uint32_t c = s[i];
    uint32_t cp = c, extra = 0;
    if (c >= 0x80) {
      if ((c >> 5) == 6) { cp = c & 0x1F; extra = 1; }
      else if ((c >> 4) == 14) { cp = c & 0x0F; extra = 2; }
      else if ((c >> 3) == 30) { cp = c & 0x07; extra = 3; }
      else return ~0u;
      if (i + extra >= len) return ~0u;
      for (uint32_t j = 1; j <= extra; ++j) {
        const uint32_t d = s[i + j];
        if ((d >> 6) != 2) return ~0u;
        cp = (cp << 6) | (d & 0x3F);
      }
    }
    uint8_t sc = 0;
    const int kind = json_escape_kind(cp, flavor, &sc);
    if (kind == 0) {
      if (WRITE)
        for (uint32_t j = 0; j <= extra; ++j) out[o + j] = s[i + j];
      o += extra + 1;
    } else if (kind == 1) {
      if (WRITE) { out[o] = '\\'; out[o + 1] = sc; }
      o += 2;
    } else if (kind == 2) {
      if (WRITE) put_u4(out + o, cp);
      o += 6;
    } else {
      if (WRITE) {
        const uint32_t v = cp - 0x10000;
        put_u4(out + o, 0xD800 + (v >> 10));
        put_u4(out + o + 6, 0xDC00 + (v & 0x3FF));
      }
      o += 12;
    }
    i += extra + 1;
  }
  return o;
}

// One dictionary entry's escaped form: where it starts in the escaped pool (in 32-bit words; every
// entry starts on a 16-byte boundary so the emitter reads it with aligned 128-bit loads) and its
// length in bytes.
struct EscEntry {
  uint32_t word_off;
  uint32_t len;
};
__host__ __device__ __forceinline__ uint32_t esc_words(uint32_t len) { return ((len + 15u) >> 4) * 4u; }

constexpr uint32_t kEscMaxLen = 1u << 20;  // longest escaped string the emitter accepts

// Pass 1 over a dictionary: escaped length of every entry.  err[0] |= 1 on malformed UTF-8 or an
// over-long entry, err[1] = one offending entry.
__global__ void esc_len_kernel(const uint32_t* off, const uint8_t* pool, uint32_t n, int flavor, uint32_t* len_out,
                               uint32_t* err) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint32_t a = off[k], b = off[k + 1];
  uint32_t l = json_escape<false>(pool + a, b - a, flavor, nullptr);
  if (l == ~0u || l > kEscMaxLen) {
    atomicOr(err, 1u);
    err[1] = k;
    l = 0;
  }
  len_out[k] = l;
}

// Pass 2 (after the scan of the word counts): writes the escaped bytes.
__global__ void esc_write_kernel(const uint32_t* off, const uint8_t* pool, uint32_t n, int flavor,
                                 const EscEntry* entry, uint32_t* esc_pool) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const uint32_t a = off[k], b = off[k + 1];
  uint8_t* dst = reinterpret_cast<uint8_t*>(esc_pool + entry[k].word_off);
  const uint32_t l = entry[k].len;
  if (l) json_escape<true>(pool + a, b - a, flavor, dst);
  for (uint32_t j = l; j < ((l + 15) & ~15u); ++j) dst[j] = 0;  // padding of the last 16 bytes
}

// ---- a small generic scan: u32 weights -> u64 exclusive offsets ------------------------------
// Level 1: one CTA per 1024 elements writes their sum; level 2: ONE CTA walks the block sums
// (n / 1024 of them) and leaves their exclusive scan in place + the grand total; the consumer
// kernels redo the cheap in-block scan themselves.
constexpr int kScanBlock = 1024;

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t* total) {
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
    if (lane >= (uint32_t)d) incl += up;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t base = 0, tot = 0;
  const uint32_t nw = (blockDim.x + 31) >> 5;
  for (uint32_t w = 0; w < nw; ++w) {
    const uint32_t x = s_warp[w];
    base += w < warp ? x : 0u;
    tot += x;
  }
  __syncthreads();  // s_warp may be reused by the caller's next scan
  if (total) *total = tot;
  return base + incl - v;
}

// weights of a dictionary: words of each escaped entry
__global__ void __launch_bounds__(kScanBlock) esc_block_sum_kernel(const uint32_t* len, uint32_t n, uint64_t* block_sum) {
  __shared__ uint32_t s_warp[32];
  const uint32_t k = blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t w = k < n ? esc_words(len[k]) : 0u;
  uint32_t tot;
  block_exclusive_scan(w, s_warp, &tot);
  if (threadIdx.x == 0) block_sum[blockIdx.x] = tot;
}

// in place: block_sum[b] <- sum of the blocks before b; block_sum[nb] <- total.  One CTA; every
// thread owns kScanPerThread consecutive sums per round.
constexpr int kScanPerThread = 8;
__global__ void __launch_bounds__(1024) scan_block_sums_kernel(uint64_t* block_sum, uint32_t nb) {
  __shared__ uint64_t s_warp[32];
  __shared__ uint64_t s_carry;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  for (uint32_t b0 = 0; b0 < nb; b0 += 1024 * kScanPerThread) {
    const uint32_t b = b0 + threadIdx.x * kScanPerThread;
    uint64_t v[kScanPerThread];
    uint64_t mine = 0;
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
      v[k] = b + k < nb ? block_sum[b + k] : 0ull;
      mine += v[k];
    }
    uint64_t incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint64_t up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= (uint32_t)d) incl += up;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint64_t run = s_carry + incl - mine;
    for (uint32_t w = 0; w < warp; ++w) run += s_warp[w];
#pragma unroll
    for (int k = 0; k < kScanPerThread; ++k) {
      if (b + k < nb) block_sum[b + k] = run;
      run += v[k];
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = run;
    __syncthreads();
  }
  if (threadIdx.x == 0) block_sum[nb] = s_carry;
}

__global__ void __launch_bounds__(kScanBlock) esc_entries_kernel(const uint32_t* len, uint32_t n, const uint64_t* block_base,
                                                                 EscEntry* entry) {
  __shared__ uint32_t s_warp[32];
  const uint32_t k = blockIdx.x * kScanBlock + threadIdx.x;
  const uint32_t l = k < n ? len[k] : 0u;
  const uint32_t local = block_exclusive_scan(esc_words(l), s_warp, nullptr);
  if (k < n) {
    EscEntry e;
    e.word_off = (uint32_t)(block_base[blockIdx.x] + local);
    e.len = l;
    entry[k] = e;
  }
}

// ---- the record ---------------------------------------------------------------------------
// {"taskId":"<36>","taskName":"..","taskCreatedBy":"..","taskCreatedOn":"..","taskDueDate":"..",
//  "taskAssignedTo":"..","isCompleted":false,"isOverDue":false}
// preceded by '[' (first record) or ',' and, for the last record, followed by ']'.
constexpr uint32_t kJsonFixed = 1 + 11 + 36 + 14 + 19 + 19 + 17 + 20 + 16 + 13 + 1;  // lead + literals + id + '}'

struct JsonParams {
  const uint32_t* rows;      // the result's row numbers
  const uint4* ids;          // the result's TaskIds (same order)
  int64_t n;
  const uint64_t* created_raw;
  const uint64_t* due_raw;
  const uint32_t* completed;
  const uint32_t* overdue;
  const uint32_t* ref[3];    // name, creator, assignee: row -> dictionary entry (null = empty strings)
  uint32_t dict_n[3];        // entries in each dictionary (a ref beyond it writes an empty string)
  const EscEntry* entry[3];
  const uint32_t* esc_pool[3];
  int32_t local_offset_min;  // UTC offset used to print Kind = Local values
  uint64_t* block_base;      // [blocks + 1]
  uint64_t* plan;            // [5][plan_stride]: what K6a gathered for every record, read back by K6b
  int64_t plan_stride;
  uint64_t* rec_off;         // [n]: where every record's piece starts in the text (may be null)
  uint8_t* out;
};

__device__ __forceinline__ uint32_t datetime_len(uint64_t raw) {
  const uint64_t ticks = raw & 0x3FFFFFFFFFFFFFFFull;
  const uint32_t kind = (uint32_t)(raw >> 62);
  uint32_t frac = (uint32_t)(ticks % 10000000ull);
  uint32_t l = 19;
  if (frac) {
    l += 8;
    while (frac % 10 == 0) { frac /= 10; --l; }
  }
  return l + (kind == 1 ? 1u : kind >= 2 ? 6u : 0u);
}

constexpr int kJsonThreads = 128;            // records per CTA (the default; the kernels are built for 64 and 256 as well)
constexpr int kJsonStagePerThread = 352;     // shared-memory staging per record: 44 KB per 128-thread CTA, 5 CTAs per SM
constexpr int json_stage_bytes(int threads) { return threads * kJsonStagePerThread; }

// A string member that holds null (a C# string may: `"taskName":null` in a stored document) is written
// as the literal null, as JsonSerializer does.  Its ref in the row's string column is kRefNull; in a
// record's plan it travels as an entry with this word offset and length 2 (= "null" minus the two quotes
// the surrounding literals carry).
constexpr uint32_t kRefNull = 0xFFFFFFFEu;
constexpr uint32_t kNullWordOff = 0xFFFFFFFFu;

struct JsonRecord {
  uint64_t created, due;
  EscEntry s[3];
  bool completed, overdue, live;
};

// K6a's gathers: five row-indexed columns, the two flag planes and three dictionary entries.
__device__ __forceinline__ JsonRecord json_gather(const JsonParams& p, int64_t i, uint64_t keep) {
  JsonRecord r;
  r.live = i < p.n;
  r.created = r.due = 0;
  r.completed = r.overdue = false;
#pragma unroll
  for (int c = 0; c < 3; ++c) r.s[c].word_off = r.s[c].len = 0;
  if (!r.live) return r;
  const uint32_t row = p.rows[i];
  r.created = ld_gather_u64(p.created_raw + row);
  r.due = ld_gather_u64(p.due_raw + row);
  uint32_t k[3] = {~0u, ~0u, ~0u};
#pragma unroll
  for (int c = 0; c < 3; ++c)
    if (p.ref[c]) k[c] = p.ref[c][row];
  r.completed = (p.completed[row >> 5] >> (row & 31)) & 1u;
  r.overdue = (p.overdue[row >> 5] >> (row & 31)) & 1u;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    if (k[c] < p.dict_n[c]) {
      const uint64_t e = ld_hint_u64(reinterpret_cast<const uint64_t*>(p.entry[c] + k[c]), keep);
      r.s[c].word_off = (uint32_t)e;
      r.s[c].len = (uint32_t)(e >> 32);
    } else if (k[c] == kRefNull) {
      r.s[c].word_off = kNullWordOff;
      r.s[c].len = 2;
    }
  }
  return r;
}

__device__ __forceinline__ uint32_t json_record_len(const JsonParams& p, const JsonRecord& r, int64_t i) {
  if (!r.live) return 0;
  return kJsonFixed + r.s[0].len + r.s[1].len + r.s[2].len + datetime_len(r.created) + datetime_len(r.due) +
         (r.completed ? 4u : 5u) + (r.overdue ? 4u : 5u) + (i == p.n - 1 ? 1u : 0u);
}

// The plan: 5 x u64 per record, one array per word (coalesced both ways).  Entry lengths are
// < 2^21 (kEscMaxLen), so the two flags ride in bit 63 of words 2 and 3.
__device__ __forceinline__ void json_plan_store(const JsonParams& p, int64_t i, const JsonRecord& r, uint64_t stream) {
  uint64_t* w = p.plan + i;
  st_hint_u64(w, r.created, stream);
  st_hint_u64(w + p.plan_stride, r.due, stream);
  st_hint_u64(w + 2 * p.plan_stride, r.s[0].word_off | ((uint64_t)r.s[0].len << 32) | ((uint64_t)r.completed << 63), stream);
  st_hint_u64(w + 3 * p.plan_stride, r.s[1].word_off | ((uint64_t)r.s[1].len << 32) | ((uint64_t)r.overdue << 63), stream);
  st_hint_u64(w + 4 * p.plan_stride, r.s[2].word_off | ((uint64_t)r.s[2].len << 32), stream);
}

__device__ __forceinline__ JsonRecord json_plan_load(const JsonParams& p, int64_t i, uint64_t stream) {
  JsonRecord r;
  r.live = i < p.n;
  r.created = r.due = 0;
  r.completed = r.overdue = false;
#pragma unroll
  for (int c = 0; c < 3; ++c) r.s[c].word_off = r.s[c].len = 0;
  if (!r.live) return r;
  const uint64_t* w = p.plan + i;
  r.created = ld_hint_u64(w, stream);
  r.due = ld_hint_u64(w + p.plan_stride, stream);
  const uint64_t a = ld_hint_u64(w + 2 * p.plan_stride, stream), b = ld_hint_u64(w + 3 * p.plan_stride, stream),
                 c = ld_hint_u64(w + 4 * p.plan_stride, stream);
  r.s[0].word_off = (uint32_t)a;
  r.s[0].len = (uint32_t)(a >> 32) & 0x7FFFFFFFu;
  r.completed = a >> 63;
  r.s[1].word_off = (uint32_t)b;
  r.s[1].len = (uint32_t)(b >> 32) & 0x7FFFFFFFu;
  r.overdue = b >> 63;
  r.s[2].word_off = (uint32_t)c;
  r.s[2].len = (uint32_t)(c >> 32);
  return r;
}

// K6a: gathers every record's fields once (the scattered reads), keeps them as the plan, and
// sizes every CTA's piece of the text
template <int THREADS>
__global__ void __launch_bounds__(THREADS) json_len_kernel(JsonParams p) {
  __shared__ uint32_t s_warp[32];
  const uint64_t keep = l2_evict_last_policy(), stream = l2_evict_first_policy();
  const int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x;
  const JsonRecord r = json_gather(p, i, keep);
  if (r.live) json_plan_store(p, i, r, stream);
  uint32_t tot;
  block_exclusive_scan(json_record_len(p, r, i), s_warp, &tot);
  if (threadIdx.x == 0) p.block_base[blockIdx.x] = tot;
}

// ---- word-wise assembly of a record -----------------------------------------------------------
// A record is ~290 bytes at an arbitrary byte offset.  Written a byte at a time it costs a store (and a constant
// load) per byte, all chained through the output pointer; JsonPacker assembles 32-bit words instead: `prev` holds the
// last source word, whose top s / 8 bytes have not been written yet, and appending a word v is one funnel shift and one
// aligned store — (v << s) | (prev >> (32 - s)).  Literals are compile-time words, the TaskId's hex digits and the
// dates' digits are computed four characters at a time, escaped strings arrive as words from the dictionary pool.
// Only a record's first and last word, which it shares with its neighbours, are written bytewise.
struct JsonPacker {
  uint32_t* wp;   // next aligned word of the output
  uint32_t prev;  // its top s / 8 bytes are pending
  uint32_t s;     // 8 * pending bytes, 0..24

  // starts at byte address o with the record's first four bytes: only the bytes from o on are stored
  __device__ __forceinline__ void start(uint8_t* o, uint32_t first) {
    const uint32_t f = (uint32_t)(reinterpret_cast<uintptr_t>(o) & 3u);
    wp = reinterpret_cast<uint32_t*>(o - f);
    s = 8u * f;
    const uint32_t out = first << s;
    for (uint32_t b = f; b < 4u; ++b) reinterpret_cast<uint8_t*>(wp)[b] = (uint8_t)(out >> (8u * b));
    ++wp;
    prev = first;
  }
  __device__ __forceinline__ void word(uint32_t v) {
    *wp++ = __funnelshift_l(prev, v, s);
    prev = v;
  }
  // k = 1..3 bytes in the low end of v, zero above
  __device__ __forceinline__ void part(uint32_t v, uint32_t k) {
    const uint32_t out = __funnelshift_l(prev, v, s);
    prev = __funnelshift_r(prev, v, 8u * k);
    s += 8u * k;
    if (s >= 32u) {
      *wp++ = out;
      s -= 32u;
    }
  }
  // the pending bytes; returns the address after the record's last byte
  __device__ __forceinline__ uint8_t* finish() {
    uint8_t* o = reinterpret_cast<uint8_t*>(wp);
    const uint32_t n = s >> 3;
    const uint32_t tail = n ? prev >> (32u - s) : 0u;
    for (uint32_t b = 0; b < n; ++b) o[b] = (uint8_t)(tail >> (8u * b));
    return o + n;
  }
};

// bytes [from, from + 4) of a literal as a little-endian word (zero beyond its end)
template <int N>
__host__ __device__ constexpr uint32_t lit_word(const char (&lit)[N], int from) {
  uint32_t w = 0;
  for (int k = 0; k < 4; ++k)
    if (from + k < N - 1) w |= (uint32_t)(uint8_t)lit[from + k] << (8 * k);
  return w;
}

// a literal as words, built at compile time
template <int N>
struct LitWords {
  uint32_t w[(N + 2) / 4 + 1];
};
template <int N>
__host__ __device__ constexpr LitWords<N> lit_words(const char (&lit)[N]) {
  LitWords<N> r = {};
  for (int j = 0; j * 4 < N - 1; ++j) r.w[j] = lit_word(lit, 4 * j);
  return r;
}

// appends a literal: its whole words, then the 1..3 bytes that are left (lengths known at compile time)
template <int N>
__device__ __forceinline__ void pack_words(JsonPacker& k, const LitWords<N> lw) {
  constexpr int len = N - 1;
#pragma unroll
  for (int j = 0; j < len / 4; ++j) k.word(lw.w[j]);
  if constexpr (len % 4 != 0) k.part(lw.w[len / 4], (uint32_t)(len % 4));
}
#define TT_PACK_LIT(K, LIT)                    \
  do {                                         \
    constexpr auto lit_w_ = lit_words(LIT);    \
    pack_words(K, lit_w_);                     \
  } while (0)

// four lower-case hex digits of two bytes (the byte in bits 0..7 first), as four characters
__device__ __forceinline__ uint32_t hex4(uint32_t v) {
  const uint32_t t = ((v & 0xF0u) >> 4) | ((v & 0x0Fu) << 8) | ((v & 0xF000u) << 4) | ((v & 0x0F00u) << 16);
  return t + 0x30303030u + (((t + 0x06060606u) >> 4) & 0x01010101u) * 39u;
}

// two decimal digits of v < 100 as two characters (tens first)
__device__ __forceinline__ uint32_t dec2(uint32_t v) {
  const uint32_t t = v / 10u;
  return (0x3030u + t) | ((v - 10u * t) << 8);
}

// "yyyy-MM-ddTHH:mm:ss[.fffffff][Z|+hh:mm]" (System.Text.Json's DateTime writer)
__device__ __forceinline__ void pack_datetime(JsonPacker& k, uint64_t raw, int32_t local_offset_min) {
  const uint64_t ticks = raw & 0x3FFFFFFFFFFFFFFFull;
  const uint32_t kind = (uint32_t)(raw >> 62);
  const uint64_t days = ticks / 864000000000ull;
  const uint64_t tod = ticks % 864000000000ull;
  // civil date from days since 0001-01-01 (proleptic Gregorian, era arithmetic)
  const uint32_t z = (uint32_t)days + 306;  // days since 0000-03-01
  const uint32_t era = z / 146097;
  const uint32_t doe = z - era * 146097;
  const uint32_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  uint32_t y = yoe + era * 400;
  const uint32_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const uint32_t mp = (5 * doy + 2) / 153;
  const uint32_t d = doy - (153 * mp + 2) / 5 + 1;
  const uint32_t m = mp < 10 ? mp + 3 : mp - 9;
  if (m <= 2) y += 1;
  const uint32_t secs = (uint32_t)(tod / 10000000ull);
  uint32_t frac = (uint32_t)(tod % 10000000ull);
  const uint32_t hh = secs / 3600, mi = secs / 60 % 60, ss = secs % 60;
  const uint32_t dh = dec2(hh);
  k.word(dec2(y / 100) | (dec2(y % 100) << 16));              // yyyy
  k.word('-' | (dec2(m) << 8) | ((uint32_t)'-' << 24));       // -MM-
  k.word(dec2(d) | ((uint32_t)'T' << 16) | (dh << 24));       // ddTH
  k.word((dh >> 8) | ((uint32_t)':' << 8) | (dec2(mi) << 16));  // H:mm
  k.part(':' | (dec2(ss) << 8), 3);                           // :ss
  if (frac) {
    uint32_t digits = 7;
    while (frac % 10 == 0) { frac /= 10; --digits; }
    // '.' and the digits, most significant first: eight characters at most
    uint64_t txt = '.';
    uint32_t pow = 1;
    for (uint32_t j = 1; j < digits; ++j) pow *= 10;
    for (uint32_t j = 0; j < digits; ++j) {
      const uint32_t q = frac / pow;
      txt |= (uint64_t)('0' + q) << (8 * (j + 1));
      frac -= q * pow;
      pow /= 10;
    }
    const uint32_t n = digits + 1;
    if (n >= 4) {
      k.word((uint32_t)txt);
      if (n == 8) k.word((uint32_t)(txt >> 32));
      else if (n > 4) k.part((uint32_t)(txt >> 32), n - 4);
    } else {
      k.part((uint32_t)txt, n);
    }
  }
  if (kind == 1) {
    k.part('Z', 1);
  } else if (kind >= 2) {
    int32_t off = local_offset_min;
    const uint32_t sign = off < 0 ? '-' : '+';
    if (off < 0) off = -off;
    const uint32_t oh = dec2((uint32_t)off / 60), om = dec2((uint32_t)off % 60);
    k.word(sign | (oh << 8) | ((uint32_t)':' << 24));  // +hh:
    k.part(om, 2);                                     // mm
  }
}

// The first 32 bytes of an escaped string, fetched before any formatting starts: the three strings
// of a record are six independent 128-bit loads in flight together instead of a chain of
// dependent word loads (each an L2 round trip) in the middle of the stores.
struct EscHead {
  uint4 q0, q1;
};

__device__ __forceinline__ EscHead esc_prefetch(const uint32_t* pool, EscEntry e, uint64_t keep) {
  EscHead h;
  h.q0 = h.q1 = make_uint4(0, 0, 0, 0);
  const uint4* q = reinterpret_cast<const uint4*>(pool + e.word_off);
  if (e.word_off == kNullWordOff) return h;
  if (e.len > 0) h.q0 = ld_hint_v4(q, keep);
  if (e.len > 16) h.q1 = ld_hint_v4(q + 1, keep);
  return h;
}

// an escaped string: whole words, then its last 1..3 bytes (the pool pads every entry with zeros)
__device__ __forceinline__ void pack_esc(JsonPacker& k, const uint32_t* pool, EscEntry e, const EscHead& h, uint64_t keep) {
  const uint32_t w[8] = {h.q0.x, h.q0.y, h.q0.z, h.q0.w, h.q1.x, h.q1.y, h.q1.z, h.q1.w};
  const uint32_t nw = e.len >> 2, rest = e.len & 3u;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if ((uint32_t)j < nw) k.word(w[j]);
    else if ((uint32_t)j == nw && rest) k.part(w[j], rest);
  }
  if (nw >= 8) {  // long strings: the words after the prefetched 32 bytes
    const uint32_t* src = pool + e.word_off;
    for (uint32_t j = 8; j < nw; ++j) k.word(ld_hint_u32(src + j, keep));
    if (rest) k.part(ld_hint_u32(src + nw, keep), rest);
  }
}

// One string member's value: "…" or, for a member that holds null, the literal null.
__device__ __forceinline__ void pack_member(JsonPacker& k, const uint32_t* pool, EscEntry e, const EscHead& h, uint64_t keep) {
  if (e.word_off == kNullWordOff) {
    k.word(lit_word("null", 0));
    return;
  }
  k.part('"', 1);
  pack_esc(k, pool, e, h, keep);
  k.part('"', 1);
}

// {"taskId":"<36>","taskName":…,"taskCreatedBy":…,"taskCreatedOn":"…","taskDueDate":"…","taskAssignedTo":…,
//  "isCompleted":…,"isOverDue":…}  preceded by '[' or ',' and, for the last record, followed by ']'
__device__ uint8_t* json_write_record(const JsonParams& p, const JsonRecord& r, uint4 id, int64_t i, uint8_t* o,
                                      const EscHead& h0, const EscHead& h1, const EscHead& h2, uint64_t keep) {
  JsonPacker k;
  k.start(o, (i == 0 ? (uint32_t)'[' : (uint32_t)',') | (lit_word("{\"t", 0) << 8));  // [{"t  or  ,{"t
  TT_PACK_LIT(k, "askId\":\"");
  // canonical bytes 0..15 in memory order: 8-4-4-4-12 hex digits
  const uint32_t a = hex4(id.y & 0xFFFFu), b = hex4(id.y >> 16), c = hex4(id.z & 0xFFFFu);
  k.word(hex4(id.x & 0xFFFFu));
  k.word(hex4(id.x >> 16));
  k.word('-' | (a << 8));
  k.word((a >> 24) | ((uint32_t)'-' << 8) | (b << 16));
  k.word((b >> 16) | ((uint32_t)'-' << 16) | (c << 24));
  k.word((c >> 8) | ((uint32_t)'-' << 24));
  k.word(hex4(id.z >> 16));
  k.word(hex4(id.w & 0xFFFFu));
  k.word(hex4(id.w >> 16));
  TT_PACK_LIT(k, "\",\"taskName\":");
  pack_member(k, p.esc_pool[0], r.s[0], h0, keep);
  TT_PACK_LIT(k, ",\"taskCreatedBy\":");
  pack_member(k, p.esc_pool[1], r.s[1], h1, keep);
  TT_PACK_LIT(k, ",\"taskCreatedOn\":\"");
  pack_datetime(k, r.created, p.local_offset_min);
  TT_PACK_LIT(k, "\",\"taskDueDate\":\"");
  pack_datetime(k, r.due, p.local_offset_min);
  TT_PACK_LIT(k, "\",\"taskAssignedTo\":");
  pack_member(k, p.esc_pool[2], r.s[2], h2, keep);
  TT_PACK_LIT(k, ",\"isCompleted\":");
  if (r.completed) TT_PACK_LIT(k, "true");
  else TT_PACK_LIT(k, "false");
  TT_PACK_LIT(k, ",\"isOverDue\":");
  if (r.overdue) TT_PACK_LIT(k, "true");
  else TT_PACK_LIT(k, "false");
  if (i == p.n - 1) TT_PACK_LIT(k, "}]");
  else k.part('}', 1);
  return k.finish();
}

// K6b: every thread formats one record; the CTA's piece of the text is contiguous in the output,
// so it is staged in shared memory (shifted so that 16-byte chunks of both line up) and flushed
// with 16-byte stores.  A CTA whose piece does not fit the staging area (very long strings)
// writes straight to global memory instead.
template <int THREADS>
__global__ void __launch_bounds__(THREADS) json_emit_kernel(JsonParams p) {
  extern __shared__ __align__(16) uint8_t s_stage[];
  __shared__ uint32_t s_warp[32];
  const uint64_t keep = l2_evict_last_policy(), stream = l2_evict_first_policy();
  const int64_t i = (int64_t)blockIdx.x * THREADS + threadIdx.x;
  const JsonRecord r = json_plan_load(p, i, stream);
  uint4 id = make_uint4(0, 0, 0, 0);
  if (r.live) id = p.ids[i];
  // the strings' first 32 bytes are requested now and land while the CTA does its scan
  const EscHead h0 = esc_prefetch(p.esc_pool[0], r.s[0], keep), h1 = esc_prefetch(p.esc_pool[1], r.s[1], keep),
                h2 = esc_prefetch(p.esc_pool[2], r.s[2], keep);
  uint32_t total;
  const uint32_t local = block_exclusive_scan(json_record_len(p, r, i), s_warp, &total);
  const uint64_t base = p.block_base[blockIdx.x];
  const uint32_t shift = (uint32_t)(base & 15u);
  const bool staged = shift + total <= (uint32_t)json_stage_bytes(THREADS);
  if (r.live && p.rec_off) st_hint_u64(p.rec_off + i, base + local, stream);
  if (r.live) json_write_record(p, r, id, i, staged ? s_stage + shift + local : p.out + base + local, h0, h1, h2, keep);
  if (!staged) return;
  __syncthreads();
  uint8_t* dst = p.out + (base - shift);
  const uint32_t lo = shift, hi = shift + total;
  for (uint32_t a = threadIdx.x * 16; a < hi; a += THREADS * 16) {
    if (a >= lo && a + 16 <= hi) {
      st_hint_v4(reinterpret_cast<uint4*>(dst + a), *reinterpret_cast<const uint4*>(s_stage + a), stream);
    } else {
      for (uint32_t e = a > lo ? a : lo; e < a + 16 && e < hi; ++e) dst[e] = s_stage[e];
    }
  }
}

}  // namespace tt
