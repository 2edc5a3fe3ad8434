// creator.cuh — K4 creator filter (hash scan + exact string confirm) and K3m mark-by-id.
//
// K4 replaces `Where(t => t.TaskCreatedBy.Equals(createdBy))`
// (Api/Services/FakeTasksManager.cs:70; store-side EQ in Api/Services/TasksStoreManager.cs:56-61):
// the 64-bit creator_hash column is streamed once (8 B/row) for a whole batch of queries; a row
// whose hash equals a query's is a candidate and its dictionary string is compared byte for byte
// with that query (ordinal, case-sensitive — equal UTF-8 bytes <=> equal UTF-16 code units), so
// hash collisions never reach the result.  With more than one query the hashes sit in a small
// open-addressing table in shared memory (1024 to 8192 slots, load factor 1/64 for up to 128
// queries): the common row costs one 32-bit shared-memory probe.
// K3m replaces the per-task upsert loop of Api/Services/TasksStoreManager.cs:143-148: each
// listed TaskId is located by binary search over the TaskId-ordered rows and its overdue bit
// is set with one atomicOr.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tt_synth.h"

namespace tt {

constexpr int kCreatorThreads = 256;
constexpr int kCreatorUnroll = 4;       // 4 x 128-bit loads (8 rows) in flight per thread
constexpr uint32_t kSlotEmpty = 0xFFFFFFFFu;
constexpr int kMaxQueries = 1024;       // per batch
constexpr uint32_t kMaxSlots = 8192;    // 8192 slots x 12 B = 96 KB of shared memory
constexpr uint32_t kListCap = 4096;     // per-query match list a single CTA can order

// A batch of queries as the kernel sees it (all pointers into one device buffer).
struct QuerySet {
  const uint64_t* hash;     // [nq]
  const uint32_t* off;      // [nq + 1] byte offsets into pool
  const uint8_t* pool;      // query strings, UTF-8
  const uint64_t* slot_key; // [slots] open-addressing table: hash -> query index
  const uint32_t* slot_val; // [slots] (21-bit hash tag << 10 | query index) or kSlotEmpty
  uint32_t slots;           // power of two
  uint32_t nq;
};

__host__ __device__ __forceinline__ uint32_t slot_of(uint64_t h, uint32_t slots) {
  return (uint32_t)((h * 0x9E3779B97F4A7C15ULL) >> 40) & (slots - 1);
}
__host__ __device__ __forceinline__ uint32_t tag_of(uint64_t h) { return (uint32_t)(h >> 43); }  // 21 bits

__device__ __forceinline__ ulonglong2 ld_stream_u64x2(const ulonglong2* p) {
  ulonglong2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
  return r;
}

// Where the matches go: per-query lists (ordered afterwards by one CTA per query) or, for result
// sets too large for that, one list of (query, row) pairs.
struct MatchSink {
  uint32_t* lists;            // [nq][kListCap] rows, or null
  uint32_t* qcount;           // [nq] matches per query (may exceed kListCap: then the lists are incomplete)
  unsigned long long* pairs;  // (query << 32 | row), or null
  int64_t pairs_cap;
  unsigned long long* counters;  // [0] exact matches, [1] hash candidates
};

__device__ __forceinline__ void creator_confirm(uint32_t row, uint32_t q, int64_t n, const QuerySet& qs,
                                                const uint32_t* creator_ref, const uint32_t* creator_off,
                                                const uint8_t* creator_pool, const MatchSink& out) {
  if ((int64_t)row >= n) return;  // padding row
  atomicAdd(out.counters + 1, 1ull);
  const uint32_t ref = creator_ref[row];
  if (ref >= 0xFFFFFFFEu) return;  // deleted row (point.cuh), or a task whose TaskCreatedBy is null: no string equals it
  const uint32_t off = creator_off[ref];
  const uint32_t len = creator_off[ref + 1] - off;
  const uint32_t qoff = qs.off[q];
  if (len != qs.off[q + 1] - qoff) return;
  for (uint32_t i = 0; i < len; ++i)
    if (creator_pool[off + i] != qs.pool[qoff + i]) return;
  const unsigned long long pos = atomicAdd(out.counters + 0, 1ull);
  if (out.lists) {
    const uint32_t k = atomicAdd(out.qcount + q, 1u);
    if (k < kListCap) out.lists[(size_t)q * kListCap + k] = row;
  } else if ((int64_t)pos < out.pairs_cap) {
    out.pairs[pos] = ((unsigned long long)q << 32) | row;
  }
}

// Rare path: the first slot probed is occupied.  Walk the chain; every query with this hash
// (there can be several) gets the candidate.
__device__ __noinline__ void creator_probe_chain(unsigned long long h, uint32_t row, uint32_t s, uint32_t v, int64_t n,
                                                 const QuerySet& qs, const uint64_t* s_key, const uint32_t* s_val,
                                                 const uint32_t* creator_ref, const uint32_t* creator_off,
                                                 const uint8_t* creator_pool, const MatchSink& out) {
  const uint32_t tag = tag_of(h);
  while (v != kSlotEmpty) {
    if ((v >> 10) == tag && s_key[s] == h)
      creator_confirm(row, v & 1023u, n, qs, creator_ref, creator_off, creator_pool, out);
    s = (s + 1) & (qs.slots - 1);
    v = s_val[s];
  }
}

template <bool SINGLE>
__global__ void __launch_bounds__(kCreatorThreads, 4) creator_scan_kernel(
    const ulonglong2* __restrict__ hash2, int64_t n, int64_t n_pairs, QuerySet qs,
    const uint32_t* __restrict__ creator_ref, const uint32_t* __restrict__ creator_off,
    const uint8_t* __restrict__ creator_pool, MatchSink out) {
  extern __shared__ __align__(8) unsigned char s_raw[];
  uint64_t* s_key = reinterpret_cast<uint64_t*>(s_raw);
  uint32_t* s_val = reinterpret_cast<uint32_t*>(s_raw + (size_t)qs.slots * 8);
  unsigned long long h0 = 0;
  if (SINGLE) {
    h0 = qs.hash[0];
  } else {
    for (uint32_t i = threadIdx.x; i < qs.slots; i += kCreatorThreads) {
      s_key[i] = qs.slot_key[i];
      s_val[i] = qs.slot_val[i];
    }
    __syncthreads();
  }
  const int64_t stride = (int64_t)gridDim.x * kCreatorThreads;
  for (int64_t base = (int64_t)blockIdx.x * kCreatorThreads + threadIdx.x; base < n_pairs;
       base += stride * kCreatorUnroll) {
    ulonglong2 v[kCreatorUnroll];
#pragma unroll
    for (int j = 0; j < kCreatorUnroll; ++j) {
      int64_t i = base + j * stride;
      // past the end: a hash that cannot be a candidate in either mode is not needed — the row
      // number check in creator_confirm rejects it; reading the last valid vector keeps it simple
      v[j] = ld_stream_u64x2(hash2 + (i < n_pairs ? i : n_pairs - 1));
    }
    if (SINGLE) {
#pragma unroll
      for (int j = 0; j < kCreatorUnroll; ++j) {
        int64_t i = base + j * stride;
        if (i >= n_pairs) continue;
        if (v[j].x == h0) creator_confirm((uint32_t)(2 * i), 0, n, qs, creator_ref, creator_off, creator_pool, out);
        if (v[j].y == h0) creator_confirm((uint32_t)(2 * i + 1), 0, n, qs, creator_ref, creator_off, creator_pool, out);
      }
    } else {
      // All first probes go out together (independent shared-memory loads).  An occupied slot
      // with a different tag and an empty successor is still a miss; only a tag match or a longer
      // chain (both rare at load factor <= 1/16) takes the out-of-line walk.
      uint32_t sl[2 * kCreatorUnroll], sv[2 * kCreatorUnroll];
#pragma unroll
      for (int j = 0; j < kCreatorUnroll; ++j) {
        sl[2 * j] = slot_of(v[j].x, qs.slots);
        sl[2 * j + 1] = slot_of(v[j].y, qs.slots);
        sv[2 * j] = s_val[sl[2 * j]];
        sv[2 * j + 1] = s_val[sl[2 * j + 1]];
      }
#pragma unroll
      for (int k = 0; k < 2 * kCreatorUnroll; ++k) {
        if (sv[k] == kSlotEmpty) continue;
        const int64_t i = base + (k >> 1) * stride;
        if (i >= n_pairs) continue;
        const unsigned long long h = (k & 1) ? v[k >> 1].y : v[k >> 1].x;
        if ((sv[k] >> 10) != tag_of(h) && s_val[(sl[k] + 1) & (qs.slots - 1)] == kSlotEmpty) continue;
        creator_probe_chain(h, (uint32_t)(2 * i + (k & 1)), sl[k], sv[k], n, qs, s_key, s_val, creator_ref, creator_off,
                            creator_pool, out);
      }
    }
  }
}

// ---- K3m ------------------------------------------------------------------------------------

// TaskIds are stored as their 16 canonical bytes; compare them as big-endian numbers.
__device__ __forceinline__ int cmp_id(uint4 a, uint4 b) {
  uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    uint32_t x = __byte_perm(aw[k], 0, 0x0123), y = __byte_perm(bw[k], 0, 0x0123);
    if (x != y) return x < y ? -1 : 1;
  }
  return 0;
}

// counters[0] += ids found (bit set), counters[1] += ids with no live row.  Rows (or id_order)
// ascend by TaskId; a deleted row keeps its id but carries creator_ref == 0xFFFFFFFF and is
// passed over, like FirstOrDefault over the live list (FakeTasksManager.cs:76) / a missing key.
__global__ void mark_by_id_kernel(const uint4* ids, int64_t n_ids, const uint4* task_id, const uint32_t* creator_ref,
                                  const uint32_t* id_order, int64_t n_rows, uint32_t* overdue,
                                  unsigned long long* counters) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_ids) return;
  const uint4 key = ids[i];
  int64_t lo = 0, hi = n_rows;  // first position whose id >= key
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    uint32_t r = id_order ? id_order[mid] : (uint32_t)mid;
    if (cmp_id(task_id[r], key) < 0) lo = mid + 1; else hi = mid;
  }
  bool hit = false;
  for (; lo < n_rows; ++lo) {
    uint32_t r = id_order ? id_order[lo] : (uint32_t)lo;
    if (cmp_id(task_id[r], key) != 0) break;
    if (creator_ref[r] != 0xFFFFFFFFu) {
      atomicOr(overdue + (r >> 5), 1u << (r & 31));
      hit = true;
      break;
    }
  }
  atomicAdd(counters + (hit ? 0 : 1), 1ull);
}

}  // namespace tt
