// order.cuh — K5: order a hit list by TaskCreatedOn and gather the TaskIds in that order.
//
// Replaces `.OrderBy(o => o.TaskCreatedOn)` (Api/Services/TasksStoreManager.cs:136) and
// `.OrderByDescending(o => o.TaskCreatedOn)` (Api/Services/FakeTasksManager.cs:70,
// TasksStoreManager.cs:66).  LINQ's OrderBy is a stable sort on DateTime.Ticks, so equal keys
// keep their source order: the hit rows arrive in ascending row order and the LSD radix sort of
// radix.cuh (stable) of the 62-bit tick key preserves that; "descending" sorts the complemented
// key, which keeps ties in source order exactly as OrderByDescending does.  Small per-query
// results are ordered by one CTA with a bitonic network on (key, row) instead.  The sort runs
// over the hits only (H rows), never over the table.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tt_synth.h"
#include "radix.cuh"
#include "sweep.cuh"

namespace tt {

__global__ void row_keys_kernel(const uint32_t* rows, unsigned long long* keys, uint32_t* vals, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    keys[i] = rows[i];
    vals[i] = rows[i];
  }
}

__global__ void gather_ids_kernel(const uint32_t* rows, const uint4* task_id, uint4* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = ld_gather_v4(task_id + rows[i]);
}

// rows_out[i] = rows_sorted[i]; ids_out[i] = task_id[rows_sorted[i]]
__global__ void emit_rows_ids_kernel(const uint32_t* rows_sorted, const uint4* task_id, uint32_t* rows_out, uint4* ids_out,
                                     int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t r = rows_sorted[i];
    rows_out[i] = r;
    if (ids_out) ids_out[i] = ld_gather_v4(task_id + r);
  }
}

// Sorts a hit list by row number and re-gathers the TaskIds (used when two row-ordered runs
// have to be merged, and as the first step for unordered candidate lists).
// Returns 0 or a cudaError_t value; synchronises the stream.
inline int sort_rows_and_gather(cudaStream_t stream, SortScratch& sc, const uint4* task_id, uint32_t* d_rows,
                                uint4* d_ids, int64_t n) {
  if (n <= 0) return 0;
  cudaError_t e = sc.reserve(n);
  if (e != cudaSuccess) return (int)e;
  const unsigned grid = (unsigned)((n + 255) / 256);
  row_keys_kernel<<<grid, 256, 0, stream>>>(d_rows, sc.keys[0], sc.vals[0], n);
  int w = 0;
  e = radix_sort_pairs(stream, sc, n, 32, &w);
  if (e != cudaSuccess) return (int)e;
  emit_rows_ids_kernel<<<grid, 256, 0, stream>>>(sc.vals[w], task_id, d_rows, d_ids, n);
  e = cudaGetLastError();
  cudaError_t s2 = cudaStreamSynchronize(stream);
  return (int)(e != cudaSuccess ? e : s2);
}

// OrderBy(TaskCreatedOn), stable (TasksStoreManager.cs:136).  The sort buffers are filled already — sc.keys[0] = the
// hits' ticks, sc.vals[0] = their rows, in row order (the placement kernel of a listing sweep gathered them:
// sweep_owned, keys_too): min / max, the sort, and the list (rows and TaskIds) written in its final order.
// Returns 0 or a cudaError_t value; synchronises the stream.
inline int order_prefilled(cudaStream_t stream, SortScratch& sc, const uint4* task_id, uint32_t* d_rows, uint4* d_ids, int64_t n,
                           RadixSortInfo* info = nullptr) {
  if (n <= 0) return 0;
  int w = 0;
  cudaError_t e = radix_sort_auto(stream, sc, n, &w, info);
  if (e != cudaSuccess) return (int)e;
  const unsigned grid = (unsigned)((n + 255) / 256);
  emit_rows_ids_kernel<<<grid, 256, 0, stream>>>(sc.vals[w], task_id, d_rows, d_ids, n);
  e = cudaGetLastError();
  cudaError_t s2 = cudaStreamSynchronize(stream);
  return (int)(e != cudaSuccess ? e : s2);
}

// ---- ordering the creator filter's (query, row) pairs ------------------------------------------
//
// `.OrderByDescending(o => o.TaskCreatedOn)` per query (FakeTasksManager.cs:70,
// TasksStoreManager.cs:66), for a whole batch at once: pairs are ordered by
// (query, complemented ticks, row) — the row as last key reproduces LINQ's stable order for equal
// ticks, because the unordered candidates all come from one table scan.  The composite key is two
// 64-bit words: hi = query << 54 | key >> 8, lo = (key & 0xFF) << 32 | row (key = 62-bit ticks).
constexpr int kSmallOrderMax = 8192;     // pairs a single CTA sorts in shared memory (128 KB)
constexpr int kSmallOrderThreads = 1024;

__device__ __forceinline__ void pair_keys(unsigned long long pair, const uint64_t* created_raw, int descending,
                                          unsigned long long& hi, unsigned long long& lo) {
  const uint32_t row = (uint32_t)pair;
  const unsigned long long q = pair >> 32;
  unsigned long long k = created_raw[row] & TT_TICKS_MASK;
  if (descending) k = ~k & TT_TICKS_MASK;
  hi = (q << 54) | (k >> 8);
  lo = ((k & 0xFFull) << 32) | row;
}

// One CTA: build keys, bitonic sort in shared memory, write rows + TaskIds + per-query counts.
__global__ void __launch_bounds__(kSmallOrderThreads) order_pairs_small_kernel(
    const unsigned long long* pairs, int n, int n2, const uint64_t* created_raw, const uint4* task_id, int descending,
    uint32_t* out_rows, uint4* out_ids, uint32_t* q_count) {
  extern __shared__ __align__(8) unsigned char s_raw[];
  unsigned long long* s_hi = reinterpret_cast<unsigned long long*>(s_raw);
  unsigned long long* s_lo = s_hi + n2;
  for (int i = threadIdx.x; i < n2; i += kSmallOrderThreads) {
    unsigned long long hi = ~0ull, lo = ~0ull;
    if (i < n) pair_keys(pairs[i], created_raw, descending, hi, lo);
    s_hi[i] = hi;
    s_lo[i] = lo;
  }
  __syncthreads();
  for (int k = 2; k <= n2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n2; i += kSmallOrderThreads) {
        const int x = i ^ j;
        if (x > i) {
          const bool up = (i & k) == 0;
          const unsigned long long ah = s_hi[i], al = s_lo[i], bh = s_hi[x], bl = s_lo[x];
          const bool gt = ah > bh || (ah == bh && al > bl);
          if (gt == up) {
            s_hi[i] = bh; s_lo[i] = bl;
            s_hi[x] = ah; s_lo[x] = al;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < n; i += kSmallOrderThreads) {
    const uint32_t row = (uint32_t)s_lo[i];
    out_rows[i] = row;
    out_ids[i] = ld_gather_v4(task_id + row);
    atomicAdd(q_count + (uint32_t)(s_hi[i] >> 54), 1u);
  }
}

// One CTA per query: order its match list (at most kListCap rows) by (complemented ticks, row)
// and write it, with the TaskIds, at the query's offset in the batch result.
constexpr int kListOrderThreads = 512;
constexpr int kListOrderCap = 4096;  // == tt::kListCap; 4096 x 12 B = 48 KB of dynamic shared memory

__global__ void __launch_bounds__(kListOrderThreads) order_lists_kernel(const uint32_t* lists, const uint32_t* qcount,
                                                                        const uint64_t* created_raw,
                                                                        const uint4* task_id, int descending,
                                                                        uint32_t* out_rows, uint4* out_ids) {
  extern __shared__ __align__(8) unsigned char s_dyn[];
  unsigned long long* s_key = reinterpret_cast<unsigned long long*>(s_dyn);
  uint32_t* s_row = reinterpret_cast<uint32_t*>(s_dyn + (size_t)kListOrderCap * 8);
  __shared__ uint32_t s_first;
  const uint32_t q = blockIdx.x;
  const int n = (int)qcount[q];
  if (threadIdx.x < 32) {  // offset of this query = matches of the queries before it
    uint32_t c = 0;
    for (uint32_t k = threadIdx.x; k < q; k += 32) c += qcount[k];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xFFFFFFFFu, c, d);
    if (threadIdx.x == 0) s_first = c;
  }
  int n2 = 2;
  while (n2 < n) n2 <<= 1;
  for (int i = threadIdx.x; i < n2; i += kListOrderThreads) {
    unsigned long long k = ~0ull;
    uint32_t r = 0xFFFFFFFFu;
    if (i < n) {
      r = lists[(size_t)q * kListOrderCap + i];
      k = created_raw[r] & TT_TICKS_MASK;
      if (descending) k = ~k & TT_TICKS_MASK;
    }
    s_key[i] = k;
    s_row[i] = r;
  }
  __syncthreads();
  for (int k = 2; k <= n2; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < n2; i += kListOrderThreads) {
        const int x = i ^ j;
        if (x > i) {
          const bool up = (i & k) == 0;
          const unsigned long long ak = s_key[i], bk = s_key[x];
          const uint32_t ar = s_row[i], br = s_row[x];
          const bool gt = ak > bk || (ak == bk && ar > br);
          if (gt == up) {
            s_key[i] = bk; s_row[i] = br;
            s_key[x] = ak; s_row[x] = ar;
          }
        }
      }
      __syncthreads();
    }
  }
  const uint32_t first = s_first;
  for (int i = threadIdx.x; i < n; i += kListOrderThreads) {
    const uint32_t r = s_row[i];
    out_rows[first + i] = r;
    out_ids[first + i] = ld_gather_v4(task_id + r);
  }
}

// Large results: two stable radix sorts over an index permutation — first by the low key word,
// then by the high one — followed by one kernel that writes rows, TaskIds and per-query counts.
__global__ void pair_lo_keys_kernel(const unsigned long long* pairs, int64_t n, const uint64_t* created_raw,
                                    int descending, unsigned long long* keys, uint32_t* vals) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    unsigned long long hi, lo;
    pair_keys(pairs[i], created_raw, descending, hi, lo);
    keys[i] = lo;
    vals[i] = (uint32_t)i;
  }
}

__global__ void pair_hi_keys_kernel(const unsigned long long* pairs, const uint32_t* perm, int64_t n,
                                    const uint64_t* created_raw, int descending, unsigned long long* keys,
                                    uint32_t* vals) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    unsigned long long hi, lo;
    const uint32_t src = perm[i];
    pair_keys(pairs[src], created_raw, descending, hi, lo);
    keys[i] = hi;
    vals[i] = src;
  }
}

__global__ void pair_finish_kernel(const unsigned long long* pairs, const uint32_t* perm, int64_t n,
                                   const uint4* task_id, uint32_t* out_rows, uint4* out_ids, uint32_t* q_count) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const unsigned long long p = pairs[perm[i]];
    const uint32_t row = (uint32_t)p;
    out_rows[i] = row;
    out_ids[i] = ld_gather_v4(task_id + row);
    atomicAdd(q_count + (uint32_t)(p >> 32), 1u);
  }
}

// q_count (device, one word per query) must be zero on entry.  Returns 0 or a cudaError_t value.
inline int order_pairs(cudaStream_t stream, SortScratch& sc, const unsigned long long* pairs, int64_t n,
                       const uint64_t* created_raw, const uint4* task_id, bool descending, uint32_t* out_rows,
                       uint4* out_ids, uint32_t* q_count) {
  if (n <= 0) return 0;
  if (n <= kSmallOrderMax) {
    int n2 = 2;
    while (n2 < n) n2 <<= 1;
    const size_t smem = (size_t)n2 * 16;
    static bool opted = false;  // per process; harmless if repeated
    if (!opted) {
      cudaError_t e = cudaFuncSetAttribute(order_pairs_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           kSmallOrderMax * 16);
      if (e != cudaSuccess) return (int)e;
      opted = true;
    }
    order_pairs_small_kernel<<<1, kSmallOrderThreads, smem, stream>>>(pairs, (int)n, n2, created_raw, task_id,
                                                                     descending ? 1 : 0, out_rows, out_ids, q_count);
    return (int)cudaGetLastError();
  }
  if (n > 0xFFFFFFFFll) return (int)cudaErrorInvalidValue;
  cudaError_t e = sc.reserve(n);
  if (e != cudaSuccess) return (int)e;
  const unsigned grid = (unsigned)((n + 255) / 256);
  int w = 0;
  pair_lo_keys_kernel<<<grid, 256, 0, stream>>>(pairs, n, created_raw, descending ? 1 : 0, sc.keys[0], sc.vals[0]);
  e = radix_sort_pairs(stream, sc, n, 40, &w);
  if (e != cudaSuccess) return (int)e;
  // the permutation so far sits in vals[w]; the next sort starts from buffer 0 again (if w == 0
  // every thread reads perm[i] and writes vals[0][i] = the same value, so aliasing is harmless)
  const uint32_t* perm = sc.vals[w];
  pair_hi_keys_kernel<<<grid, 256, 0, stream>>>(pairs, perm, n, created_raw, descending ? 1 : 0, sc.keys[0], sc.vals[0]);
  e = radix_sort_pairs(stream, sc, n, 64, &w);
  if (e != cudaSuccess) return (int)e;
  pair_finish_kernel<<<grid, 256, 0, stream>>>(pairs, sc.vals[w], n, task_id, out_rows, out_ids, q_count);
  e = cudaGetLastError();
  cudaError_t s2 = cudaStreamSynchronize(stream);
  return (int)(e != cudaSuccess ? e : s2);
}

}  // namespace tt
