// order.cuh — K5: order a hit list by TaskCreatedOn and gather the TaskIds in that order.
//
// Replaces `.OrderBy(o => o.TaskCreatedOn)` (Api/Services/TasksStoreManager.cs:136) and
// `.OrderByDescending(o => o.TaskCreatedOn)` (Api/Services/FakeTasksManager.cs:70,
// TasksStoreManager.cs:66).  LINQ's OrderBy is a stable sort on DateTime.Ticks, so equal keys
// keep their source order: the hit rows arrive in ascending row order and an LSD radix sort
// (stable) of the 62-bit tick key preserves that; "descending" sorts the complemented key,
// which keeps ties in source order exactly as OrderByDescending does.
// The radix sort itself is CUB's DeviceRadixSort (library code, compiled here for sm_100a);
// it runs over the hits only (H rows), never over the table.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cub/device/device_radix_sort.cuh>

#include "../../include/tt_synth.h"
#include "sweep.cuh"

namespace tt {

__global__ void order_keys_kernel(const uint32_t* rows, const uint64_t* created_raw, uint64_t* keys, int64_t n,
                                  int descending) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    uint64_t k = created_raw[rows[i]] & TT_TICKS_MASK;
    keys[i] = descending ? (~k & TT_TICKS_MASK) : k;
  }
}

__global__ void gather_ids_kernel(const uint32_t* rows, const uint4* task_id, uint4* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = ld_gather_v4(task_id + rows[i]);
}

// Returns 0 or a cudaError_t value.
inline int order_by_created(cudaStream_t stream, const uint64_t* created_raw, const uint4* task_id, uint32_t* d_rows,
                            uint4* d_ids, int64_t n, bool descending, bool rows_unordered = false) {
  if (n <= 0) return 0;
  if (n > 0x7FFFFFFF) return (int)cudaErrorInvalidValue;
  const int num = (int)n;
  uint64_t *keys_in = nullptr, *keys_out = nullptr;
  uint32_t* rows_tmp = nullptr;
  void* temp = nullptr;
  size_t temp_pairs = 0, temp_keys = 0;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, temp_pairs, keys_in, keys_out, d_rows, rows_tmp, num, 0, 62, stream);
  if (e == cudaSuccess) e = cub::DeviceRadixSort::SortKeys(nullptr, temp_keys, d_rows, rows_tmp, num, 0, 32, stream);
  size_t temp_bytes = temp_pairs > temp_keys ? temp_pairs : temp_keys;
  if (e == cudaSuccess) e = cudaMalloc(&keys_in, (size_t)n * 8);
  if (e == cudaSuccess) e = cudaMalloc(&keys_out, (size_t)n * 8);
  if (e == cudaSuccess) e = cudaMalloc(&rows_tmp, (size_t)n * 4);
  if (e == cudaSuccess) e = cudaMalloc(&temp, temp_bytes ? temp_bytes : 16);
  const unsigned grid = (unsigned)((n + 255) / 256);
  if (e == cudaSuccess && rows_unordered) {
    e = cub::DeviceRadixSort::SortKeys(temp, temp_bytes, d_rows, rows_tmp, num, 0, 32, stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_rows, rows_tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream);
  }
  if (e == cudaSuccess) {
    order_keys_kernel<<<grid, 256, 0, stream>>>(d_rows, created_raw, keys_in, n, descending ? 1 : 0);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess)
    e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, d_rows, rows_tmp, num, 0, 62, stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_rows, rows_tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream);
  if (e == cudaSuccess) {
    gather_ids_kernel<<<grid, 256, 0, stream>>>(d_rows, task_id, d_ids, n);
    e = cudaGetLastError();
  }
  cudaError_t s = cudaStreamSynchronize(stream);
  if (e == cudaSuccess) e = s;
  cudaFree(keys_in);
  cudaFree(keys_out);
  cudaFree(rows_tmp);
  cudaFree(temp);
  return (int)e;
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

__global__ void pair_keys_kernel(const unsigned long long* pairs, int64_t n, const uint64_t* created_raw, int descending,
                                 unsigned long long* hi, unsigned long long* lo) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pair_keys(pairs[i], created_raw, descending, hi[i], lo[i]);
}

__global__ void pair_finish_kernel(const unsigned long long* hi, const unsigned long long* lo, int64_t n,
                                   const uint4* task_id, uint32_t* out_rows, uint4* out_ids, uint32_t* q_count) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t row = (uint32_t)lo[i];
    out_rows[i] = row;
    out_ids[i] = ld_gather_v4(task_id + row);
    atomicAdd(q_count + (uint32_t)(hi[i] >> 54), 1u);
  }
}

// q_count (device, one word per query) must be zero on entry.  Returns 0 or a cudaError_t value.
inline int order_pairs(cudaStream_t stream, const unsigned long long* pairs, int64_t n, const uint64_t* created_raw,
                       const uint4* task_id, bool descending, uint32_t* out_rows, uint4* out_ids, uint32_t* q_count) {
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
  if (n > 0x7FFFFFFF) return (int)cudaErrorInvalidValue;
  // large result: two stable LSD passes over the two key words (CUB DeviceRadixSort, hits only)
  const int num = (int)n;
  unsigned long long *hi = nullptr, *lo = nullptr, *hi2 = nullptr, *lo2 = nullptr;
  void* temp = nullptr;
  size_t temp_bytes = 0;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, lo, lo2, hi, hi2, num, 0, 64, stream);
  if (e == cudaSuccess) e = cudaMalloc(&hi, (size_t)n * 8);
  if (e == cudaSuccess) e = cudaMalloc(&lo, (size_t)n * 8);
  if (e == cudaSuccess) e = cudaMalloc(&hi2, (size_t)n * 8);
  if (e == cudaSuccess) e = cudaMalloc(&lo2, (size_t)n * 8);
  if (e == cudaSuccess) e = cudaMalloc(&temp, temp_bytes ? temp_bytes : 16);
  const unsigned grid = (unsigned)((n + 255) / 256);
  if (e == cudaSuccess) {
    pair_keys_kernel<<<grid, 256, 0, stream>>>(pairs, n, created_raw, descending ? 1 : 0, hi, lo);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, lo, lo2, hi, hi2, num, 0, 40, stream);
  if (e == cudaSuccess) e = cub::DeviceRadixSort::SortPairs(temp, temp_bytes, hi2, hi, lo2, lo, num, 0, 64, stream);
  if (e == cudaSuccess) {
    pair_finish_kernel<<<grid, 256, 0, stream>>>(hi, lo, n, task_id, out_rows, out_ids, q_count);
    e = cudaGetLastError();
  }
  cudaError_t s2 = cudaStreamSynchronize(stream);
  if (e == cudaSuccess) e = s2;
  cudaFree(hi);
  cudaFree(lo);
  cudaFree(hi2);
  cudaFree(lo2);
  cudaFree(temp);
  return (int)e;
}

// Sorts a hit list by row number and re-gathers the TaskIds (used when two row-ordered runs
// have to be merged).  Returns 0 or a cudaError_t value.
inline int sort_rows_and_gather(cudaStream_t stream, const uint4* task_id, uint32_t* d_rows, uint4* d_ids, int64_t n) {
  if (n <= 1) return 0;
  if (n > 0x7FFFFFFF) return (int)cudaErrorInvalidValue;
  const int num = (int)n;
  uint32_t* tmp = nullptr;
  void* temp = nullptr;
  size_t temp_bytes = 0;
  cudaError_t e = cub::DeviceRadixSort::SortKeys(nullptr, temp_bytes, d_rows, tmp, num, 0, 32, stream);
  if (e == cudaSuccess) e = cudaMalloc(&tmp, (size_t)n * 4);
  if (e == cudaSuccess) e = cudaMalloc(&temp, temp_bytes ? temp_bytes : 16);
  if (e == cudaSuccess) e = cub::DeviceRadixSort::SortKeys(temp, temp_bytes, d_rows, tmp, num, 0, 32, stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_rows, tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream);
  if (e == cudaSuccess) {
    gather_ids_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(d_rows, task_id, d_ids, n);
    e = cudaGetLastError();
  }
  cudaError_t s = cudaStreamSynchronize(stream);
  if (e == cudaSuccess) e = s;
  cudaFree(tmp);
  cudaFree(temp);
  return (int)e;
}

}  // namespace tt
