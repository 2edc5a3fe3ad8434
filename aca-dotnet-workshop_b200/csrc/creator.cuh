// creator.cuh — K4 creator filter (hash scan + exact string confirm) and K3 mark-by-id.
//
// K4 replaces `Where(t => t.TaskCreatedBy.Equals(createdBy))`
// (Api/Services/FakeTasksManager.cs:70; store-side EQ in Api/Services/TasksStoreManager.cs:56-61):
// the 64-bit creator_hash column is streamed once (8 B/row); a row whose hash equals the
// query's is a candidate and its dictionary string is compared byte for byte (ordinal,
// case-sensitive — equal UTF-8 bytes <=> equal UTF-16 code units), so hash collisions never
// reach the result.
// K3 replaces the per-task upsert loop of Api/Services/TasksStoreManager.cs:143-148: each
// listed TaskId is located by binary search over the TaskId-ordered rows and its overdue bit
// is set with one atomicOr.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tt_synth.h"

namespace tt {

constexpr int kCreatorThreads = 256;
constexpr int kCreatorUnroll = 4;  // 4 x 128-bit loads (8 rows) in flight per thread

__device__ __forceinline__ ulonglong2 ld_stream_u64x2(const ulonglong2* p) {
  ulonglong2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(r.x), "=l"(r.y) : "l"(p));
  return r;
}

__device__ __forceinline__ void creator_confirm(uint32_t row, int64_t n, const uint32_t* creator_ref,
                                                const uint32_t* creator_off, const uint8_t* creator_pool,
                                                const uint8_t* query, uint32_t qlen, uint32_t* out_rows, int64_t cap,
                                                unsigned long long* counters) {
  if ((int64_t)row >= n) return;  // padding row
  atomicAdd(counters + 1, 1ull);  // hash candidates (diagnostic)
  const uint32_t ref = creator_ref[row];
  const uint32_t off = creator_off[ref];
  const uint32_t len = creator_off[ref + 1] - off;
  if (len != qlen) return;
  for (uint32_t i = 0; i < len; ++i)
    if (creator_pool[off + i] != query[i]) return;
  unsigned long long pos = atomicAdd(counters + 0, 1ull);
  if ((int64_t)pos < cap) out_rows[pos] = row;
}

__global__ void __launch_bounds__(kCreatorThreads) creator_scan_kernel(
    const ulonglong2* __restrict__ hash2, int64_t n, int64_t n_pairs, unsigned long long qhash,
    const uint32_t* __restrict__ creator_ref, const uint32_t* __restrict__ creator_off,
    const uint8_t* __restrict__ creator_pool, const uint8_t* __restrict__ query, uint32_t qlen, uint32_t* out_rows,
    int64_t cap, unsigned long long* counters) {
  const int64_t stride = (int64_t)gridDim.x * kCreatorThreads;
  // n_pairs is a multiple of kCreatorThreads * kCreatorUnroll (the table is padded to 8192 rows)
  for (int64_t base = (int64_t)blockIdx.x * kCreatorThreads + threadIdx.x; base < n_pairs;
       base += stride * kCreatorUnroll) {
    ulonglong2 v[kCreatorUnroll];
#pragma unroll
    for (int j = 0; j < kCreatorUnroll; ++j) {
      int64_t i = base + j * stride;
      v[j] = i < n_pairs ? ld_stream_u64x2(hash2 + i) : make_ulonglong2(~qhash, ~qhash);
    }
#pragma unroll
    for (int j = 0; j < kCreatorUnroll; ++j) {
      int64_t i = base + j * stride;
      if (v[j].x == qhash)
        creator_confirm((uint32_t)(2 * i), n, creator_ref, creator_off, creator_pool, query, qlen, out_rows, cap, counters);
      if (v[j].y == qhash)
        creator_confirm((uint32_t)(2 * i + 1), n, creator_ref, creator_off, creator_pool, query, qlen, out_rows, cap,
                        counters);
    }
  }
}

// Runs the filter; *found = number of exact matches (may exceed cap: then only cap rows were
// stored and the caller retries with a larger buffer).  Output order is unspecified.
// d_small holds the two counters, d_query (>= qlen bytes) receives the query string.
inline cudaError_t creator_filter(cudaStream_t stream, int sm_count, const uint64_t* creator_hash,
                                  const uint32_t* creator_ref, const uint32_t* creator_off,
                                  const uint8_t* creator_pool, int64_t n, int64_t n_pad, const uint8_t* query,
                                  uint32_t qlen, uint8_t* d_query, uint32_t* out_rows, int64_t cap,
                                  unsigned long long* d_small, unsigned long long* h_small, int64_t* found) {
  *found = 0;
  if (n == 0) return cudaSuccess;
  cudaError_t e = cudaMemsetAsync(d_small, 0, 16, stream);
  if (e != cudaSuccess) return e;
  if (qlen) {
    e = cudaMemcpyAsync(d_query, query, qlen, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) return e;
  }
  const unsigned long long qhash = tt_hash64(query, qlen);
  const int64_t n_pairs = n_pad / 2;
  int64_t blocks = (n_pairs + (int64_t)kCreatorThreads * kCreatorUnroll - 1) / ((int64_t)kCreatorThreads * kCreatorUnroll);
  const int64_t max_blocks = (int64_t)sm_count * 8;
  if (blocks > max_blocks) blocks = max_blocks;
  creator_scan_kernel<<<(unsigned)blocks, kCreatorThreads, 0, stream>>>(
      reinterpret_cast<const ulonglong2*>(creator_hash), n, n_pairs, qhash, creator_ref, creator_off, creator_pool,
      d_query, qlen, out_rows, cap, d_small);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  e = cudaMemcpyAsync(h_small, d_small, 16, cudaMemcpyDeviceToHost, stream);
  if (e != cudaSuccess) return e;
  e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) return e;
  *found = (int64_t)h_small[0];
  return cudaSuccess;
}

// ---- K3 -------------------------------------------------------------------------------------

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

// counters[0] += ids found (bit set), counters[1] += ids with no row
__global__ void mark_by_id_kernel(const uint4* ids, int64_t n_ids, const uint4* task_id, const uint32_t* id_order,
                                  int64_t n_rows, uint32_t* overdue, unsigned long long* counters) {
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
  if (lo < n_rows) {
    uint32_t r = id_order ? id_order[lo] : (uint32_t)lo;
    if (cmp_id(task_id[r], key) == 0) {
      atomicOr(overdue + (r >> 5), 1u << (r & 31));
      hit = true;
    }
  }
  atomicAdd(counters + (hit ? 0 : 1), 1ull);
}

}  // namespace tt
