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
