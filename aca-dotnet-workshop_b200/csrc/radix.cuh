// radix.cuh — stable LSD radix sort of (64-bit key, 32-bit value) pairs, written for the hit
// lists of this library (K5).  LINQ's OrderBy / OrderByDescending are stable sorts
// (Api/Services/TasksStoreManager.cs:136,66; Api/Services/FakeTasksManager.cs:70), so the sort
// must be stable too: equal TaskCreatedOn values keep their list order.
//
// One pass per 8-bit digit, three launches per pass:
//   radix_hist_kernel    per-tile digit counts, digit-major (shared-memory atomics), and the scan's carries
//   tile_offsets_kernel (sweep.cuh)  exclusive scan of those counts
//   radix_scatter_kernel stable scatter: inside a tile every warp owns a contiguous chunk and
//                        walks it 32 elements at a time; eight ballots (one per digit bit) group
//                        the lanes of a round by digit, the group leader bumps the warp's digit
//                        counter, so an element's rank = counter before the round + its position
//                        in the group.  (match.any does the same grouping but measured 4x slower
//                        on high-entropy digits.)
// Digits on which all keys agree are skipped (the caller passes the number of significant bits),
// e.g. 7 passes instead of 8 for creation times that span two years.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "sweep.cuh"

namespace tt {

constexpr int kRadixThreads = 256;
constexpr int kRadixWarps = kRadixThreads / 32;
constexpr int kRadixItems = 8;                                 // rounds per warp
constexpr int kRadixTile = kRadixThreads * kRadixItems;        // 2048 elements per CTA
constexpr int kRadixDigits = 256;

// Lanes of the warp (among `live`) whose 8-bit digit equals this lane's: one ballot per bit.
__device__ __forceinline__ uint32_t digit_peers(uint32_t d, uint32_t live) {
  uint32_t peers = live;
#pragma unroll
  for (int b = 0; b < 8; ++b) {
    const bool bit = (d >> b) & 1u;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, bit);
    peers &= bit ? m : ~m;
  }
  return peers;
}

// hist[d * num_tiles + tile] = elements of `tile` whose digit is d (digit-major, the order of
// the exclusive scan that turns counts into output offsets).  The scan's carries are accumulated on the
// way: group_sum[e / 1024] += hist entry e (zeroed by the caller), so no separate reduction pass runs.
// LOW_ENTROPY = false: plain shared-memory atomics (the counts need no order).  LOW_ENTROPY = true: the
// pass's digit has only a few significant bits (the top digit of the keys), nearly every key of a tile
// falls into the same few bins and plain atomics would serialise, so the lanes of a warp are grouped by
// digit first (ballots) and each group adds once.
template <bool LOW_ENTROPY>
__global__ void __launch_bounds__(kRadixThreads) radix_hist_kernel(const unsigned long long* keys, int64_t n, int shift,
                                                                    uint32_t* hist, uint32_t num_tiles, uint32_t* group_sum) {
  __shared__ uint32_t s_hist[kRadixDigits];
  const uint32_t tile = blockIdx.x;
  s_hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)tile * kRadixTile;
  unsigned long long k[kRadixItems];
#pragma unroll
  for (int r = 0; r < kRadixItems; ++r) {
    const int64_t i = base + r * kRadixThreads + threadIdx.x;
    k[r] = i < n ? keys[i] : 0ull;
  }
#pragma unroll
  for (int r = 0; r < kRadixItems; ++r) {
    const int64_t i = base + r * kRadixThreads + threadIdx.x;
    const bool valid = i < n;
    const uint32_t d = (uint32_t)(k[r] >> shift) & 255u;
    if (LOW_ENTROPY) {
      const uint32_t peers = digit_peers(d, __ballot_sync(0xFFFFFFFFu, valid));
      if (valid && (threadIdx.x & 31u) == (uint32_t)(__ffs(peers) - 1)) atomicAdd(&s_hist[d], (uint32_t)__popc(peers));
    } else if (valid) {
      atomicAdd(&s_hist[d], 1u);
    }
  }
  __syncthreads();
  const uint32_t c = s_hist[threadIdx.x];
  const uint32_t e = threadIdx.x * num_tiles + tile;
  hist[e] = c;
  if (c) atomicAdd(group_sum + e / kTilesPerGroup, c);
}

// group_sum[g] = sum of counts[g*1024 .. g*1024+1023] (the scan kernel's carries)
__global__ void __launch_bounds__(kTilesPerGroup) group_reduce_kernel(const uint32_t* counts, uint32_t num,
                                                                      uint32_t* group_sum) {
  __shared__ uint32_t s_w[kTilesPerGroup / 32];
  const uint32_t e = blockIdx.x * kTilesPerGroup + threadIdx.x;
  uint32_t c = warp_sum(e < num ? counts[e] : 0u);
  if ((threadIdx.x & 31u) == 0) s_w[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x < 32) {
    c = warp_sum(s_w[threadIdx.x]);
    if (threadIdx.x == 0) group_sum[blockIdx.x] = c;
  }
}

__global__ void __launch_bounds__(kRadixThreads) radix_scatter_kernel(
    const unsigned long long* keys_in, const uint32_t* vals_in, unsigned long long* keys_out, uint32_t* vals_out,
    int64_t n, int shift, const uint32_t* offsets, uint32_t num_tiles) {
  __shared__ uint32_t s_count[kRadixWarps][kRadixDigits];
  __shared__ unsigned long long s_keys[kRadixTile];
  __shared__ uint32_t s_vals[kRadixTile];
  __shared__ uint32_t s_gbase[kRadixDigits];  // output position of staged element j = s_gbase[digit] + j
  __shared__ uint32_t s_wtot[kRadixWarps];
  const uint32_t tile = blockIdx.x, lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < kRadixWarps * kRadixDigits; i += kRadixThreads) (&s_count[0][0])[i] = 0;
  __syncthreads();

  const int64_t tile_base = (int64_t)tile * kRadixTile;
  const int64_t base = tile_base + (int64_t)warp * (32 * kRadixItems);
  unsigned long long key[kRadixItems];
  uint32_t val[kRadixItems], rank[kRadixItems];
#pragma unroll
  for (int r = 0; r < kRadixItems; ++r) {
    const int64_t i = base + r * 32 + lane;
    const bool valid = i < n;
    key[r] = valid ? keys_in[i] : 0ull;
    val[r] = valid ? vals_in[i] : 0u;
  }
#pragma unroll
  for (int r = 0; r < kRadixItems; ++r) {
    const bool valid = base + r * 32 + lane < n;
    const uint32_t live = __ballot_sync(0xFFFFFFFFu, valid);
    const uint32_t d = (uint32_t)(key[r] >> shift) & 255u;
    const uint32_t peers = digit_peers(d, live);
    rank[r] = 0;
    if (valid) {
      const uint32_t leader = __ffs(peers) - 1;
      uint32_t before = 0;
      if (lane == leader) {
        before = s_count[warp][d];
        s_count[warp][d] = before + __popc(peers);
      }
      before = __shfl_sync(peers, before, leader);
      rank[r] = before + __popc(peers & ((1u << lane) - 1u));
    }
    __syncwarp();
  }
  __syncthreads();
  {
    // thread d: this tile's count of digit d -> start of the digit inside the staged tile
    // (exclusive scan over the digits), then the start of each warp's share of it
    const uint32_t d = threadIdx.x;
    uint32_t tot = 0;
#pragma unroll
    for (int w = 0; w < kRadixWarps; ++w) tot += s_count[w][d];
    uint32_t incl = tot;
#pragma unroll
    for (int k = 1; k < 32; k <<= 1) {
      const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, k);
      if (lane >= (uint32_t)k) incl += up;
    }
    if (lane == 31) s_wtot[warp] = incl;
    __syncthreads();
    uint32_t start = incl - tot;
    for (uint32_t w = 0; w < warp; ++w) start += s_wtot[w];
    s_gbase[d] = offsets[d * num_tiles + tile] - start;
    uint32_t run = start;
#pragma unroll
    for (int w = 0; w < kRadixWarps; ++w) {
      const uint32_t c = s_count[w][d];
      s_count[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
  // stage the tile in digit order, then write it out with consecutive threads on consecutive
  // elements: every digit's run lands in one contiguous piece of the output
#pragma unroll
  for (int r = 0; r < kRadixItems; ++r) {
    if (base + r * 32 + lane < n) {
      const uint32_t j = s_count[warp][(uint32_t)(key[r] >> shift) & 255u] + rank[r];
      s_keys[j] = key[r];
      s_vals[j] = val[r];
    }
  }
  __syncthreads();
  const int64_t left = n - tile_base;
  const uint32_t count = left < kRadixTile ? (uint32_t)left : (uint32_t)kRadixTile;
  for (uint32_t j = threadIdx.x; j < count; j += kRadixThreads) {
    const unsigned long long k = s_keys[j];
    const uint32_t pos = s_gbase[(uint32_t)(k >> shift) & 255u] + j;
    keys_out[pos] = k;
    vals_out[pos] = s_vals[j];
  }
}

// min / max of the keys -> mm[0], mm[1] (mm[0] preset to ~0, mm[1] to 0)
__global__ void radix_minmax_kernel(const unsigned long long* keys, int64_t n, unsigned long long* mm) {
  unsigned long long lo = ~0ull, hi = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const unsigned long long k = keys[i];
    lo = k < lo ? k : lo;
    hi = k > hi ? k : hi;
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const unsigned long long a = __shfl_xor_sync(0xFFFFFFFFu, lo, d), b = __shfl_xor_sync(0xFFFFFFFFu, hi, d);
    lo = a < lo ? a : lo;
    hi = b > hi ? b : hi;
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mm + 0, lo);
    atomicMax(mm + 1, hi);
  }
}

// Grow-only device scratch of the sort, owned by the table.
struct SortScratch {
  unsigned long long* keys[2] = {nullptr, nullptr};
  uint32_t* vals[2] = {nullptr, nullptr};
  uint32_t* hist = nullptr;     // [256][tiles]
  uint32_t* offsets = nullptr;  // same shape, exclusive scan of hist
  uint32_t* group_sum = nullptr;
  unsigned long long* mm = nullptr;  // 2 words (device) for min / max
  int64_t cap = 0;
  void release() {
    for (int k = 0; k < 2; ++k) {
      cudaFree(keys[k]);
      cudaFree(vals[k]);
      keys[k] = nullptr;
      vals[k] = nullptr;
    }
    cudaFree(hist);
    cudaFree(offsets);
    cudaFree(group_sum);
    cudaFree(mm);
    hist = offsets = group_sum = nullptr;
    mm = nullptr;
    cap = 0;
  }
  cudaError_t reserve(int64_t n) {
    if (n <= cap) return cudaSuccess;
    release();
    const int64_t want = n + n / 4 + 4096;
    const size_t tiles = (size_t)((want + kRadixTile - 1) / kRadixTile);
    const size_t entries = tiles * kRadixDigits;
    cudaError_t e = cudaSuccess;
    for (int k = 0; k < 2 && e == cudaSuccess; ++k) {
      e = cudaMalloc(&keys[k], (size_t)want * 8);
      if (e == cudaSuccess) e = cudaMalloc(&vals[k], (size_t)want * 4);
    }
    if (e == cudaSuccess) e = cudaMalloc(&hist, entries * 4);
    if (e == cudaSuccess) e = cudaMalloc(&offsets, entries * 4);
    if (e == cudaSuccess) e = cudaMalloc(&group_sum, (entries / kTilesPerGroup + 2) * 4);
    if (e == cudaSuccess) e = cudaMalloc(&mm, 16);
    if (e != cudaSuccess) {
      release();
      return e;
    }
    cap = want;
    return cudaSuccess;
  }
};

// Sorts the n pairs in (s.keys[0], s.vals[0]) by the low `bits` bits of the key, stably.
// *which = index (0 or 1) of the buffer pair that holds the result.  Asynchronous on `stream`.
inline cudaError_t radix_sort_pairs(cudaStream_t stream, SortScratch& s, int64_t n, int bits, int* which) {
  *which = 0;
  if (n <= 1 || bits <= 0) return cudaSuccess;
  const uint32_t tiles = (uint32_t)((n + kRadixTile - 1) / kRadixTile);
  const uint32_t entries = tiles * kRadixDigits;
  const uint32_t groups = (entries + kTilesPerGroup - 1) / kTilesPerGroup;
  int cur = 0;
  for (int shift = 0; shift < bits; shift += 8) {
    cudaError_t em = cudaMemsetAsync(s.group_sum, 0, (size_t)groups * 4, stream);
    if (em != cudaSuccess) return em;
    // the ballot-grouped form for every pass: measured on B200 at 25 M keys 0.090 ms per pass against 0.107 ms for
    // plain shared-memory atomics on full-entropy digits (and far worse on the top digit, where a tile's keys fall
    // into a few bins); what bounds either is the digit-major store of 256 counts per tile, not the counting
    radix_hist_kernel<true><<<tiles, kRadixThreads, 0, stream>>>(s.keys[cur], n, shift, s.hist, tiles, s.group_sum);
    tile_offsets_kernel<<<groups, kTilesPerGroup, 0, stream>>>(s.hist, s.group_sum, entries, s.offsets);
    radix_scatter_kernel<<<tiles, kRadixThreads, 0, stream>>>(s.keys[cur], s.vals[cur], s.keys[cur ^ 1], s.vals[cur ^ 1], n,
                                                              shift, s.offsets, tiles);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    cur ^= 1;
  }
  *which = cur;
  return cudaSuccess;
}

// s.mm = {~0, 0}: the start state of a min / max collected by the kernel that produces the keys.
inline cudaError_t radix_minmax_reset(cudaStream_t stream, SortScratch& s) {
  static const unsigned long long init[2] = {~0ull, 0ull};
  return cudaMemcpyAsync(s.mm, init, 16, cudaMemcpyHostToDevice, stream);
}

inline cudaError_t radix_bits_from_minmax(cudaStream_t stream, SortScratch& s, int* bits);

// Number of low bits on which the keys in s.keys[0][0..n) can differ (synchronises the stream).
inline cudaError_t radix_significant_bits(cudaStream_t stream, SortScratch& s, int64_t n, int* bits) {
  cudaError_t e = radix_minmax_reset(stream, s);
  if (e != cudaSuccess) return e;
  const unsigned grid = (unsigned)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  radix_minmax_kernel<<<grid ? grid : 1, 256, 0, stream>>>(s.keys[0], n, s.mm);
  return radix_bits_from_minmax(stream, s, bits);
}

// The same from a min / max already collected in s.mm (synchronises the stream).
inline cudaError_t radix_bits_from_minmax(cudaStream_t stream, SortScratch& s, int* bits) {
  unsigned long long mm[2];
  cudaError_t e = cudaMemcpyAsync(mm, s.mm, 16, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) return e;
  unsigned long long diff = mm[0] ^ mm[1];
  int b = 0;
  while (diff) {
    ++b;
    diff >>= 1;
  }
  *bits = b;
  return cudaSuccess;
}

}  // namespace tt
