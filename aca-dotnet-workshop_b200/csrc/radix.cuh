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
// e.g. 7 passes instead of 8 for creation times that span two years.  radix_sort_by_prefix goes further
// for wide keys: LSD passes over the top 24-40 significant bits of (key - smallest key) only, then a check of
// neighbours with equal prefixes, whose (rare) out-of-order groups are put right in place.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <algorithm>

#include "sweep.cuh"

namespace tt {

constexpr int kRadixThreads = 256;
constexpr int kRadixWarps = kRadixThreads / 32;
constexpr int kRadixItemsMax = 16;                             // rounds per warp: 8 (2048 elements per CTA) or 16 (4096)
constexpr int kRadixTileMax = kRadixThreads * kRadixItemsMax;
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
template <bool LOW_ENTROPY, int kRadixItems>
__global__ void __launch_bounds__(kRadixThreads) radix_hist_kernel(const unsigned long long* keys, int64_t n,
                                                                    unsigned long long base0, int shift, uint32_t* hist,
                                                                    uint32_t num_tiles, uint32_t* group_sum) {
  constexpr int kRadixTile = kRadixThreads * kRadixItems;
  __shared__ uint32_t s_hist[kRadixDigits];
  const uint32_t tile = blockIdx.x;
  s_hist[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)tile * kRadixTile;
  unsigned long long k[kRadixItems];
#pragma unroll
  for (int r = 0; r < kRadixItems; ++r) {
    const int64_t i = base + r * kRadixThreads + threadIdx.x;
    k[r] = i < n ? keys[i] - base0 : 0ull;
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

template <int kRadixItems>
__global__ void __launch_bounds__(kRadixThreads) radix_scatter_kernel(
    const unsigned long long* keys_in, const uint32_t* vals_in, unsigned long long* keys_out, uint32_t* vals_out,
    int64_t n, unsigned long long base0, int shift, const uint32_t* offsets, uint32_t num_tiles) {
  constexpr int kRadixTile = kRadixThreads * kRadixItems;
  extern __shared__ __align__(16) unsigned char s_radix[];
  unsigned long long* s_keys = reinterpret_cast<unsigned long long*>(s_radix);                  // [kRadixTile]
  uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys + kRadixTile);                          // [kRadixTile]
  uint32_t(*s_count)[kRadixDigits] = reinterpret_cast<uint32_t(*)[kRadixDigits]>(s_vals + kRadixTile);  // [warps][256]
  uint32_t* s_gbase = &s_count[kRadixWarps][0];  // [256]: output position of staged element j = s_gbase[digit] + j
  uint32_t* s_wtot = s_gbase + kRadixDigits;     // [warps]
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
    key[r] = valid ? keys_in[i] - base0 : 0ull;
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
    keys_out[pos] = k + base0;
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
    const size_t tiles = (size_t)((want + kRadixThreads * 8 - 1) / (kRadixThreads * 8));  // the smaller tile
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

// Elements per CTA of the two pass kernels: 8 per thread (2048 per CTA).  TT_RADIX_ITEMS=16 selects 4096 per CTA
// (a digit's run in the output is 16 keys on average instead of 8 and the count table half as large), which
// measured SLOWER on B200 — 5.45 against 4.47 ms for the ordered list of 25 M hits: 109 registers per thread leave
// two CTAs per SM.
inline int radix_items() {
  static const int items = [] {
    const char* e = getenv("TT_RADIX_ITEMS");
    return e && atoi(e) == 16 ? 16 : 8;
  }();
  return items;
}

template <int ITEMS>
inline cudaError_t radix_pass(cudaStream_t stream, SortScratch& s, int64_t n, unsigned long long base0, int shift, int cur) {
  constexpr int kTile = kRadixThreads * ITEMS;
  constexpr size_t kSmem = (size_t)kTile * 12 + (size_t)(kRadixWarps + 1) * kRadixDigits * 4 + kRadixWarps * 4;
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (kSmem > 48 * 1024 && dev >= 0 && dev < 64 && !attr_set[dev]) {
    cudaError_t ea = cudaFuncSetAttribute(radix_scatter_kernel<ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
    if (ea != cudaSuccess) return ea;
    attr_set[dev] = true;
  }
  const uint32_t tiles = (uint32_t)((n + kTile - 1) / kTile);
  const uint32_t entries = tiles * kRadixDigits;
  const uint32_t groups = (entries + kTilesPerGroup - 1) / kTilesPerGroup;
  cudaError_t em = cudaMemsetAsync(s.group_sum, 0, (size_t)groups * 4, stream);
  if (em != cudaSuccess) return em;
  // the ballot-grouped form for every pass: measured on B200 at 25 M keys 0.090 ms per pass against 0.107 ms for
  // plain shared-memory atomics on full-entropy digits (and far worse on the top digit, where a tile's keys fall
  // into a few bins); what bounds either is the digit-major store of 256 counts per tile, not the counting
  radix_hist_kernel<true, ITEMS><<<tiles, kRadixThreads, 0, stream>>>(s.keys[cur], n, base0, shift, s.hist, tiles, s.group_sum);
  tile_offsets_kernel<<<groups, kTilesPerGroup, 0, stream>>>(s.hist, s.group_sum, entries, s.offsets);
  radix_scatter_kernel<ITEMS><<<tiles, kRadixThreads, kSmem, stream>>>(s.keys[cur], s.vals[cur], s.keys[cur ^ 1], s.vals[cur ^ 1],
                                                                      n, base0, shift, s.offsets, tiles);
  return cudaGetLastError();
}

// Stable LSD passes over bits [lo, hi) of (key - base0), from buffer pair *cur; *cur follows the data.
inline cudaError_t radix_passes(cudaStream_t stream, SortScratch& s, int64_t n, unsigned long long base0, int lo, int hi,
                                int* cur) {
  for (int shift = lo; shift < hi; shift += 8) {
    const cudaError_t e = radix_items() == 8 ? radix_pass<8>(stream, s, n, base0, shift, *cur)
                                             : radix_pass<16>(stream, s, n, base0, shift, *cur);
    if (e != cudaSuccess) return e;
    *cur ^= 1;
  }
  return cudaSuccess;
}

// Sorts the n pairs in (s.keys[0], s.vals[0]) by the low `bits` bits of the key, stably.
// *which = index (0 or 1) of the buffer pair that holds the result.  Asynchronous on `stream`.
inline cudaError_t radix_sort_pairs(cudaStream_t stream, SortScratch& s, int64_t n, int bits, int* which) {
  *which = 0;
  if (n <= 1 || bits <= 0) return cudaSuccess;
  return radix_passes(stream, s, n, 0ull, 0, bits, which);
}

// s.mm = {~0, 0}: the start state of a min / max collected by the kernel that produces the keys.
inline cudaError_t radix_minmax_reset(cudaStream_t stream, SortScratch& s) {
  static const unsigned long long init[2] = {~0ull, 0ull};
  return cudaMemcpyAsync(s.mm, init, 16, cudaMemcpyHostToDevice, stream);
}

// ---- wide keys: sort a prefix of the significant bits, then put right what the low bits disturb ------------
// After stable passes over the top `prefix` bits [shift0, bits) of (key - base0) the pairs are in order wherever the prefixes
// differ.  Elements with EQUAL prefixes still stand in their input order: that is already the final order when they
// are equal altogether (ties) or were ascending on input (a table filled in creation order), and otherwise the group
// is tiny — 25 M keys drawn from two years of 100 ns ticks meet about 70 000 times in 2^32 prefixes.  One kernel looks
// for neighbours that are out of order and reports the head of each such group (the thread at the group's FIRST
// inversion does, so a group is reported once); a second kernel sorts every reported group by insertion (stable).
// Groups longer than kFixMaxGroup or more reports than the list holds raise ctl[1]: the caller then finishes with
// ordinary passes.
constexpr int kFixMaxGroup = 32;

__global__ void radix_fix_find_kernel(const unsigned long long* keys, int64_t n, unsigned long long base0, int shift0,
                                      uint32_t* list, uint32_t list_cap, unsigned long long* ctl) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= n) return;
  const unsigned long long a = keys[i] - base0, b = keys[i + 1] - base0;
  if ((a >> shift0) != (b >> shift0) || a <= b) return;
  const unsigned long long prefix = a >> shift0;
  int64_t j = i;
  unsigned long long right = a;
  for (int steps = 0; j > 0; ++steps) {
    const unsigned long long c = keys[j - 1] - base0;
    if ((c >> shift0) != prefix) break;
    if (c > right) return;  // an earlier inversion of the same group: its thread reports
    if (steps >= kFixMaxGroup) {
      ctl[1] = 1;
      return;
    }
    right = c;
    --j;
  }
  const unsigned long long slot = atomicAdd(ctl, 1ull);
  if (slot < list_cap) list[slot] = (uint32_t)j;
  else ctl[1] = 1;
}

__global__ void radix_fix_sort_kernel(unsigned long long* keys, uint32_t* vals, int64_t n, unsigned long long base0, int shift0,
                                      const uint32_t* list, uint32_t list_cap, unsigned long long* ctl) {
  const unsigned long long count = ctl[0] < list_cap ? ctl[0] : list_cap;
  for (unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; t < count;
       t += (unsigned long long)gridDim.x * blockDim.x) {
    const int64_t h = list[t];
    const unsigned long long prefix = (keys[h] - base0) >> shift0;
    int64_t e = h + 1;
    while (e < n && ((keys[e] - base0) >> shift0) == prefix && e - h <= kFixMaxGroup) ++e;
    if (e - h > kFixMaxGroup) {
      ctl[1] = 1;
      continue;
    }
    for (int64_t x = h + 1; x < e; ++x) {
      const unsigned long long kx = keys[x];
      const uint32_t vx = vals[x];
      int64_t y = x;
      while (y > h && keys[y - 1] > kx) {
        keys[y] = keys[y - 1];
        vals[y] = vals[y - 1];
        --y;
      }
      keys[y] = kx;
      vals[y] = vx;
    }
  }
}

struct RadixSortInfo {
  int bits = 0, passes = 0;
  unsigned long long fixed = 0;  // out-of-order groups put right after the prefix passes
  bool fell_back = false;
};

// Sorts the n pairs in (s.keys[0], s.vals[0]) by the whole key, stably, given the smallest and largest key.
// *which = the buffer pair that holds the result.  Synchronises the stream when the keys are wider than the prefix.
inline cudaError_t radix_sort_by_prefix(cudaStream_t stream, SortScratch& s, int64_t n, unsigned long long lo,
                                        unsigned long long hi, int* which, RadixSortInfo* info = nullptr) {
  *which = 0;
  RadixSortInfo local;
  if (!info) info = &local;
  *info = RadixSortInfo();
  if (n <= 1 || hi <= lo) return cudaSuccess;
  int bits = 0;
  for (unsigned long long d = hi - lo; d; d >>= 1) ++bits;
  info->bits = bits;
  static const bool no_prefix = [] {
    const char* e = getenv("TT_RADIX_PREFIX");
    return e && atoi(e) == 0;
  }();
  // Width of the prefix: the fewest whole digits that keep the expected number of keys sharing a prefix with another
  // key (n^2 / 2^(P+1) pairs for uniform keys) below n / 64 — P >= log2(n) + 5: 32 bits for 25 M keys (about 70 000
  // pairs), 24 for 60 000.  Keys that crowd more than that only cost the fall-back's extra passes, never the order.
  int log2n = 0;
  while ((1ll << log2n) < n) ++log2n;
  int prefix = (log2n + 5 + 7) / 8 * 8;
  prefix = prefix < 16 ? 16 : prefix > 40 ? 40 : prefix;
  if (bits <= prefix || no_prefix) {
    info->passes = (bits + 7) / 8;
    return radix_passes(stream, s, n, lo, 0, bits, which);
  }
  const int shift0 = bits - prefix;
  cudaError_t e = radix_passes(stream, s, n, lo, shift0, bits, which);
  if (e != cudaSuccess) return e;
  info->passes = prefix / 8;
  const size_t tiles = (size_t)((s.cap + kRadixThreads * 8 - 1) / (kRadixThreads * 8));
  const uint32_t list_cap = (uint32_t)std::min<size_t>(tiles * kRadixDigits, 1u << 24);
  e = cudaMemsetAsync(s.mm, 0, 16, stream);
  if (e != cudaSuccess) return e;
  radix_fix_find_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(s.keys[*which], n, lo, shift0, s.hist, list_cap, s.mm);
  radix_fix_sort_kernel<<<256, 128, 0, stream>>>(s.keys[*which], s.vals[*which], n, lo, shift0, s.hist, list_cap, s.mm);
  e = cudaGetLastError();
  unsigned long long ctl[2] = {0, 0};
  if (e == cudaSuccess) e = cudaMemcpyAsync(ctl, s.mm, 16, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  if (e != cudaSuccess) return e;
  info->fixed = ctl[0];
  if (ctl[1]) {  // long or too many groups: low digits, then the prefix again (all passes stable)
    info->fell_back = true;
    info->passes += (shift0 + 7) / 8 + prefix / 8;
    e = radix_passes(stream, s, n, lo, 0, shift0, which);
    if (e == cudaSuccess) e = radix_passes(stream, s, n, lo, shift0, bits, which);
  }
  return e;
}

// min / max in s.mm -> host (synchronises the stream)
inline cudaError_t radix_minmax_read(cudaStream_t stream, SortScratch& s, unsigned long long mm[2]) {
  cudaError_t e = cudaMemcpyAsync(mm, s.mm, 16, cudaMemcpyDeviceToHost, stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
  return e;
}

// Sorts the n pairs in (s.keys[0], s.vals[0]) by the whole key, stably (collects min / max first).
inline cudaError_t radix_sort_auto(cudaStream_t stream, SortScratch& s, int64_t n, int* which, RadixSortInfo* info = nullptr) {
  *which = 0;
  if (n <= 1) return cudaSuccess;
  cudaError_t e = radix_minmax_reset(stream, s);
  if (e != cudaSuccess) return e;
  const unsigned grid = (unsigned)((n + 255) / 256 < 1184 ? (n + 255) / 256 : 1184);
  radix_minmax_kernel<<<grid ? grid : 1, 256, 0, stream>>>(s.keys[0], n, s.mm);
  unsigned long long mm[2];
  e = radix_minmax_read(stream, s, mm);
  if (e != cudaSuccess) return e;
  return radix_sort_by_prefix(stream, s, n, mm[0], mm[1], which, info);
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
