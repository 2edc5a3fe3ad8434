// shards.cuh — root-side kernels of the sharded task table.
//
// The table is split by TaskId range over G shards, one per GPU (or several per GPU in the
// tests).  Every shard runs the ordinary tick (sweep.cuh: K1, K3g, tail) and K3g stores the shard's
// result straight into the shard's SLOT of one receive buffer in the root GPU's HBM — over NVLink
// when the shard lives on another GPU: per 8192-row tile its hit count and its offset in the
// shard's row-ordered list, per hit a 13-bit in-tile offset, and optionally the hit's TaskId and
// TaskCreatedOn.  The tail kernel then publishes the tick's epoch in the slot header.  The root
// needs no collective: shard_wait_kernel polls the G epoch words (acquire loads at system scope),
// lays the slots end to end (exclusive scan of the G hit counts), and shard_expand_kernel
// rebuilds the row numbers.  The list `GetYesterdaysDueTasks` returns
// (Api/Services/ITasksManager.cs:14) is that concatenation — shards are TaskId ranges, so
// shard-major row order is TaskId-range order — or, for the store-backed manager, the same hits
// ordered by TaskCreatedOn (Api/Services/TasksStoreManager.cs:136), which is one radix sort of
// the pushed keys on the root.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tt_synth.h"
#include "sweep.cuh"

namespace tt {

constexpr int kMaxShards = 32;  // one lane of the waiting warp per shard

// First 256 bytes of every slot.
struct SlotHeader {
  unsigned long long epoch;     // ticks this shard has published (written last, release at system scope)
  unsigned long long count;     // hits stored in the slot by the current tick
  unsigned long long consumed;  // slot 0 only: epochs the root has finished reading
  unsigned long long pad[29];
};
static_assert(sizeof(SlotHeader) == 256, "slot header is 256 bytes");

// What the root learns from a wait: where each slot's hits start in the concatenated list.
struct CollectMeta {
  unsigned long long base[kMaxShards + 1];  // exclusive scan of the counts; base[world] = total
  unsigned int timed_out;                   // bit r set: shard r did not publish the epoch in time
  unsigned int pad;
  unsigned long long done;                  // host copy only: number of the last collect whose kernels have all run
};

__device__ __forceinline__ unsigned long long ld_acquire_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// One warp; lane r waits until the word at base + r * stride reaches `expected`, or gives up after
// timeout_ns (a shard that never ticks must not hang the GPU).  With meta != null the lanes then
// read their slot's hit count and publish the exclusive scan (device copy for the kernels that
// follow, mapped host copy for the caller).  word_off selects the awaited word inside the header.
__global__ void shard_wait_kernel(const unsigned char* base, size_t stride, int world, size_t word_off,
                                  unsigned long long expected, unsigned long long timeout_ns, CollectMeta* meta_dev,
                                  CollectMeta* meta_host) {
  const int lane = threadIdx.x;
  bool late = false;
  unsigned long long count = 0;
  if (lane < world) {
    const unsigned char* slot = base + (size_t)lane * stride;
    const unsigned long long* word = reinterpret_cast<const unsigned long long*>(slot + word_off);
    const unsigned long long t0 = global_timer_ns();
    while (ld_acquire_sys_u64(word) < expected) {
      if (global_timer_ns() - t0 > timeout_ns) {
        late = true;
        break;
      }
      __nanosleep(200);
    }
    if (meta_dev && !late) count = reinterpret_cast<const SlotHeader*>(slot)->count;
  }
  const unsigned int late_mask = __ballot_sync(0xFFFFFFFFu, late);
  unsigned long long incl = count;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
    if (lane >= d) incl += up;
  }
  if (meta_dev) {
    if (lane < world) {
      meta_dev->base[lane] = incl - count;
      if (meta_host) meta_host->base[lane] = incl - count;
    }
    if (lane == world - 1) {
      meta_dev->base[world] = incl;
      if (meta_host) meta_host->base[world] = incl;
    }
  }
  if (lane == 0) {
    if (meta_dev) meta_dev->timed_out = late_mask;
    if (meta_host) meta_host->timed_out = late_mask;
  }
}

// Byte offsets of a slot's parts (identical for every slot of an exchange).
struct SlotLayout {
  size_t tile_count;  // u32[max_tiles]
  size_t tile_first;  // u32[max_tiles]
  size_t off16;       // u16[cap]
  size_t ids;         // uint4[cap], 0 = not exchanged
  size_t created;     // u64[cap],  0 = not exchanged
  size_t bytes;       // slot stride
};

// Rebuilds shard-local row numbers for every slot at once: grid (tiles / 8, world), one warp per
// tile.  rows_out is the concatenated list (slot r starts at meta->base[r]).
__global__ void __launch_bounds__(256) shard_expand_kernel(const unsigned char* base, SlotLayout lay, uint32_t max_tiles,
                                                            const CollectMeta* meta, uint32_t* rows_out, uint32_t first_slot) {
  const uint32_t r = blockIdx.y + first_slot;  // first_slot = 1: the root's own placement kernel wrote slot 0's rows in place
  const uint32_t t = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (t >= max_tiles) return;
  const unsigned char* slot = base + (size_t)r * lay.bytes;
  const uint32_t cnt = reinterpret_cast<const uint32_t*>(slot + lay.tile_count)[t];
  if (cnt == 0) return;
  const uint32_t first = reinterpret_cast<const uint32_t*>(slot + lay.tile_first)[t];
  const uint16_t* src = reinterpret_cast<const uint16_t*>(slot + lay.off16) + first;
  uint32_t* dst = rows_out + meta->base[r] + first;
  // all of a lane's loads go out before its first store: a tile of the usual density (about 200 hits) is in
  // flight at once instead of one 64-byte piece per warp (measured at 8 GPUs: the dependent form ran at 2.7 TB/s)
  const uint32_t lane = threadIdx.x & 31u, row0 = t * TT_TILE_ROWS;
  uint16_t v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t k = lane + 32u * j;
    v[j] = k < cnt ? src[k] : (uint16_t)0;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t k = lane + 32u * j;
    if (k < cnt) dst[k] = row0 + v[j];
  }
  for (uint32_t k = lane + 256u; k < cnt; k += 32) dst[k] = row0 + src[k];  // denser tiles: the rest
}

// Sort keys of the cross-shard OrderBy(TaskCreatedOn): key = the pushed ticks, value = position in
// the concatenated (shard-major = TaskId-range) list, so the stable sort keeps ties in that order.
__global__ void shard_order_keys_kernel(const unsigned char* base, SlotLayout lay, const CollectMeta* meta,
                                        unsigned long long* keys, uint32_t* vals, int descending) {
  const uint32_t r = blockIdx.y;
  const unsigned long long n = meta->base[r + 1] - meta->base[r];
  const unsigned long long* created = reinterpret_cast<const unsigned long long*>(base + (size_t)r * lay.bytes + lay.created);
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long k = created[i] & TT_TICKS_MASK;
    keys[meta->base[r] + i] = descending ? (~k & TT_TICKS_MASK) : k;
    vals[meta->base[r] + i] = (uint32_t)(meta->base[r] + i);
  }
}

__device__ __forceinline__ int slot_of_position(const CollectMeta* meta, int world, unsigned long long p) {
  int r = 0;
  while (r + 1 < world && meta->base[r + 1] <= p) ++r;
  return r;
}

// The list in its final order: out_rows[j] = global row, out_ids[j] = TaskId of the hit that the
// sort (or, with perm == null, the concatenation) puts at position j.  first_row[r] = global row
// of shard r's row 0.
__global__ void shard_emit_kernel(const unsigned char* base, SlotLayout lay, const CollectMeta* meta, int world,
                                  const unsigned long long* first_row, const uint32_t* rows_cat, const uint32_t* perm,
                                  unsigned long long n, unsigned long long* out_rows, uint4* out_ids) {
  for (unsigned long long j = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; j < n;
       j += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long p = perm ? perm[j] : j;
    const int r = slot_of_position(meta, world, p);
    out_rows[j] = first_row[r] + rows_cat[p];
    if (out_ids && lay.ids)
      out_ids[j] = reinterpret_cast<const uint4*>(base + (size_t)r * lay.bytes + lay.ids)[p - meta->base[r]];
  }
}

// Order-free digest over a list of (global row, id): XOR of the ids as big-endian hi / lo words, sum
// of the rows -> out[0..2] (zeroed before launch).
__global__ void shard_digest_kernel(const unsigned long long* rows, const uint4* ids, unsigned long long n,
                                    unsigned long long* out) {
  unsigned long long xh = 0, xl = 0, sr = 0;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    if (ids) {
      const uint4 id = ids[i];
      xh ^= ((unsigned long long)__byte_perm(id.x, 0, 0x0123) << 32) | __byte_perm(id.y, 0, 0x0123);
      xl ^= ((unsigned long long)__byte_perm(id.z, 0, 0x0123) << 32) | __byte_perm(id.w, 0, 0x0123);
    }
    sr += rows[i];
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    xh ^= __shfl_xor_sync(0xFFFFFFFFu, xh, d);
    xl ^= __shfl_xor_sync(0xFFFFFFFFu, xl, d);
    sr += __shfl_xor_sync(0xFFFFFFFFu, sr, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicXor(out + 0, xh);
    atomicXor(out + 1, xl);
    atomicAdd(out + 2, sr);
  }
}

// An empty shard's tick: nothing to sweep, the slot still has to report "0 hits" for this epoch.
__global__ void shard_publish_empty_kernel(SlotHeader* hdr, unsigned long long epoch) {
  hdr->count = 0;
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&hdr->epoch), "l"(epoch) : "memory");
}

// slot 0's `consumed` word: the root has finished reading epoch `value`
// (and, for a collect that ends here, its number into the root's mapped host word, which the host watches)
__global__ void shard_consumed_kernel(SlotHeader* slot0, unsigned long long value, unsigned long long* host_done = nullptr,
                                      unsigned long long done = 0) {
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&slot0->consumed), "l"(value) : "memory");
  if (host_done) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(host_done), "l"(done) : "memory");
  }
}

// Creator filter across shards: one shard's matches (already ordered by that shard) as sort input on
// the root — keys[i] = complemented ticks (descending order), glob[i] = global row, with the ids copied
// alongside.  Runs on the SHARD's device, writing into a staging area there.
__global__ void shard_stage_matches_kernel(const uint32_t* rows, const uint4* ids, const uint64_t* created_raw, int64_t n,
                                           unsigned long long first_row, unsigned long long* keys,
                                           unsigned long long* glob, uint4* ids_out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = rows[i];
  keys[i] = ~created_raw[r] & TT_TICKS_MASK;
  glob[i] = first_row + r;
  ids_out[i] = ids[i];
}

__global__ void iota_kernel(uint32_t* v, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = (uint32_t)i;
}

__global__ void permute_matches_kernel(const uint32_t* perm, const unsigned long long* glob, const uint4* ids, int64_t n,
                                       unsigned long long* glob_out, uint4* ids_out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t p = perm[i];
  glob_out[i] = glob[p];
  ids_out[i] = ids[p];
}

}  // namespace tt
