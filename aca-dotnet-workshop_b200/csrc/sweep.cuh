// sweep.cuh — K1/K2: fused predicate + ordered stream compaction + overdue flag write.
//
// One pass over the hot columns replaces the reference's GET -> filter -> POST round trip
// (Svc/Controllers/ScheduledTasksManagerController.cs:19-46 calling
//  Api/Services/TasksStoreManager.cs:104-149 or Api/Services/FakeTasksManager.cs:107-112).
//
// Tile = 8192 rows per CTA (256 threads).  Each warp owns 1024 consecutive rows = 32 flag
// words: every lane issues 8 coalesced 128-bit loads of due_code up front (4 KB in flight
// per warp), evaluates the predicate to 8 nibbles, and an 8x8 nibble transpose across each
// group of 8 lanes (3 shuffles) turns them into one 32-row match word per lane, aligned with
// the completed / overdue bit-plane words.  Hits are ranked with popc + warp scan + CTA scan,
// the tile's global offset comes from a decoupled look-back over per-tile status words (tiles
// are handed out by an atomic counter so a tile's predecessors are always running), and the
// hit rows are staged in shared memory so the row-index store and the 16-byte TaskId gather
// are coalesced on the output side.  Output order = ascending row = the reference's list order.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "tt_bits.h"

namespace tt {

constexpr int kSweepThreads = 256;
constexpr int kSweepWarps = kSweepThreads / 32;
constexpr int kSweepCtasPerSm = 4;

// tile status word: state in the top 2 bits, value (hit count) in the low 32
constexpr unsigned long long kStInvalid = 0ull;
constexpr unsigned long long kStAggregate = 1ull << 62;
constexpr unsigned long long kStPrefix = 2ull << 62;
constexpr unsigned long long kStValueMask = 0xFFFFFFFFull;

struct SweepCtl {
  unsigned int tile_counter;         // next tile to hand out (relative to first_tile)
  unsigned int pad;
  unsigned long long total;          // hits in all tiles (inclusive prefix of the last tile)
  unsigned long long overflow;       // min over tiles that did not fit of (tile << 32 | base); ~0 if none
};

struct SweepParams {
  const uint32_t* due_code;
  const uint32_t* completed;
  uint32_t* overdue;
  const uint4* task_id;
  uint32_t* out_rows;                // may be null (count only)
  uint4* out_ids;                    // may be null
  unsigned long long* tile_status;   // one word per tile, zeroed before launch
  SweepCtl* ctl;                     // zeroed (overflow = ~0) before launch
  unsigned long long cap;            // capacity of out_rows / out_ids in hits
  uint32_t first_tile;
  uint32_t num_tiles;                // tiles to process starting at first_tile
  uint32_t mask;                     // TT_MASK_FAKE / TT_MASK_STORE
  uint32_t yesterday;                // day number of DateTime.Today.AddDays(-1)
  uint32_t mark_enable;              // cron predicate runAt.Date > TaskDueDate.Date (uniform, see below)
};

__device__ __forceinline__ uint4 ld_stream_v4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ unsigned long long ld_status(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void st_status(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
  return v;
}

// Decoupled look-back, run by one full warp.  Publishes this tile's aggregate, walks back over
// predecessors 32 at a time until one with a final prefix is found, publishes this tile's
// inclusive prefix and returns the exclusive one.  A status word carries state and value in
// one 64-bit store, so no fence is needed.
__device__ __forceinline__ uint32_t tile_lookback(unsigned long long* status, uint32_t rel_tile, uint32_t aggregate,
                                                  uint32_t lane) {
  if (rel_tile == 0) {
    if (lane == 0) st_status(status, kStPrefix | aggregate);
    return 0;
  }
  if (lane == 0) st_status(status + rel_tile, kStAggregate | aggregate);
  uint32_t excl = 0;
  int look = (int)rel_tile - 1;
  while (true) {
    int idx = look - (int)lane;
    unsigned long long s = kStPrefix;  // virtual predecessor of tile 0: prefix 0
    if (idx >= 0) {
      do {
        s = ld_status(status + idx);
      } while ((s >> 62) == 0);
    }
    __syncwarp();
    uint32_t is_prefix = __ballot_sync(0xFFFFFFFFu, (s >> 62) == 2);
    uint32_t val = (uint32_t)(s & kStValueMask);
    if (is_prefix) {
      uint32_t first = __ffs(is_prefix) - 1;  // nearest predecessor with a final prefix
      excl += warp_sum(lane <= first ? val : 0u);
      break;
    }
    excl += warp_sum(val);
    look -= 32;
  }
  if (lane == 0) st_status(status + rel_tile, kStPrefix | (unsigned long long)(excl + aggregate));
  return excl;
}

// USE_FLAGS: STORE mode's `IsCompleted == false && IsOverDue == false` (TasksStoreManager.cs:135).
// MARK     : set IsOverDue for the listed rows (TasksStoreManager.cs:146) when the cron predicate
//            holds.  For a listed row TaskDueDate.Date is yesterday's date, so the per-row test
//            `runAt.Date > TaskDueDate.Date` (ScheduledTasksManagerController.cs:34) has the
//            same value for every listed row: it is passed in as mark_enable.
template <bool USE_FLAGS, bool MARK>
__global__ void __launch_bounds__(kSweepThreads, kSweepCtasPerSm) sweep_kernel(SweepParams p) {
  __shared__ uint32_t s_list[TT_TILE_ROWS];
  __shared__ uint32_t s_wtot[kSweepWarps];
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_base;

  const uint32_t tid = threadIdx.x;
  const uint32_t lane = tid & 31u;
  const uint32_t warp = tid >> 5;
  const uint32_t src_lane = tt_word_src_lane(lane);

  for (;;) {
    if (tid == 0) s_tile = atomicAdd(&p.ctl->tile_counter, 1u);
    __syncthreads();
    const uint32_t rel = s_tile;
    if (rel >= p.num_tiles) break;
    const uint32_t tile = p.first_tile + rel;

    // ---- loads: 8 x 128-bit per lane, all issued before first use ----
    const uint4* src = reinterpret_cast<const uint4*>(p.due_code) +
                       ((size_t)tile * (TT_TILE_ROWS / 4) + (size_t)warp * (TT_WARP_ROWS / 4) + lane);
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = ld_stream_v4(src + j * 32);
    const uint32_t wi = tile * TT_TILE_WORDS + warp * 32u + lane;  // this lane's flag word
    uint32_t cw = 0, ow = 0;
    if (USE_FLAGS) {
      cw = __ldg(p.completed + wi);
      ow = p.overdue[wi];
    }

    // ---- predicate -> nibbles -> 32-row match word ----
    uint32_t packed = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      uint32_t nib = ((v[j].x & p.mask) == p.yesterday ? 1u : 0u) | ((v[j].y & p.mask) == p.yesterday ? 2u : 0u) |
                     ((v[j].z & p.mask) == p.yesterday ? 4u : 0u) | ((v[j].w & p.mask) == p.yesterday ? 8u : 0u);
      packed |= nib << (4 * j);
    }
#pragma unroll
    for (uint32_t d = 4; d > 0; d >>= 1) {
      uint32_t other = __shfl_xor_sync(0xFFFFFFFFu, packed, d);
      packed = tt_nibble_xpose_step(packed, other, lane, d);
    }
    const uint32_t match = __shfl_sync(0xFFFFFFFFu, packed, src_lane);
    const uint32_t hits = USE_FLAGS ? (match & ~(cw | ow)) : match;

    // ---- rank: lane -> warp -> CTA ----
    const uint32_t cnt = __popc(hits);
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= (uint32_t)d) incl += up;
    }
    if (lane == 31) s_wtot[warp] = incl;
    __syncthreads();
    uint32_t wbase = 0, total = 0;
#pragma unroll
    for (int k = 0; k < kSweepWarps; ++k) {
      uint32_t x = s_wtot[k];
      wbase += (k < (int)warp) ? x : 0u;
      total += x;
    }

    // ---- stage hit rows in shared memory (ascending row order) ----
    {
      uint32_t pos = wbase + incl - cnt;
      const uint32_t row0 = tile * TT_TILE_ROWS + warp * TT_WARP_ROWS + lane * 32u;
      uint32_t m = hits;
      while (m) {
        uint32_t b = __ffs(m) - 1;
        m &= m - 1;
        s_list[pos++] = row0 + b;
      }
    }
    if (warp == 0) {
      uint32_t excl = tile_lookback(p.tile_status, rel, total, lane);
      if (lane == 0) s_base = excl;
    }
    __syncthreads();
    const uint32_t base = s_base;

    // ---- emit ----
    if ((unsigned long long)base + total <= p.cap) {
      if (MARK && p.mark_enable && hits) p.overdue[wi] = ow | hits;
      if (p.out_rows) {
        for (uint32_t i = tid; i < total; i += kSweepThreads) {
          uint32_t r = s_list[i];
          p.out_rows[base + i] = r;
          if (p.out_ids) p.out_ids[base + i] = ld_stream_v4(p.task_id + r);
        }
      }
    } else if (tid == 0) {
      atomicMin(&p.ctl->overflow, ((unsigned long long)tile << 32) | base);
    }
    if (rel == p.num_tiles - 1 && tid == 0) p.ctl->total = (unsigned long long)base + total;
  }
}

}  // namespace tt
