// sweep.cuh — K1/K2: fused predicate + stream compaction + overdue flag write, and the
// two small kernels that put the hit list in row order and fetch the TaskIds.
//
// One pass over the hot columns replaces the reference's GET -> filter -> POST round trip
// (Svc/Controllers/ScheduledTasksManagerController.cs:19-46 calling
//  Api/Services/TasksStoreManager.cs:104-149 or Api/Services/FakeTasksManager.cs:107-112).
//
// K1 sweep_kernel.  Tile = 8192 rows per CTA iteration (256 threads, 3 CTAs per SM,
// persistent).  Each warp owns 1024 consecutive rows = 32 flag words: every lane issues 8
// coalesced 128-bit loads of due_code (4 KB in flight per warp), evaluates the predicate to 8
// nibbles, and an 8x8 nibble transpose across each group of 8 lanes (3 shuffles) turns them
// into one 32-row match word per lane, aligned with the completed / overdue bit-plane words.
// Hits are ranked with popc + warp scan + CTA scan; the tile then takes a contiguous slot in
// the hit list with ONE atomicAdd on a global cursor.  Software pipeline: the loads of the
// CTA's next tile are in flight while the current one is ranked, and a tile is written one
// iteration after its slot was requested, so neither DRAM nor atomic latency is exposed.
// No CTA ever waits for another one (measured on B200: an in-order variant with per-tile
// status words spent most of its time polling them; this one streams at ~6.3 TB/s).
//
// The slots are handed out in completion order, so the list K1 writes is ordered inside a
// tile but not across tiles.  K2 tile_offsets_kernel turns the per-tile counts into the
// offsets of the row-ordered list, and K3 place_hits_kernel moves every hit to its final
// position and gathers its 16-byte TaskId on the way (out_ids[i] = task_id[out_rows[i]]).
// Final order = ascending row = the reference's list order, deterministic.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "tt_bits.h"

namespace tt {

// Programmatic dependent launch: the three kernels of a tick are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so the next kernel's launch and its first CTAs' set-up
// overlap the tail of the one before; a dependent kernel reads nothing its predecessor wrote before
// grid_dependency_wait() returns (= the predecessor has completed and its writes are visible).  Both are
// no-ops in a kernel launched the ordinary way.
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_dependent(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t stream, bool programmatic,
                                    Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = programmatic ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}


constexpr int kSweepThreads = 256;
constexpr int kSweepWarps = kSweepThreads / 32;
constexpr int kSweepCtasPerSm = 3;  // measured: 0.648 ms per 1 B rows; 4 (64 registers) 0.680 ms; 2 0.803 ms

struct SweepCtl {
  unsigned int tile_counter;         // next tile to hand out (relative to first_tile)
  unsigned int pad;
  unsigned long long cursor;         // hits so far = next free slot of the unordered list; total at the end
  unsigned long long stored;         // min slot of a tile that did not fit (= hits stored); ~0 if all fit
};

// What a tick reports to the host: written by sweep_tail_kernel into mapped pinned host memory,
// so the call needs no device-to-host copy of its own.
struct SweepOut {
  unsigned long long cursor;
  unsigned long long stored;
  unsigned long long seq;  // the tick's number, written last: a host that reads it has the two words above
};

struct SweepParams {
  const uint32_t* due_code;
  const uint32_t* completed;
  uint32_t* overdue;
  uint32_t* list;                    // unordered hit rows (scratch, cap entries); null = count only
  uint32_t* tile_slot;               // per tile: first slot in `list`
  uint32_t* tile_count;              // per tile: hits stored (0 if the tile did not fit)
  uint32_t* group_sum;               // per kTilesPerGroup tiles: sum of tile_count (zero before launch)
  uint32_t* sub_sum;                 // per kTilesPerSub tiles: the same at a finer grain (zero before launch); may be null
  SweepCtl* ctl;                     // tile_counter = cursor = 0, stored = ~0 before launch
  unsigned long long cap;            // capacity of `list` in hits
  uint32_t first_tile;
  uint32_t num_tiles;                // tiles to process starting at first_tile
  uint32_t mask;                     // TT_MASK_FAKE / TT_MASK_STORE
  uint32_t yesterday;                // day number of DateTime.Today.AddDays(-1)
  uint32_t mark_enable;              // cron predicate runAt.Date > TaskDueDate.Date (uniform, see below)
};

constexpr int kTilesPerGroup = 1024;  // K2: one CTA per group of tiles
constexpr int kTilesPerSub = 32;      // K3g finds a tile's offset from group sums + sub sums + <= 31 tile counts

__device__ __forceinline__ uint4 ld_stream_v4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// Random 16-byte gathers: without a hint the L2 fetches 128 B from DRAM per miss on B200
// (measured 132 B/hit); the 64 B prefetch-size hint halves that (68 B/hit) — there is no
// smaller fetch size.
__device__ __forceinline__ uint4 ld_gather_v4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ uint64_t ld_gather_u64(const uint64_t* p) {
  uint64_t r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::64B.u64 %0, [%1];" : "=l"(r) : "l"(p));
  return r;
}

__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, d);
  return v;
}

// USE_FLAGS: STORE mode's `IsCompleted == false && IsOverDue == false` (TasksStoreManager.cs:135).
// MARK     : set IsOverDue for the listed rows (TasksStoreManager.cs:146) when the cron predicate
//            holds.  For a listed row TaskDueDate.Date is yesterday's date, so the per-row test
//            `runAt.Date > TaskDueDate.Date` (ScheduledTasksManagerController.cs:34) has the
//            same value for every listed row: it is passed in as mark_enable.
//
// Software pipeline: while a tile is ranked, looked back and written, the 8 x 128-bit loads
// and the two flag words of the CTA's NEXT tile are already in flight (it was claimed from the
// atomic counter at the top of the iteration), so every CTA keeps 32 KB outstanding almost all
// the time and the look-back latency is not exposed to DRAM.
struct TileLoads {
  uint4 v[8];
  uint32_t cw, ow;
};

template <bool USE_FLAGS, int THREADS>
__device__ __forceinline__ void issue_tile_loads(const SweepParams& p, uint32_t tile, uint32_t warp, uint32_t lane,
                                                 TileLoads& t) {
  constexpr uint32_t kTileRows = (THREADS / 32) * TT_WARP_ROWS;
  const uint4* src = reinterpret_cast<const uint4*>(p.due_code) +
                     ((size_t)tile * (kTileRows / 4) + (size_t)warp * (TT_WARP_ROWS / 4) + lane);
#pragma unroll
  for (int j = 0; j < 8; ++j) t.v[j] = ld_stream_v4(src + j * 32);
  t.cw = 0;
  t.ow = 0;
  if (USE_FLAGS) {
    const uint32_t wi = tile * (kTileRows / 32) + warp * 32u + lane;
    t.cw = __ldg(p.completed + wi);
    t.ow = p.overdue[wi];
  }
}

// What a thread remembers about the tile of the previous iteration until its slot is known.
struct PendingTile {
  uint32_t rel;     // tile number relative to first_tile; ~0u = none
  uint32_t hits;    // this lane's 32-row hit word
  uint32_t ow;      // the overdue word it was computed against
  uint32_t offset;  // rank of this lane's first hit inside the tile
  uint32_t total;   // hits in the whole tile
};

template <bool MARK, int THREADS>
__device__ __forceinline__ void emit_tile(const SweepParams& p, const PendingTile& t, uint32_t slot, uint32_t warp,
                                          uint32_t lane, uint32_t tid) {
  constexpr uint32_t kTileRows = (THREADS / 32) * TT_WARP_ROWS;
  const uint32_t tile = p.first_tile + t.rel;
  const bool fits = (unsigned long long)slot + t.total <= p.cap;
  if (fits) {
    if (p.list) {
      uint32_t* dst = p.list + slot + t.offset;
      const uint32_t row0 = tile * kTileRows + warp * TT_WARP_ROWS + lane * 32u;
      uint32_t m = t.hits;
      while (m) {
        uint32_t b = __ffs(m) - 1;
        m &= m - 1;
        *dst++ = row0 + b;
      }
    }
    if (MARK && p.mark_enable && t.hits) p.overdue[tile * (kTileRows / 32) + warp * 32u + lane] = t.ow | t.hits;
  }
  if (tid == 0) {
    if (p.list) {
      p.tile_slot[t.rel] = slot;
      p.tile_count[t.rel] = fits ? t.total : 0u;
      if (fits && t.total) {
        atomicAdd(p.group_sum + t.rel / kTilesPerGroup, t.total);
        if (p.sub_sum) atomicAdd(p.sub_sum + t.rel / kTilesPerSub, t.total);
      }
    }
    if (!fits) atomicMin(&p.ctl->stored, (unsigned long long)slot);
  }
}

// RANGE    : the generalised predicate of SURVEY.md §8 f4 — "due on or before yesterday" instead of
//            "due exactly yesterday at 00:00:00": every live row whose due DAY is <= yesterday
//            matches, whatever its time of day or Kind (padding rows and tombstones carry bits
//            above bit 24 and never match).  No reference counterpart: it covers the cases the
//            tutorial says its exact match leaves out (docs/aca/07-aca-cron-bindings/index.md:106).
template <bool USE_FLAGS, bool MARK, int THREADS, int CTAS_PER_SM, bool RANGE = false>
__global__ void __launch_bounds__(THREADS, CTAS_PER_SM) sweep_kernel(SweepParams p) {
  grid_launch_dependents();  // the placement kernel's first CTAs may take their seats now (they wait for this grid's end)
  constexpr int kWarps = THREADS / 32;
  __shared__ uint32_t s_wtot[kWarps];
  __shared__ uint32_t s_tile;
  __shared__ uint32_t s_slot;

  const uint32_t tid = threadIdx.x;
  const uint32_t lane = tid & 31u;
  const uint32_t warp = tid >> 5;
  const uint32_t src_lane = tt_word_src_lane(lane);

  if (tid == 0) s_tile = atomicAdd(&p.ctl->tile_counter, 1u);
  __syncthreads();
  uint32_t cur = s_tile;
  __syncthreads();  // s_tile is read before iteration 0 overwrites it

  PendingTile pend;
  pend.rel = ~0u;
  pend.hits = pend.ow = pend.offset = pend.total = 0;
  uint32_t pend_slot = 0;  // thread 0: the slot the cursor atomic returned for the pending tile

  TileLoads ld;
  if (cur < p.num_tiles) issue_tile_loads<USE_FLAGS, THREADS>(p, p.first_tile + cur, warp, lane, ld);

  while (cur < p.num_tiles) {
    const uint32_t rel = cur;
    // claim the next tile now; its number is needed only after the barrier
    uint32_t claimed = 0;
    if (tid == 0) claimed = atomicAdd(&p.ctl->tile_counter, 1u);

    // ---- predicate -> nibbles -> 32-row match word (waits for this tile's loads) ----
    uint32_t packed = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const uint4 v = ld.v[j];
      uint32_t nib;
      if (RANGE) {
        // valid codes are < 2^25; (code & day mask) <= yesterday  <=>  the due date's day has passed
        auto due = [&](uint32_t c) { return c < (1u << 25) && (c & TT_DUE_DAY_MASK) <= p.yesterday; };
        nib = (due(v.x) ? 1u : 0u) | (due(v.y) ? 2u : 0u) | (due(v.z) ? 4u : 0u) | (due(v.w) ? 8u : 0u);
      } else {
        nib = ((v.x & p.mask) == p.yesterday ? 1u : 0u) | ((v.y & p.mask) == p.yesterday ? 2u : 0u) |
              ((v.z & p.mask) == p.yesterday ? 4u : 0u) | ((v.w & p.mask) == p.yesterday ? 8u : 0u);
      }
      packed |= nib << (4 * j);
    }
    const uint32_t cw = ld.cw, ow = ld.ow;
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
    if (tid == 0) {
      s_tile = claimed;
      s_slot = pend_slot;  // the atomic issued one iteration ago has long returned
    }
    __syncthreads();

    // ---- next tile's loads go out now; they land while the previous tile is written ----
    const uint32_t nxt = s_tile;
    if (nxt < p.num_tiles) issue_tile_loads<USE_FLAGS, THREADS>(p, p.first_tile + nxt, warp, lane, ld);

    uint32_t wbase = 0, total = 0;
#pragma unroll
    for (int k = 0; k < kWarps; ++k) {
      uint32_t x = s_wtot[k];
      wbase += (k < (int)warp) ? x : 0u;
      total += x;
    }
    // ---- ask for this tile's slot (used next iteration), write the previous tile ----
    if (tid == 0) pend_slot = (uint32_t)atomicAdd(&p.ctl->cursor, (unsigned long long)total);
    if (pend.rel != ~0u) emit_tile<MARK, THREADS>(p, pend, s_slot, warp, lane, tid);
    pend.rel = rel;
    pend.hits = hits;
    pend.ow = ow;
    pend.offset = wbase + incl - cnt;
    pend.total = total;
    cur = nxt;
    __syncthreads();  // s_wtot / s_tile / s_slot are rewritten by the next iteration
  }

  // ---- drain: the last tile this CTA processed ----
  if (pend.rel != ~0u) {
    if (tid == 0) s_slot = pend_slot;
    __syncthreads();
    emit_tile<MARK, THREADS>(p, pend, s_slot, warp, lane, tid);
  }
}

// K2: offsets of the row-ordered list.  tile_first[t] = number of stored hits in tiles < t.
// One CTA per group of kTilesPerGroup tiles: its carry is the sum of the earlier groups'
// totals (accumulated by K1), then a block-wide exclusive scan of the group's counts.
__global__ void __launch_bounds__(kTilesPerGroup) tile_offsets_kernel(const uint32_t* tile_count,
                                                                      const uint32_t* group_sum, uint32_t num_tiles,
                                                                      uint32_t* tile_first) {
  __shared__ uint32_t s_warp[kTilesPerGroup / 32];
  __shared__ uint32_t s_carry;
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t g = blockIdx.x;
  if (warp == 0) {
    uint32_t c = 0;
    for (uint32_t k = lane; k < g; k += 32) c += group_sum[k];
    c = warp_sum(c);
    if (lane == 0) s_carry = c;
  }
  const uint32_t t = g * kTilesPerGroup + tid;
  const uint32_t cnt = t < num_tiles ? tile_count[t] : 0u;
  uint32_t incl = cnt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
    if (lane >= (uint32_t)d) incl += up;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  uint32_t wbase = 0;
  for (uint32_t k = 0; k < warp; ++k) wbase += s_warp[k];
  if (t < num_tiles) tile_first[t] = s_carry + wbase + incl - cnt;
}

// K3: every stored hit goes to its place in the row-ordered list; its TaskId is gathered on the
// way (16 B random read, coalesced writes).  The hit count is read from the sweep's control
// block on the device, so K1 -> K2 -> K3 are queued back to back without a host round trip.
constexpr int kPlaceThreads = 256;
constexpr int kPlaceUnroll = 4;

template <int TILE_ROWS>
__global__ void __launch_bounds__(kPlaceThreads) place_hits_kernel(const SweepCtl* ctl, const uint32_t* list,
                                                                   const uint32_t* tile_slot,
                                                                   const uint32_t* tile_first, uint32_t first_tile,
                                                                   const uint4* task_id, uint32_t* out_rows,
                                                                   uint4* out_ids) {
  const unsigned long long st = ctl->stored;
  const unsigned long long n = st == ~0ull ? ctl->cursor : st;
  const unsigned long long stride = (unsigned long long)gridDim.x * kPlaceThreads;
  for (unsigned long long i0 = (unsigned long long)blockIdx.x * kPlaceThreads + threadIdx.x; i0 < n;
       i0 += stride * kPlaceUnroll) {
    uint32_t r[kPlaceUnroll];
    uint4 id[kPlaceUnroll];
#pragma unroll
    for (int j = 0; j < kPlaceUnroll; ++j) {
      unsigned long long i = i0 + j * stride;
      r[j] = i < n ? list[i] : 0u;
    }
    if (out_ids) {
#pragma unroll
      for (int j = 0; j < kPlaceUnroll; ++j) id[j] = ld_gather_v4(task_id + r[j]);
    }
#pragma unroll
    for (int j = 0; j < kPlaceUnroll; ++j) {
      unsigned long long i = i0 + j * stride;
      if (i < n) {
        const uint32_t t = r[j] / TILE_ROWS - first_tile;
        const uint32_t dst = tile_first[t] + ((uint32_t)i - tile_slot[t]);
        out_rows[dst] = r[j];
        if (out_ids) out_ids[dst] = id[j];
      }
    }
  }
}

// K3g: the same placement, driven by the tiles in row order instead of by the hits in list order,
// and fused with the cross-shard exchange.  One CTA owns GT consecutive tiles
// (GT = 1 and one hit per thread is what ships: measured fastest, see sweep_run): their hits are
// consecutive in the row-ordered list (from tile_first[first tile]), so hit h of the group is
// found with GT - 1 compares against the group's prefix counts, read from the tile's slot of the
// unordered list (coalesced), its TaskId gathered, and both written at base + h (coalesced).  While the CTA waits for the 16-byte gathers it also builds the COMPACT form of the
// hits — a 13-bit in-tile offset per hit plus one count per tile — in shared memory and stores it
// with 16-byte vectors into `push`, which may be this GPU's memory or the root GPU's receive
// buffer mapped over NVLink (tt_peer_open): the exchange then costs no pass of its own and no
// collective, only a barrier after the kernel.  Remote stores are fire-and-forget; the vector
// stores keep the NVLink packets at full size (2-byte scalar stores would send 64 B per warp).
constexpr int kGroupTiles = 1;    // defaults of the template; sweep_run picks the shape
constexpr int kGroupThreads = 256;
constexpr int kGroupUnroll = 1;
constexpr int kGroupStage = 4096;  // offsets staged per group; denser groups store theirs directly

struct PushTarget {
  uint16_t* off16;              // [cap] in-tile offsets of the row-ordered hits; 16-byte aligned; null = no push
  uint32_t* tile_count;         // [num_tiles] hits per tile
  unsigned long long* count;    // [1] hits stored
  uint4* ids;                   // [cap] TaskIds of the hits in row order, or null (the sharded GetYesterdaysDueTasks)
  unsigned long long* created;  // [cap] TaskCreatedOn (.NET dateData) of the hits, or null (the cross-shard OrderBy)
  uint32_t* tile_first;         // [num_tiles] offset of each tile's hits in the row-ordered list, or null
  int created_as_key;           // 0: `created` receives the raw dateData; 1: its ticks (the sort key of OrderBy(TaskCreatedOn))
  int shape;                    // tiles per placement CTA for this push (1, 2, 4, 8, 16, 32); 0 = the library's default
};

// What a sharded tick adds around K3g (shards.cuh): before it, wait until the root has finished
// reading this shard's slot (wait_word >= wait_value); after it, publish `epoch` in the slot.
struct ShardHook {
  unsigned long long* epoch_word;
  unsigned long long epoch;
  const unsigned char* wait_slot0;  // null = no wait (the root's own shard)
  unsigned long long wait_value;
  unsigned long long timeout_ns;
};

// Where the row-ordered list of the CTA's first tile starts: hits stored in all earlier tiles =
// whole groups (group_sum) + whole sub-groups of this group (sub_sum) + the <= 31 tiles of this
// sub-group before t0 — about ten independent loads per lane of one warp and a warp sum, issued
// together with the loads of the tile's own slot and count.  (Until round 2 a separate scan
// kernel, K2 tile_offsets_kernel, produced these offsets between K1 and K3g.)
__device__ __forceinline__ uint32_t tile_list_base(const uint32_t* group_sum, const uint32_t* sub_sum,
                                                   const uint32_t* tile_count, uint32_t t0, uint32_t lane) {
  const uint32_t g = t0 / kTilesPerGroup, s = t0 / kTilesPerSub;
  uint32_t acc = 0;
  for (uint32_t k = lane; k < g; k += 32) acc += group_sum[k];
  const uint32_t s0 = g * (kTilesPerGroup / kTilesPerSub);
  if (s0 + lane < s) acc += sub_sum[s0 + lane];
  const uint32_t f = s * kTilesPerSub;
  if (f + lane < t0) acc += tile_count[f + lane];
  return warp_sum(acc);
}

template <int TILE_ROWS, int GT = kGroupTiles, int UNROLL = kGroupUnroll>
__global__ void __launch_bounds__(kGroupThreads) place_group_kernel(const uint32_t* list, const uint32_t* tile_slot,
                                                                    const uint32_t* tile_count,
                                                                    const uint32_t* group_sum, const uint32_t* sub_sum,
                                                                    uint32_t num_tiles, const uint4* task_id,
                                                                    const uint64_t* created_raw, uint32_t* out_rows,
                                                                    uint4* out_ids, PushTarget push) {
  static_assert(kTilesPerSub % GT == 0 || GT % kTilesPerSub == 0, "a CTA's tiles must not straddle the prefix levels unevenly");
  grid_launch_dependents();
  grid_dependency_wait();  // everything K1 wrote is visible from here on
  __shared__ uint32_t s_slot[GT];
  __shared__ uint32_t s_pre[GT + 1];
  __shared__ uint32_t s_base;
  constexpr int kStage = GT * 512 > kGroupStage ? GT * 512 : kGroupStage;  // about 2.5 x the usual hits of GT tiles
  __shared__ __align__(16) uint16_t s_off[kStage + 8];
  const uint32_t tid = threadIdx.x;
  const uint32_t t0 = blockIdx.x * GT;
  if (tid < 32) {
    const uint32_t t = t0 + tid;
    const bool live = tid < GT && t < num_tiles;
    const uint32_t c = live ? tile_count[t] : 0u;
    const uint32_t slot = live ? tile_slot[t] : 0u;
    const uint32_t base = tile_list_base(group_sum, sub_sum, tile_count, t0, tid);
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < GT; d <<= 1) {
      uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (tid >= (uint32_t)d) incl += up;
    }
    if (tid < GT) {
      s_slot[tid] = slot;
      s_pre[tid + 1] = incl;
      if (live && push.tile_count) push.tile_count[t] = c;
      if (live && push.tile_first) push.tile_first[t] = base + incl - c;
    }
    if (tid == 0) {
      s_pre[0] = 0;
      s_base = base;
    }
  }
  __syncthreads();
  const uint32_t total = s_pre[GT];
  if (total == 0) return;
  const uint32_t base = s_base;
  const uint32_t shift = base & 7u;  // staged index = list index - (base - shift): 16-byte chunks line up
  const bool staged = push.off16 != nullptr && total <= (uint32_t)kStage;

  for (uint32_t h0 = tid; h0 < total; h0 += kGroupThreads * UNROLL) {
    uint32_t r[UNROLL];
    uint4 id[UNROLL];
    unsigned long long cr[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t h = h0 + j * kGroupThreads;
      r[j] = 0;
      if (h < total) {
        uint32_t k = 0;
#pragma unroll
        for (int q = 1; q < GT; ++q) k += h >= s_pre[q] ? 1u : 0u;
        r[j] = list[s_slot[k] + (h - s_pre[k])];
      }
    }
    if (out_ids || push.ids) {
#pragma unroll
      for (int j = 0; j < UNROLL; ++j) id[j] = ld_gather_v4(task_id + r[j]);
    }
    if (push.created) {
#pragma unroll
      for (int j = 0; j < UNROLL; ++j) {
        cr[j] = ld_gather_u64(created_raw + r[j]);
        if (push.created_as_key) cr[j] &= TT_TICKS_MASK;
      }
    }
#pragma unroll
    for (int j = 0; j < UNROLL; ++j) {
      const uint32_t h = h0 + j * kGroupThreads;
      if (h < total) {
        const uint16_t off = (uint16_t)(r[j] & (TILE_ROWS - 1));
        if (staged)
          s_off[shift + h] = off;
        else if (push.off16)
          push.off16[base + h] = off;
        if (out_rows) out_rows[base + h] = r[j];
        if (out_ids) out_ids[base + h] = id[j];
        if (push.ids) push.ids[base + h] = id[j];
        if (push.created) push.created[base + h] = cr[j];
      }
    }
  }
  if (!staged) return;
  __syncthreads();
  uint16_t* dst = push.off16 + (base - shift);
  const uint32_t lo = shift, hi = shift + total;  // staged entries [lo, hi) are valid
  for (uint32_t v = tid; v * 8 < hi; v += kGroupThreads) {
    const uint32_t a = v * 8;
    if (a >= lo && a + 8 <= hi) {
      *reinterpret_cast<uint4*>(dst + a) = *reinterpret_cast<const uint4*>(s_off + a);
    } else {
      for (uint32_t e = a > lo ? a : lo; e < a + 8 && e < hi; ++e) dst[e] = s_off[e];
    }
  }
}

// The tick's epilogue, one small CTA after the last kernel: reports the outcome (hit count,
// stored count) into mapped pinned host memory and into the push target, puts the control block
// and the partial sums back into their start state for the next tick (so a tick issues no
// memset and no copy of its own), and — for a sharded table — publishes the tick's epoch in the
// root GPU's slot: the kernel boundary has made this shard's stores into the root's memory
// complete, the system-scope fence orders the epoch word after them, and the root waits on the
// word (shard_wait_kernel) instead of on a collective.
__global__ void sweep_tail_kernel(SweepCtl* ctl, SweepOut* out, uint32_t* group_sum, uint32_t n_groups, uint32_t* sub_sum,
                                  uint32_t n_subs, unsigned long long* push_count, unsigned long long* epoch_word,
                                  unsigned long long epoch, unsigned long long seq) {
  grid_dependency_wait();  // the placement kernel has finished with the partial sums this kernel clears
  const unsigned long long cursor = ctl->cursor, stored = ctl->stored;
  for (uint32_t k = threadIdx.x; k < n_groups; k += blockDim.x) group_sum[k] = 0;
  if (sub_sum)
    for (uint32_t k = threadIdx.x; k < n_subs; k += blockDim.x) sub_sum[k] = 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    ctl->tile_counter = 0;
    ctl->pad = 0;
    ctl->cursor = 0;
    ctl->stored = ~0ull;
    if (out) {
      out->cursor = cursor;
      out->stored = stored;
    }
    if (push_count) *push_count = stored == ~0ull ? cursor : stored;
    if (epoch_word) {
      __threadfence_system();
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(epoch_word), "l"(epoch) : "memory");
    }
    if (out) {
      __threadfence_system();
      asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(&out->seq), "l"(seq) : "memory");
    }
  }
}

// ---- compact form of a row-ordered hit list (for the cross-shard exchange) --------------------
// A hit is its tile (implied by per-tile counts) and a 13-bit offset inside the tile: 2 bytes per
// hit instead of 4, plus 4 bytes per tile.  pack: off16[i] = rows[i] % tile; the per-tile counts
// are the ones K1 recorded.  expand: one warp per tile rebuilds the row numbers.
__global__ void pack_hits16_kernel(const uint32_t* rows, int64_t n, uint16_t* off16) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) off16[i] = (uint16_t)(rows[i] & (TT_TILE_ROWS - 1));
}

__global__ void expand_hits16_kernel(const uint32_t* tile_count, const uint32_t* tile_first, uint32_t num_tiles,
                                     const uint16_t* off16, uint32_t* rows_out) {
  const uint32_t t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (t >= num_tiles) return;
  const uint32_t first = tile_first[t], cnt = tile_count[t];
  for (uint32_t k = threadIdx.x & 31u; k < cnt; k += 32) rows_out[first + k] = t * TT_TILE_ROWS + off16[first + k];
}

}  // namespace tt
