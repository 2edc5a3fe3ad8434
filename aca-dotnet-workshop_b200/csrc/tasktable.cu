// tasktable.cu — the C ABI of include/tasktable.h over the sm_100a kernels.
//
// Host side: table lifetime, synthetic fill / SoA load, result objects, the resumable sweep
// driver.  No CPU implementation of any ITasksManager operation lives here: if CUDA is not
// usable every call fails with TT_ERR_CUDA.

#include <cuda_runtime.h>
#include <unistd.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <chrono>
#include <memory>
#include <condition_variable>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/tasktable.h"
#include "creator.cuh"
#include "ingest.cuh"
#include "json.cuh"
#include "order.cuh"
#include "point.cuh"
#include "shards.cuh"
#include "sweep.cuh"
#include "tt_bits.h"

namespace {

thread_local std::string g_err = "";

int32_t fail(int32_t code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define TT_CUDA(expr)                                                                              \
  do {                                                                                             \
    cudaError_t e_ = (expr);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      cudaGetLastError();                                                                          \
      return fail(e_ == cudaErrorMemoryAllocation ? TT_ERR_OOM : TT_ERR_CUDA, "%s: %s (%s:%d)", #expr, \
                  cudaGetErrorString(e_), __FILE__, __LINE__);                                     \
    }                                                                                              \
  } while (0)

constexpr int64_t kMaxRows = 1ll << 31;

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

}  // namespace

// Shared between a table and the results it handed out, so a result outliving its table can
// still be freed (and refuses everything else).
struct TableCore {
  int device = 0;
  std::atomic<bool> alive{true};
  // one spare result allocation: the block of the last freed result is kept for the next one, so a
  // steady stream of queries does not pay cudaMalloc / cudaFree (both synchronise the device)
  std::mutex pool_mu;
  void* spare = nullptr;
  size_t spare_bytes = 0;
  // the same for the text of a JSON emission (tt_result_json): sizing and emitting are two kernels with the allocation
  // of the text between them, and cudaMalloc of gigabytes took longer than either kernel's launch gap
  void* json_spare = nullptr;
  size_t json_spare_bytes = 0;
};

struct tt_table {
  std::shared_ptr<TableCore> core = std::make_shared<TableCore>();
  int device = 0;
  std::mutex mu;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
  int sm_count = 148;

  int64_t n = 0, n_pad = 0;
  // hot columns
  uint32_t* due_code = nullptr;
  uint32_t* completed = nullptr;
  uint32_t* overdue = nullptr;
  uint64_t* creator_hash = nullptr;
  // cold columns
  uint4* task_id = nullptr;
  uint64_t* created_raw = nullptr;
  uint64_t* due_raw = nullptr;
  uint32_t* creator_ref = nullptr;
  // creator dictionary
  uint32_t n_creators = 0;
  uint32_t* creator_off = nullptr;
  uint8_t* creator_pool = nullptr;
  uint64_t pool_bytes = 0;
  bool ids_sorted = false;
  uint32_t* id_order = nullptr;  // rows sorted by TaskId when the table itself is not
  // point operations: live-row count, capacity of the creator dictionary on the device and its
  // host mirror (built on the first insert)
  int64_t n_live = 0;
  size_t creator_off_cap = 0, creator_pool_cap = 0;
  bool dict_mirrored = false;
  std::vector<uint32_t> h_creator_off;
  std::vector<uint8_t> h_creator_pool;
  std::unordered_map<std::string, uint32_t> h_creator_index;
  uint8_t last_id[16] = {};
  bool last_id_known = false;

  // sweep scratch: per-tile slot / count / final offset (3 words per tile), group sums, the
  // unordered hit list (grown on demand), and the control block with its pinned mirror
  uint32_t* tile_words = nullptr;   // [3][n_tiles]
  uint32_t* group_sum = nullptr;    // [n_groups] then [n_subs]: K1's partial sums of the per-tile counts
  int64_t n_tiles = 0, n_groups = 0, n_subs = 0;
  bool ctl_clean = false;           // control block and partial sums are in their start state (the last tick's tail kernel ran)
  tt::SweepOut* h_out = nullptr;    // the tick's outcome: mapped pinned host memory the tail kernel writes
  tt::SweepOut* d_out = nullptr;    // its device address
  uint64_t tick_seq = 0;            // ticks launched so far (the tail kernel reports the number back)
  struct PendingSweep {             // a tick queued with TT_SWEEP_ASYNC and not yet collected by tt_sweep_wait
    bool active = false, listing = false, timing = false, trivial = false;  // trivial: nothing was launched
    uint32_t mark_enable = 0;
    int64_t first_row = 0, row_end = -1, cap = 0;
    int32_t launches = 0;
    cudaStream_t stream = nullptr;
    uint64_t seq = 0;  // the tick's number (0: nothing to watch for)
  } pending;
  uint32_t* hit_list = nullptr;
  int64_t hit_list_cap = 0;
  tt::SweepCtl* ctl = nullptr;
  tt::SweepCtl* h_ctl = nullptr;
  // small device/pinned scratch for reductions and the creator filter
  unsigned long long* d_small = nullptr;  // 16 words
  unsigned long long* h_small = nullptr;
  uint8_t* d_query = nullptr;  // creator filter: the query set (hashes, offsets, strings, slot table)
  size_t d_query_cap = 0;
  unsigned long long* d_pairs = nullptr;  // creator filter: unordered (query, row) matches (large results)
  int64_t d_pairs_cap = 0;
  uint32_t* d_lists = nullptr;            // creator filter: per-query match lists [queries][kListCap]
  int64_t d_lists_queries = 0;
  uint32_t* d_qcount = nullptr;  // per-query match counts (kMaxQueries words) + pinned mirror
  uint32_t* h_qcount = nullptr;
  tt_op_stats last_op = {};
  // Request combining for GetTasksByCreator: request threads that arrive while a table pass is running queue
  // their query; whoever finds no pass running takes everything queued (up to kMaxQueries) as ONE batch —
  // one read of the hash column answers all of them — and hands each waiter its own result.
  struct CreatorRequest {
    const char* utf8;
    int32_t len;
    tt_result* out = nullptr;
    int32_t rc = TT_OK;
    std::string err;
    bool done = false;
  };
  std::mutex cq_mu;
  std::condition_variable cq_cv;
  std::vector<CreatorRequest*> cq;
  bool cq_busy = false;
  int64_t cq_batches = 0, cq_requests = 0;  // passes made / requests answered through the queue
  tt::SortScratch sort;  // K5 radix sort buffers, grow-only
  tt::RadixSortInfo last_order;  // how the last OrderBy(TaskCreatedOn) went (tt_last_order_info)
  bool creator_smem_opted = false, json_smem_opted = false;  // large dynamic shared memory enabled on this device

  // payload: TaskName and TaskAssignedTo as dictionary-encoded string columns.  No kernel of the
  // sweep or the filter reads them; they exist so a result can be written out as JSON (K6).
  // ref == 0xFFFFFFFF (rows inserted later, or no payload loaded) prints an empty string.
  struct StrColumn {
    uint32_t* ref = nullptr;  // [n_pad]
    uint32_t n = 0;           // dictionary entries
    uint32_t* off = nullptr;  // [n + 1]
    uint8_t* pool = nullptr;
    // single-row updates (tt_set_task_strings): device capacities and a host mirror built on first use
    size_t off_cap = 0, pool_cap = 0;
    uint64_t pool_bytes = 0;
    bool mirrored = false;
    std::vector<uint32_t> h_off;
    std::vector<uint8_t> h_pool;
    std::unordered_map<std::string, uint32_t> h_index;
  } name_col, assignee_col;
  // escaped dictionaries for K6, built on first use: [name, creator, assignee][response, document]
  struct EscCache {
    tt::EscEntry* entry = nullptr;
    uint32_t* pool = nullptr;
    uint32_t n = 0;
    bool ready = false;
  } esc[3][2];
  // tt_tick_host: transfer streams, events and the two device staging buffers of the pipelined tick
  cudaStream_t tick_up = nullptr, tick_down = nullptr;
  cudaEvent_t tick_up_ev[2] = {nullptr, nullptr}, tick_down_ev[2] = {nullptr, nullptr}, tick_done_ev = nullptr;
  uint32_t* tick_rows[2] = {nullptr, nullptr};
  void* tick_ids[2] = {nullptr, nullptr};
  int64_t tick_cap[2] = {0, 0};
  uint64_t* json_blocks = nullptr;  // K6 block sums / bases
  int64_t json_blocks_cap = 0;
  uint64_t* json_plan = nullptr;    // K6 plan: 5 words per record, grow-only
  int64_t json_plan_cap = 0;
  tt_json_stats last_json = {};
};

// One device allocation, shared by the results that point into it (a batch of creator queries
// returns one result per query over a single block).
struct DeviceBlock {
  void* p = nullptr;
  size_t bytes = 0;
  std::shared_ptr<TableCore> core;
  ~DeviceBlock() {
    if (!p) return;
    if (core) {
      std::lock_guard<std::mutex> lk(core->pool_mu);
      if (core->alive && (!core->spare || core->spare_bytes < bytes)) {  // keep the larger of the two
        std::swap(core->spare, p);
        std::swap(core->spare_bytes, bytes);
      }
    }
    if (p) cudaFree(p);
  }
};

struct tt_result {
  std::shared_ptr<TableCore> core;
  tt_table* t = nullptr;  // valid only while core->alive
  int64_t n = 0;
  int64_t cap = 0;
  uint32_t* d_rows = nullptr;
  uint4* d_ids = nullptr;
  std::shared_ptr<DeviceBlock> block;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    cudaGetLastError();  // do not inherit a stale error from earlier, unrelated runtime calls
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

void free_columns(tt_table* t) {
  cudaFree(t->due_code);
  cudaFree(t->completed);
  cudaFree(t->overdue);
  cudaFree(t->creator_hash);
  cudaFree(t->task_id);
  cudaFree(t->created_raw);
  cudaFree(t->due_raw);
  cudaFree(t->creator_ref);
  cudaFree(t->creator_off);
  cudaFree(t->creator_pool);
  cudaFree(t->id_order);
  cudaFree(t->tile_words);
  cudaFree(t->group_sum);
  cudaFree(t->hit_list);
  for (tt_table::StrColumn* c : {&t->name_col, &t->assignee_col}) {
    cudaFree(c->ref);
    cudaFree(c->off);
    cudaFree(c->pool);
    *c = tt_table::StrColumn();
  }
  for (auto& per_col : t->esc)
    for (auto& e : per_col) {
      cudaFree(e.entry);
      cudaFree(e.pool);
      e = tt_table::EscCache();
    }
  cudaFree(t->json_blocks);
  cudaFree(t->json_plan);
  t->json_blocks = t->json_plan = nullptr;
  t->json_blocks_cap = t->json_plan_cap = 0;
  t->due_code = t->completed = t->overdue = nullptr;
  t->creator_hash = nullptr;
  t->task_id = nullptr;
  t->created_raw = t->due_raw = nullptr;
  t->creator_ref = t->creator_off = t->id_order = nullptr;
  t->creator_pool = nullptr;
  t->tile_words = t->group_sum = t->hit_list = nullptr;
  t->n_tiles = t->n_groups = t->n_subs = t->hit_list_cap = 0;
  t->ctl_clean = false;
  t->n = t->n_pad = 0;
}

// Per-tile words and K1's partial sums for a table of np padded rows (the sweep's fixed scratch).
int32_t alloc_sweep_scratch(tt_table* t, int64_t np) {
  cudaFree(t->tile_words);
  cudaFree(t->group_sum);
  t->tile_words = t->group_sum = nullptr;
  t->ctl_clean = false;
  t->n_tiles = np / TT_TILE_ROWS;
  t->n_groups = (t->n_tiles + tt::kTilesPerGroup - 1) / tt::kTilesPerGroup;
  t->n_subs = (t->n_tiles + tt::kTilesPerSub - 1) / tt::kTilesPerSub;
  TT_CUDA(cudaMalloc(&t->tile_words, 3 * t->n_tiles * 4));
  TT_CUDA(cudaMalloc(&t->group_sum, (t->n_groups + t->n_subs) * 4));
  return TT_OK;
}

int32_t alloc_columns(tt_table* t, int64_t n) {
  if (t->n_pad != 0) return fail(TT_ERR_STATE, "table already holds %lld rows", (long long)t->n);
  if (n < 0 || n > kMaxRows) return fail(TT_ERR_ARG, "row count %lld outside [0, 2^31]", (long long)n);
  int64_t np = std::max<int64_t>(round_up(n, TT_TILE_ROWS), TT_TILE_ROWS);
  TT_CUDA(cudaMalloc(&t->due_code, np * 4));
  TT_CUDA(cudaMalloc(&t->completed, np / 8));
  TT_CUDA(cudaMalloc(&t->overdue, np / 8));
  TT_CUDA(cudaMalloc(&t->creator_hash, np * 8));
  TT_CUDA(cudaMalloc(&t->task_id, np * 16));
  TT_CUDA(cudaMalloc(&t->created_raw, np * 8));
  TT_CUDA(cudaMalloc(&t->due_raw, np * 8));
  TT_CUDA(cudaMalloc(&t->creator_ref, np * 4));
  {
    int32_t rc = alloc_sweep_scratch(t, np);
    if (rc) return rc;
  }
  // padding rows never match and carry no flags
  TT_CUDA(cudaMemsetAsync(t->due_code, 0xFF, np * 4, t->stream));
  TT_CUDA(cudaMemsetAsync(t->completed, 0, np / 8, t->stream));
  TT_CUDA(cudaMemsetAsync(t->overdue, 0, np / 8, t->stream));
  TT_CUDA(cudaMemsetAsync(t->creator_hash, 0, np * 8, t->stream));
  TT_CUDA(cudaMemsetAsync(t->task_id, 0, np * 16, t->stream));
  TT_CUDA(cudaMemsetAsync(t->created_raw, 0, np * 8, t->stream));
  TT_CUDA(cudaMemsetAsync(t->due_raw, 0, np * 8, t->stream));
  TT_CUDA(cudaMemsetAsync(t->creator_ref, 0, np * 4, t->stream));
  t->n = n;
  t->n_live = n;
  t->n_pad = np;
  return TT_OK;
}

int32_t upload_dictionary(tt_table* t, uint32_t n_creators, const uint32_t* off, const uint8_t* pool) {
  t->n_creators = n_creators;
  t->pool_bytes = n_creators ? off[n_creators] : 0;
  t->creator_off_cap = (size_t)n_creators + 1;
  t->creator_pool_cap = std::max<uint64_t>(t->pool_bytes, 16);
  TT_CUDA(cudaMalloc(&t->creator_off, t->creator_off_cap * 4));
  TT_CUDA(cudaMalloc(&t->creator_pool, t->creator_pool_cap));
  TT_CUDA(cudaMemcpyAsync(t->creator_off, off, ((size_t)n_creators + 1) * 4, cudaMemcpyHostToDevice, t->stream));
  if (t->pool_bytes)
    TT_CUDA(cudaMemcpyAsync(t->creator_pool, pool, t->pool_bytes, cudaMemcpyHostToDevice, t->stream));
  return TT_OK;
}

// ---- small kernels --------------------------------------------------------------------------

// Synthetic fill: one thread per row, a warp covers one flag word.
__global__ void fill_synth_kernel(tt_synth_cfg cfg, uint64_t first_row, int64_t n, uint32_t* due_code,
                                  uint32_t* completed, uint32_t* overdue, uint64_t* creator_hash, uint4* task_id,
                                  uint64_t* created_raw, uint64_t* due_raw, uint32_t* creator_ref,
                                  const uint64_t* dict_hash) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;  // grid covers a multiple of 32 rows
  bool live = i < n;
  bool c = false, o = false;
  if (live) {
    tt_synth_row r = tt_synth_make(&cfg, first_row + (uint64_t)i);
    due_code[i] = tt_pack_due(r.due_raw);
    due_raw[i] = r.due_raw;
    created_raw[i] = r.created_raw;
    creator_ref[i] = r.creator;
    creator_hash[i] = dict_hash[r.creator];
    // canonical big-endian bytes 0..15 laid out in memory order: x = bytes 0..3, ...
    uint4 id;
    id.x = __byte_perm((uint32_t)(r.id_hi >> 32), 0, 0x0123);
    id.y = __byte_perm((uint32_t)r.id_hi, 0, 0x0123);
    id.z = __byte_perm((uint32_t)(r.id_lo >> 32), 0, 0x0123);
    id.w = __byte_perm((uint32_t)r.id_lo, 0, 0x0123);
    task_id[i] = id;
    c = r.completed;
    o = r.overdue;
  }
  uint32_t cw = __ballot_sync(0xFFFFFFFFu, c), ow = __ballot_sync(0xFFFFFFFFu, o);
  if ((threadIdx.x & 31) == 0 && i < ((n + 31) & ~31ll)) {
    completed[i >> 5] = cw;
    overdue[i >> 5] = ow;
  }
}

__global__ void hash_rows_kernel(const uint32_t* creator_ref, const uint64_t* dict_hash, uint64_t* creator_hash,
                                 int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) creator_hash[i] = dict_hash[creator_ref[i]];
}

// popcount of both flag planes -> out[0], out[1]
__global__ void count_flags_kernel(const uint32_t* completed, const uint32_t* overdue, int64_t words,
                                   unsigned long long* out) {
  unsigned long long c = 0, o = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (int64_t)gridDim.x * blockDim.x) {
    c += __popc(completed[i]);
    o += __popc(overdue[i]);
  }
  for (int d = 16; d > 0; d >>= 1) {
    c += __shfl_xor_sync(0xFFFFFFFFu, c, d);
    o += __shfl_xor_sync(0xFFFFFFFFu, o, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(out + 0, c);
    atomicAdd(out + 1, o);
  }
}

__device__ __forceinline__ uint64_t be64(uint32_t a, uint32_t b) {
  return ((uint64_t)__byte_perm(a, 0, 0x0123) << 32) | __byte_perm(b, 0, 0x0123);
}

// XOR of ids (as big-endian hi/lo u64) and sum of rows -> out[0..2]
__global__ void digest_kernel(const uint32_t* rows, const uint4* ids, int64_t n, unsigned long long* out) {
  unsigned long long xh = 0, xl = 0, sr = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint4 id = ids[i];
    xh ^= be64(id.x, id.y);
    xl ^= be64(id.z, id.w);
    sr += rows[i];
  }
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

template <typename T>
__global__ void gather_field_kernel(const uint32_t* rows, const T* col, T* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = col[rows[i]];
}

__global__ void gather_bit_kernel(const uint32_t* rows, const uint32_t* plane, uint8_t* out, int64_t n) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    uint32_t r = rows[i];
    out[i] = (plane[r >> 5] >> (r & 31)) & 1u;
  }
}

inline unsigned grid_for(int64_t n, int block) { return (unsigned)std::max<int64_t>(1, (n + block - 1) / block); }

// ---- the resumable sweep driver -------------------------------------------------------------

int32_t check_clock(int64_t today_local_ticks, uint32_t* yesterday) {
  if (today_local_ticks < TT_TICKS_PER_DAY || today_local_ticks % TT_TICKS_PER_DAY != 0)
    return fail(TT_ERR_ARG, "today_local_ticks must be a midnight (DateTime.Today) on or after 0001-01-02");
  int64_t day = today_local_ticks / TT_TICKS_PER_DAY - 1;
  if (day > (int64_t)TT_DUE_DAY_MASK) return fail(TT_ERR_ARG, "today_local_ticks beyond DateTime.MaxValue");
  *yesterday = (uint32_t)day;
  return TT_OK;
}

// Which K3 a plain sweep uses: the tile-ordered place_group_kernel (default; measured 0.49 ms for
// 25 M hits on B200) or the older hit-ordered place_hits_kernel (0.55 ms; TT_PLACE_KERNEL=hits,
// read once, kept for A/B timing).
bool place_by_group() {
  static const bool by_group = [] {
    const char* e = getenv("TT_PLACE_KERNEL");
    return e ? strcmp(e, "hits") != 0 : true;
  }();
  return by_group;
}

// Persistent CTAs per SM for K1 (TT_SWEEP_CTAS=1..4, read once; the kernel is built for up to 4).
int sweep_ctas_per_sm() {
  static const int ctas = [] {
    const char* e = getenv("TT_SWEEP_CTAS");
    const int v = e ? atoi(e) : tt::kSweepCtasPerSm;
    return v >= 1 && v <= tt::kSweepCtasPerSm ? v : tt::kSweepCtasPerSm;
  }();
  return ctas;
}

// Waits for the tick queued by sweep_run and turns the tail kernel's report into *st.
int32_t sweep_finish(tt_table* t, tt_sweep_stats* st) {
  const tt_table::PendingSweep pd = t->pending;
  t->pending.active = false;
  memset(st, 0, sizeof *st);
  st->launches = pd.launches;
  if (pd.trivial) {
    st->resume_row = pd.row_end;
    return TT_OK;
  }
  // An untimed tick ends when the tail kernel's last store — the tick's number, in mapped host memory — is seen:
  // watching that word for a short while costs less than the wake-up of cudaStreamSynchronize (what is left of a
  // 10 M-row tick is launch and wake-up time).  Long ticks, failed ones and timed ones wait on the stream.
  bool seen = false;
  if (!pd.timing && pd.seq) {
    const volatile unsigned long long* word = &t->h_out->seq;
    const auto t0 = std::chrono::steady_clock::now();
    for (int spin = 0; !(seen = *word == pd.seq); ++spin)
      if ((spin & 63) == 63 && std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(150)) break;
    if (seen) std::atomic_thread_fence(std::memory_order_acquire);
  }
  if (!seen) TT_CUDA(cudaStreamSynchronize(pd.stream));
  t->ctl_clean = true;
  if (pd.timing) {
    TT_CUDA(cudaEventElapsedTime(&st->sweep_ms, t->ev0, t->ev1));
    TT_CUDA(cudaEventElapsedTime(&st->gather_ms, t->ev1, t->ev2));
    st->kernel_ms = st->sweep_ms + st->gather_ms;
  }
  const tt::SweepOut done = *t->h_out;
  st->n_listed = (int64_t)done.cursor;
  const bool complete = done.stored == ~0ull;
  st->n_stored = pd.listing ? (complete ? st->n_listed : (int64_t)done.stored) : 0;
  st->resume_row = complete ? (pd.row_end >= 0 ? std::min<int64_t>(pd.row_end, t->n) : t->n) : pd.first_row;
  st->n_marked = pd.mark_enable ? (pd.listing ? st->n_stored : st->n_listed) : 0;
  if (!complete)
    return fail(TT_ERR_CAPACITY, "%lld hits do not fit in %lld (%lld stored%s)", (long long)st->n_listed,
                (long long)pd.cap, (long long)st->n_stored, st->n_marked ? " and marked; call again for the rest" : "");
  return TT_OK;
}

// Launches the sweep kernel over rows [first_row, n) on `stream`, waits, fills *st.
int32_t sweep_run(tt_table* t, int32_t mode, int32_t flags, int64_t today_local_ticks, int64_t run_at_utc_ticks,
                  int64_t first_row, uint32_t* d_rows, uint4* d_ids, int64_t cap, tt_sweep_stats* st,
                  cudaStream_t stream, const tt::PushTarget* push = nullptr, int64_t row_end = -1,
                  const tt::ShardHook* hook = nullptr) {
  memset(st, 0, sizeof *st);
  if (t->pending.active) return fail(TT_ERR_STATE, "a tick queued with TT_SWEEP_ASYNC is pending: call tt_sweep_wait first");
  if (mode != TT_MODE_FAKE && mode != TT_MODE_STORE) return fail(TT_ERR_ARG, "mode must be TT_MODE_FAKE or TT_MODE_STORE");
  uint32_t yesterday = 0;
  int32_t rc = check_clock(today_local_ticks, &yesterday);
  if (rc) return rc;
  if (first_row < 0 || first_row % TT_TILE_ROWS != 0 || first_row > t->n_pad)
    return fail(TT_ERR_ARG, "first_row must be a multiple of %d inside the table", TT_TILE_ROWS);
  if (cap < 0) return fail(TT_ERR_ARG, "negative capacity");
  const bool mark = (flags & TT_SWEEP_MARK) != 0;
  const bool range = (flags & TT_SWEEP_DUE_ON_OR_BEFORE) != 0;
  // ScheduledTasksManagerController.cs:34 — runAt.Date > TaskDueDate.Date; for listed rows the
  // right-hand side is yesterday's day number (see sweep.cuh).  In range mode the listed rows have
  // different due days: the predicate is uniform only while runAt is past yesterday, which is the
  // cron's situation (it runs at 00:05 of today); a clock that far behind is refused.
  const bool survive = run_at_utc_ticks / TT_TICKS_PER_DAY > (int64_t)yesterday;
  if (range && mark && !survive)
    return fail(TT_ERR_ARG, "range sweep: run_at_utc_ticks must lie after yesterday (runAt.Date > TaskDueDate.Date would differ per row)");
  auto nothing_to_do = [&](int64_t resume) {  // an empty range: no kernel runs, an async caller still gets its wait
    st->resume_row = resume;
    if (flags & TT_SWEEP_ASYNC) {
      t->pending = tt_table::PendingSweep();
      t->pending.active = t->pending.trivial = true;
      t->pending.row_end = resume;
    }
    return TT_OK;
  };
  if (t->n == 0 || first_row >= t->n) return nothing_to_do(t->n);
  const uint32_t first_tile = (uint32_t)(first_row / TT_TILE_ROWS);
  uint32_t num_tiles = (uint32_t)(t->n_pad / TT_TILE_ROWS) - first_tile;
  if (row_end >= 0) {  // a chunk [first_row, row_end) of the table, whole tiles
    if (row_end % TT_TILE_ROWS != 0 || row_end < first_row || row_end > t->n_pad)
      return fail(TT_ERR_ARG, "row_end must be a tile multiple in [first_row, padded rows]");
    num_tiles = (uint32_t)((row_end - first_row) / TT_TILE_ROWS);
    if (num_tiles == 0) return nothing_to_do(row_end);
  }

  // scratch for the unordered hit list K1 writes before K3 puts it in row order
  const bool listing = d_rows != nullptr || push != nullptr;
  if (listing && cap > t->hit_list_cap) {
    cudaFree(t->hit_list);
    t->hit_list = nullptr;
    t->hit_list_cap = 0;
    TT_CUDA(cudaMalloc(&t->hit_list, (size_t)std::max<int64_t>(cap, 1) * 4));
    t->hit_list_cap = cap;
  }

  tt::SweepParams p;
  p.due_code = t->due_code;
  p.completed = t->completed;
  p.overdue = t->overdue;
  p.list = listing ? t->hit_list : nullptr;
  p.tile_slot = t->tile_words;
  p.tile_count = t->tile_words + t->n_tiles;
  uint32_t* tile_first = t->tile_words + 2 * t->n_tiles;
  p.group_sum = t->group_sum;
  p.sub_sum = t->group_sum + t->n_groups;
  p.ctl = t->ctl;
  p.cap = listing ? (unsigned long long)cap : ~0ull;
  p.first_tile = first_tile;
  p.num_tiles = num_tiles;
  p.mask = mode == TT_MODE_FAKE ? TT_MASK_FAKE : TT_MASK_STORE;
  p.yesterday = yesterday;
  // FAKE mode never mutates: MarkOverdueTasks throws before touching the list (FakeTasksManager.cs:104)
  p.mark_enable = (mark && survive && mode == TT_MODE_STORE) ? 1u : 0u;

  // The control block and the partial sums are left in their start state by the previous tick's
  // tail kernel; only the first tick of a table (or one after a failed call) sets them from the host.
  const bool timing = (flags & TT_SWEEP_NO_TIMING) == 0;
  // programmatic dependent launch of K3g and the tail kernel (sweep.cuh); TT_NO_PDL=1 launches them the ordinary way
  static const bool pdl = getenv("TT_NO_PDL") == nullptr;
  const bool was_clean = t->ctl_clean;
  t->ctl_clean = false;
  if (!was_clean) {
    tt::SweepCtl init;
    init.tile_counter = 0;
    init.pad = 0;
    init.cursor = 0;
    init.stored = ~0ull;
    *t->h_ctl = init;
    TT_CUDA(cudaMemcpyAsync(t->ctl, t->h_ctl, sizeof init, cudaMemcpyHostToDevice, stream));
    TT_CUDA(cudaMemsetAsync(t->group_sum, 0, (size_t)(t->n_groups + t->n_subs) * 4, stream));
  }
  const unsigned n_groups = (num_tiles + tt::kTilesPerGroup - 1) / tt::kTilesPerGroup;
  const unsigned n_subs = (num_tiles + tt::kTilesPerSub - 1) / tt::kTilesPerSub;
  const unsigned grid = (unsigned)std::min<int64_t>(num_tiles, (int64_t)t->sm_count * sweep_ctas_per_sm());
  if (timing) TT_CUDA(cudaEventRecord(t->ev0, stream));
  constexpr int kT = tt::kSweepThreads, kC = tt::kSweepCtasPerSm;
  if (range) {
    if (mode == TT_MODE_STORE) {
      if (p.mark_enable) tt::sweep_kernel<true, true, kT, kC, true><<<grid, kT, 0, stream>>>(p);
      else tt::sweep_kernel<true, false, kT, kC, true><<<grid, kT, 0, stream>>>(p);
    } else {
      tt::sweep_kernel<false, false, kT, kC, true><<<grid, kT, 0, stream>>>(p);
    }
  } else if (mode == TT_MODE_STORE) {
    if (p.mark_enable) tt::sweep_kernel<true, true, kT, kC><<<grid, kT, 0, stream>>>(p);
    else tt::sweep_kernel<true, false, kT, kC><<<grid, kT, 0, stream>>>(p);
  } else {
    tt::sweep_kernel<false, false, kT, kC><<<grid, kT, 0, stream>>>(p);
  }
  TT_CUDA(cudaGetLastError());
  if (timing) TT_CUDA(cudaEventRecord(t->ev1, stream));
  st->launches = 1;
  if (listing) {
    // K3: row order and TaskIds; the hit count is read from the control block on the device
    if (push || place_by_group()) {
      // K3g: tile-ordered placement, optionally fused with the push of the compact form (sweep.cuh); it
      // derives every tile's offset in the row-ordered list from K1's partial sums itself (no K2)
      const tt::PushTarget none = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, 0};
      if (hook && hook->wait_slot0) {  // the root must have read this shard's slot of the previous tick
        tt::shard_wait_kernel<<<1, 32, 0, stream>>>(hook->wait_slot0, 0, 1, offsetof(tt::SlotHeader, consumed),
                                                    hook->wait_value, hook->timeout_ns, nullptr, nullptr);
        TT_CUDA(cudaGetLastError());
        st->launches += 1;
      }
      // Tiles per CTA and gathers per thread, measured on B200 at 25 M hits (plain placement): 1 tile x 1
      // gather 0.439 ms, 2 x 2 0.464, 4 x 4 0.474, 8 x 4 0.488, 16 x 4 0.503, 1 tile on 128 threads x 2 0.470 —
      // one tile per CTA, one hit per thread wins: all of a tile's gathers fall into one 128 KB span of the
      // TaskId column and are in flight together.  With the push fused in, the shape that is best on
      // one GPU is the worst at scale: K3g time on 2 GPUs 1 x 1 0.472 ms, 8 x 4 0.509; but the whole
      // tick on 8 GPUs (rank 0 takes the stores of 7 peers) 1 x 1 1.683 ms, 2 x 2 1.511, 4 x 4 1.393,
      // 8 x 4 1.349 — larger contiguous pieces mean fewer, larger NVLink packets into rank 0, so the
      // push keeps 8 x 4.  TT_PLACE_VARIANT / TT_PUSH_VARIANT (read once) pick another shape for A/B
      // runs: 1 = 1 x 1, 2 = 2 x 2, 4 = 4 x 4, 8 = 8 x 4.
      static const int plain_variant = [] { const char* e = getenv("TT_PLACE_VARIANT"); return e ? atoi(e) : 1; }();
      const char* push_env = push ? getenv("TT_PUSH_VARIANT") : nullptr;  // read per tick: A/B scripts change it between runs
      const int push_variant = push_env ? atoi(push_env) : 8;
      const bool push_forced = push_env != nullptr;
#define TT_PLACE_LAUNCH(GT, UN)                                                                                     \
  TT_CUDA(tt::launch_dependent(tt::place_group_kernel<TT_TILE_ROWS, GT, UN>, dim3((num_tiles + GT - 1) / GT),           \
                               dim3(tt::kGroupThreads), stream, pdl, (const uint32_t*)t->hit_list,                      \
                               (const uint32_t*)p.tile_slot, (const uint32_t*)p.tile_count,                             \
                               (const uint32_t*)p.group_sum, (const uint32_t*)p.sub_sum, (uint32_t)num_tiles,           \
                               (const uint4*)t->task_id, (const uint64_t*)t->created_raw, d_rows, d_ids,                \
                               push ? *push : none))
      // (a push of creation times only stays on this GPU; a push to ONE receiver that takes nobody else's stores asks for
      // the single-GPU shape: push->shape)
      switch (push && push->off16 ? (push->shape && !push_forced ? push->shape : push_variant) : plain_variant) {
        case 2: TT_PLACE_LAUNCH(2, 2); break;
        case 4: TT_PLACE_LAUNCH(4, 4); break;
        case 8: TT_PLACE_LAUNCH(8, 4); break;
        case 16: TT_PLACE_LAUNCH(16, 4); break;
        case 32: TT_PLACE_LAUNCH(32, 4); break;
        default: TT_PLACE_LAUNCH(1, 1); break;
      }
#undef TT_PLACE_LAUNCH
      st->launches += 1;
    } else {
      tt::tile_offsets_kernel<<<n_groups, tt::kTilesPerGroup, 0, stream>>>(p.tile_count, p.group_sum, num_tiles, tile_first);
      TT_CUDA(cudaGetLastError());
      const int64_t per_cta = (int64_t)tt::kPlaceThreads * tt::kPlaceUnroll;
      const unsigned pgrid =
          (unsigned)std::max<int64_t>(1, std::min<int64_t>((cap + per_cta - 1) / per_cta, (int64_t)t->sm_count * 8));
      tt::place_hits_kernel<TT_TILE_ROWS><<<pgrid, tt::kPlaceThreads, 0, stream>>>(
          t->ctl, t->hit_list, p.tile_slot, tile_first, first_tile, t->task_id, d_rows, d_ids);
      st->launches = 3;
    }
    TT_CUDA(cudaGetLastError());
  }
  // the epilogue: outcome -> mapped host memory (and the push target), scratch back to its start state,
  // the epoch word of a sharded tick
  TT_CUDA(tt::launch_dependent(tt::sweep_tail_kernel, dim3(1), dim3(256), stream, pdl, t->ctl, t->d_out, t->group_sum,
                               (uint32_t)n_groups, p.sub_sum, (uint32_t)n_subs, push ? push->count : nullptr,
                               hook ? hook->epoch_word : nullptr, hook ? hook->epoch : 0ull,
                               (unsigned long long)++t->tick_seq));
  st->launches += 1;
  if (timing) TT_CUDA(cudaEventRecord(t->ev2, stream));
  t->pending = tt_table::PendingSweep();
  t->pending.active = true;
  t->pending.listing = listing;
  t->pending.timing = timing;
  t->pending.mark_enable = p.mark_enable;
  t->pending.first_row = first_row;
  t->pending.row_end = row_end;
  t->pending.cap = cap;
  t->pending.launches = st->launches;
  t->pending.stream = stream;
  t->pending.seq = t->tick_seq;
  if (flags & TT_SWEEP_ASYNC) return TT_OK;  // the caller collects the outcome with tt_sweep_wait
  return sweep_finish(t, st);
}

int32_t result_alloc(tt_table* t, int64_t cap, tt_result** out) {
  tt_result* r = new (std::nothrow) tt_result();
  if (!r) return fail(TT_ERR_OOM, "out of host memory");
  r->t = t;
  r->core = t->core;
  r->cap = std::max<int64_t>(cap, 1);
  r->block = std::make_shared<DeviceBlock>();
  r->block->core = t->core;
  r->block->bytes = (size_t)r->cap * 20;  // ids (16 B each) first, then rows
  {
    std::lock_guard<std::mutex> lk(t->core->pool_mu);
    if (t->core->spare && t->core->spare_bytes >= r->block->bytes && t->core->spare_bytes <= 4 * r->block->bytes + (1u << 20)) {
      r->block->p = t->core->spare;
      r->block->bytes = t->core->spare_bytes;
      t->core->spare = nullptr;
      t->core->spare_bytes = 0;
    }
  }
  cudaError_t e = r->block->p ? cudaSuccess : cudaMalloc(&r->block->p, r->block->bytes);
  if (e != cudaSuccess) {
    cudaGetLastError();
    delete r;
    return fail(TT_ERR_OOM, "result buffers for %lld hits: %s", (long long)cap, cudaGetErrorString(e));
  }
  r->d_ids = static_cast<uint4*>(r->block->p);
  r->d_rows = reinterpret_cast<uint32_t*>(static_cast<unsigned char*>(r->block->p) + (size_t)r->cap * 16);
  *out = r;
  return TT_OK;
}

// A result over [first, first + n) of another result's buffers (shares the allocation).
tt_result* result_slice(const tt_result* whole, int64_t first, int64_t n) {
  tt_result* r = new (std::nothrow) tt_result();
  if (!r) return nullptr;
  r->t = whole->t;
  r->core = whole->core;
  r->n = r->cap = n;
  r->d_rows = whole->d_rows + first;
  r->d_ids = whole->d_ids + first;
  r->block = whole->block;
  return r;
}

void result_release(tt_result* r) { delete r; }  // the block goes with its last result

// Sweep into a library-owned result.  The first guess at the hit count is n/16; when it is too
// small the result is regrown to the exact size and the sweep runs again: from scratch if the
// first pass changed nothing, otherwise for the hits the first pass did not reach (the rows it
// marked no longer match), after which the two row-ordered runs are merged by a sort.
// with_ids = false: the TaskIds are not gathered here — the caller re-orders the list and gathers them in the
// final order (a gather per hit saved).
// keys_too (listing sweeps only): the placement kernel also gathers every hit's TaskCreatedOn ticks — into
// t->sort.keys[0], with the rows in t->sort.vals[0] instead of the result — so that OrderBy(TaskCreatedOn) starts
// from a filled sort buffer instead of gathering the keys in a pass of its own.
int32_t sweep_owned(tt_table* t, int32_t mode, int32_t flags, int64_t today, int64_t run_at, tt_result** out,
                    tt_sweep_stats* total, bool with_ids = true, bool keys_too = false) {
  memset(total, 0, sizeof *total);
  if (keys_too && ((flags & TT_SWEEP_MARK) || with_ids)) return fail(TT_ERR_ARG, "keys_too goes with a listing sweep without ids");
  tt_result* r = nullptr;
  int64_t guess = std::min<int64_t>(t->n, std::max<int64_t>(65536, t->n / 16));
  int32_t rc = result_alloc(t, guess, &r);
  if (rc) return rc;
  bool merged = false;
  for (int pass = 0; pass < 8; ++pass) {
    tt_sweep_stats st;
    if (keys_too) {
      const cudaError_t e = t->sort.reserve(r->cap);
      if (e != cudaSuccess) {
        result_release(r);
        cudaGetLastError();
        return fail(TT_ERR_OOM, "sort buffers: %s", cudaGetErrorString(e));
      }
      tt::PushTarget keys = {nullptr, nullptr, nullptr, nullptr, t->sort.keys[0], nullptr, 1, 0};
      rc = sweep_run(t, mode, flags, today, run_at, 0, t->sort.vals[0], nullptr, r->cap, &st, t->stream, &keys);
    } else {
      rc = sweep_run(t, mode, flags, today, run_at, 0, r->d_rows + r->n, with_ids ? r->d_ids + r->n : nullptr, r->cap - r->n,
                     &st, t->stream);
    }
    total->kernel_ms += st.kernel_ms;
    total->sweep_ms += st.sweep_ms;
    total->gather_ms += st.gather_ms;
    total->launches += st.launches;
    if (rc != TT_OK && rc != TT_ERR_CAPACITY) {
      result_release(r);
      return rc;
    }
    total->n_marked += st.n_marked;
    if (rc == TT_OK) {
      r->n += st.n_stored;
      if (merged && r->n > 1) {
        int e = tt::sort_rows_and_gather(t->stream, t->sort, t->task_id, r->d_rows, with_ids ? r->d_ids : nullptr, r->n);
        if (e != 0) {
          result_release(r);
          return fail(TT_ERR_CUDA, "sort_rows_and_gather: %s", cudaGetErrorString((cudaError_t)e));
        }
      }
      total->n_listed = total->n_stored = r->n;
      total->resume_row = t->n;
      *out = r;
      return TT_OK;
    }
    // too small: regrow to what is still needed
    const bool mutated = st.n_marked > 0;
    const int64_t keep = mutated ? r->n + st.n_stored : 0;
    tt_result* g = nullptr;
    rc = result_alloc(t, keep + (st.n_listed - (mutated ? st.n_stored : 0)), &g);
    if (rc) {
      result_release(r);
      return rc;
    }
    if (keep) {
      cudaMemcpyAsync(g->d_rows, r->d_rows, keep * 4, cudaMemcpyDeviceToDevice, t->stream);
      if (with_ids) cudaMemcpyAsync(g->d_ids, r->d_ids, keep * 16, cudaMemcpyDeviceToDevice, t->stream);
      cudaStreamSynchronize(t->stream);
      merged = true;
    }
    g->n = keep;
    result_release(r);
    r = g;
  }
  result_release(r);
  return fail(TT_ERR_STATE, "sweep did not converge");
}

// The TaskId index: rows in TaskId order (t->id_order) for a table whose rows are not.  Built on
// the device: one stable 64-bit radix sort of the HIGH 8 canonical bytes orders the rows whenever those
// are all distinct (Guid.NewGuid() ids: 58 random bits up there) — a count of equal neighbours checks it;
// otherwise two sorts, the low 8 bytes first, then the high 8 bytes of the rows in that order.  No-op while
// the table itself is in TaskId order or the index still covers every row (tt_insert_task drops it when
// it appends an out-of-order id).
int32_t ensure_id_order(tt_table* t) {
  if (t->ids_sorted || t->id_order || t->n == 0) return TT_OK;
  const int64_t n = t->n;
  cudaStream_t st = t->stream;
  cudaError_t e = t->sort.reserve(n);
  int w = 0;
  unsigned long long ties = 1;
  if (e == cudaSuccess) {
    tt::id_half_keys_kernel<<<grid_for(n, 256), 256, 0, st>>>(t->task_id, nullptr, n, 1, t->sort.keys[0], t->sort.vals[0]);
    e = tt::radix_sort_auto(st, t->sort, n, &w);  // random ids: four passes over the top 32 bits at 25 M rows, the groups of equal prefixes put right
  }
  if (e == cudaSuccess) e = cudaMemsetAsync(t->d_small, 0, 8, st);
  if (e == cudaSuccess) {
    tt::adjacent_equal_kernel<<<grid_for(n, 256), 256, 0, st>>>(t->sort.keys[w], n, t->d_small);
    e = cudaMemcpyAsync(&ties, t->d_small, 8, cudaMemcpyDeviceToHost, st);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess && ties != 0) {  // some ids share their high half: the full two-key sort
    if (w != 0) {  // radix_sort_pairs starts from buffer pair 0
      std::swap(t->sort.keys[0], t->sort.keys[1]);
      std::swap(t->sort.vals[0], t->sort.vals[1]);
    }
    tt::id_half_keys_kernel<<<grid_for(n, 256), 256, 0, st>>>(t->task_id, nullptr, n, 0, t->sort.keys[0], t->sort.vals[0]);
    e = tt::radix_sort_pairs(st, t->sort, n, 64, &w);
    if (e == cudaSuccess) {
      tt::id_half_keys_kernel<<<grid_for(n, 256), 256, 0, st>>>(t->task_id, t->sort.vals[w], n, 1, t->sort.keys[w ^ 1],
                                                                t->sort.vals[w ^ 1]);
      if (w == 0) {
        std::swap(t->sort.keys[0], t->sort.keys[1]);
        std::swap(t->sort.vals[0], t->sort.vals[1]);
      }
      e = tt::radix_sort_pairs(st, t->sort, n, 64, &w);
    }
  }
  uint32_t* order = nullptr;
  if (e == cudaSuccess) e = cudaMalloc(&order, (size_t)n * 4);
  if (e == cudaSuccess) e = cudaMemcpyAsync(order, t->sort.vals[w], (size_t)n * 4, cudaMemcpyDeviceToDevice, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) {
    cudaFree(order);
    cudaGetLastError();
    return fail(e == cudaErrorMemoryAllocation ? TT_ERR_OOM : TT_ERR_CUDA, "TaskId index: %s", cudaGetErrorString(e));
  }
  t->id_order = order;
  return TT_OK;
}

}  // namespace

// =============================================================================================
extern "C" {

const char* tt_last_error(void) { return g_err.c_str(); }

int32_t tt_create(int32_t device, tt_table** out) {
  if (!out) return fail(TT_ERR_ARG, "out is null");
  *out = nullptr;
  int count = 0;
  TT_CUDA(cudaGetDeviceCount(&count));
  if (device < 0 || device >= count) return fail(TT_ERR_ARG, "device %d not in [0, %d)", device, count);
  DeviceGuard g(device);
  tt_table* t = new (std::nothrow) tt_table();
  if (!t) return fail(TT_ERR_OOM, "out of host memory");
  t->device = device;
  t->core->device = device;
  cudaDeviceProp prop;
  TT_CUDA(cudaGetDeviceProperties(&prop, device));
  t->sm_count = prop.multiProcessorCount;
  TT_CUDA(cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking));
  TT_CUDA(cudaEventCreate(&t->ev0));
  TT_CUDA(cudaEventCreate(&t->ev1));
  TT_CUDA(cudaEventCreate(&t->ev2));
  TT_CUDA(cudaMalloc(&t->ctl, sizeof(tt::SweepCtl)));
  TT_CUDA(cudaMallocHost(&t->h_ctl, sizeof(tt::SweepCtl)));
  TT_CUDA(cudaHostAlloc(&t->h_out, sizeof(tt::SweepOut), cudaHostAllocMapped));
  memset(t->h_out, 0, sizeof(tt::SweepOut));
  TT_CUDA(cudaHostGetDevicePointer(&t->d_out, t->h_out, 0));
  TT_CUDA(cudaMalloc(&t->d_small, 16 * 8));
  TT_CUDA(cudaMallocHost(&t->h_small, 16 * 8));
  TT_CUDA(cudaMalloc(&t->d_qcount, tt::kMaxQueries * 4));
  TT_CUDA(cudaMallocHost(&t->h_qcount, tt::kMaxQueries * 4));
  *out = t;
  return TT_OK;
}

int32_t tt_destroy(tt_table* t) {
  if (!t) return TT_OK;
  {
    DeviceGuard g(t->device);
    std::lock_guard<std::mutex> lk(t->mu);
    std::lock_guard<std::mutex> lk2(t->core->pool_mu);
    t->core->alive = false;
    cudaFree(t->core->spare);
    t->core->spare = nullptr;
    t->core->spare_bytes = 0;
    cudaFree(t->core->json_spare);
    t->core->json_spare = nullptr;
    t->core->json_spare_bytes = 0;
  }
  {
    DeviceGuard g(t->device);
    cudaStreamSynchronize(t->stream);
    free_columns(t);
    cudaFree(t->ctl);
    cudaFreeHost(t->h_ctl);
    cudaFreeHost(t->h_out);
    cudaFree(t->d_small);
    cudaFreeHost(t->h_small);
    t->sort.release();
    cudaFree(t->d_query);
    cudaFree(t->d_pairs);
    cudaFree(t->d_lists);
    cudaFree(t->d_qcount);
    cudaFreeHost(t->h_qcount);
    cudaEventDestroy(t->ev0);
    cudaEventDestroy(t->ev1);
    cudaEventDestroy(t->ev2);
    for (int k = 0; k < 2; ++k) {
      cudaFree(t->tick_rows[k]);
      cudaFree(t->tick_ids[k]);
      if (t->tick_up_ev[k]) cudaEventDestroy(t->tick_up_ev[k]);
      if (t->tick_down_ev[k]) cudaEventDestroy(t->tick_down_ev[k]);
    }
    if (t->tick_done_ev) cudaEventDestroy(t->tick_done_ev);
    if (t->tick_up) cudaStreamDestroy(t->tick_up);
    if (t->tick_down) cudaStreamDestroy(t->tick_down);
    cudaStreamDestroy(t->stream);
  }
  delete t;
  return TT_OK;
}

int32_t tt_row_count(tt_table* t, int64_t* n) {
  if (!t || !n) return fail(TT_ERR_ARG, "null argument");
  *n = t->n;
  return TT_OK;
}

int32_t tt_generate(tt_table* t, const tt_synth_cfg* cfg, uint64_t first_row, uint64_t n_rows) {
  if (!t || !cfg) return fail(TT_ERR_ARG, "null argument");
  if (cfg->n_creators == 0) return fail(TT_ERR_ARG, "n_creators must be > 0");
  if (cfg->today_local_ticks < 400 * TT_TICKS_PER_DAY || cfg->today_local_ticks % TT_TICKS_PER_DAY)
    return fail(TT_ERR_ARG, "cfg.today_local_ticks must be a midnight");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  int32_t rc = alloc_columns(t, (int64_t)n_rows);
  if (rc) return rc;
  // creator dictionary: strings, offsets and hashes built once on the host (ingest, not hot path)
  const uint32_t nc = cfg->n_creators;
  std::vector<uint32_t> off(nc + 1);
  std::vector<uint8_t> pool;
  pool.reserve((size_t)nc * 26);
  std::vector<uint64_t> hash(nc);
  char buf[48];
  for (uint32_t k = 0; k < nc; ++k) {
    int len = tt_synth_creator(k, buf);
    off[k] = (uint32_t)pool.size();
    pool.insert(pool.end(), buf, buf + len);
    hash[k] = tt_hash64((const uint8_t*)buf, (uint32_t)len);
  }
  off[nc] = (uint32_t)pool.size();
  rc = upload_dictionary(t, nc, off.data(), pool.data());
  if (rc) return rc;
  uint64_t* d_hash = nullptr;
  TT_CUDA(cudaMalloc(&d_hash, (size_t)nc * 8));
  TT_CUDA(cudaMemcpyAsync(d_hash, hash.data(), (size_t)nc * 8, cudaMemcpyHostToDevice, t->stream));
  if (n_rows) {
    int64_t threads = round_up((int64_t)n_rows, 32);
    fill_synth_kernel<<<grid_for(threads, 256), 256, 0, t->stream>>>(
        *cfg, first_row, (int64_t)n_rows, t->due_code, t->completed, t->overdue, t->creator_hash, t->task_id,
        t->created_raw, t->due_raw, t->creator_ref, d_hash);
    TT_CUDA(cudaGetLastError());
  }
  TT_CUDA(cudaStreamSynchronize(t->stream));
  cudaFree(d_hash);
  t->ids_sorted = true;  // tt_synth_make: TaskId grows with the row
  return TT_OK;
}

int32_t tt_load_soa(tt_table* t, const tt_soa_view* v) {
  if (!t || !v) return fail(TT_ERR_ARG, "null argument");
  if (v->n < 0) return fail(TT_ERR_ARG, "negative row count");
  if (v->n > 0 && (!v->task_id || !v->created_raw || !v->due_raw || !v->completed || !v->overdue || !v->creator_ref))
    return fail(TT_ERR_ARG, "null column");
  if (v->n > 0 && (v->n_creators == 0 || !v->creator_off || (!v->creator_pool && v->creator_off[v->n_creators] != 0)))
    return fail(TT_ERR_ARG, "creator dictionary missing");
  for (int64_t i = 0; i < v->n; ++i)
    if (v->creator_ref[i] >= v->n_creators) return fail(TT_ERR_ARG, "creator_ref[%lld] out of range", (long long)i);
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  const int64_t n = v->n;
  int32_t rc = alloc_columns(t, n);
  if (rc) return rc;
  // encode the hot columns on the host (ingest path)
  std::vector<uint32_t> code(n), cw((n + 31) / 32, 0), ow((n + 31) / 32, 0);
  for (int64_t i = 0; i < n; ++i) {
    code[i] = tt_pack_due(v->due_raw[i]);
    if (v->completed[i]) cw[i >> 5] |= 1u << (i & 31);
    if (v->overdue[i]) ow[i >> 5] |= 1u << (i & 31);
  }
  const uint32_t nc = v->n_creators;
  std::vector<uint64_t> hash(std::max<uint32_t>(nc, 1));
  for (uint32_t k = 0; k < nc; ++k) {
    if (v->creator_off[k + 1] < v->creator_off[k]) return fail(TT_ERR_ARG, "creator_off not ascending");
    hash[k] = tt_hash64(v->creator_pool + v->creator_off[k], v->creator_off[k + 1] - v->creator_off[k]);
  }
  static const uint32_t zero_off[1] = {0};
  rc = upload_dictionary(t, nc, nc ? v->creator_off : zero_off, v->creator_pool);
  if (rc) return rc;
  // is the table already in TaskId order?  (then mark-by-id can search it directly)
  bool sorted = true;
  for (int64_t i = 1; i < n && sorted; ++i) sorted = memcmp(v->task_id + 16 * (i - 1), v->task_id + 16 * i, 16) < 0;
  t->ids_sorted = sorted;
  if (!sorted && n) {
    std::vector<uint32_t> order(n);
    for (int64_t i = 0; i < n; ++i) order[i] = (uint32_t)i;
    const uint8_t* ids = v->task_id;
    std::stable_sort(order.begin(), order.end(),
                     [ids](uint32_t a, uint32_t b) { return memcmp(ids + 16 * (size_t)a, ids + 16 * (size_t)b, 16) < 0; });
    TT_CUDA(cudaMalloc(&t->id_order, n * 4));
    TT_CUDA(cudaMemcpyAsync(t->id_order, order.data(), n * 4, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaStreamSynchronize(t->stream));
  }
  if (n) {
    uint64_t* d_hash = nullptr;
    TT_CUDA(cudaMalloc(&d_hash, hash.size() * 8));
    TT_CUDA(cudaMemcpyAsync(d_hash, hash.data(), hash.size() * 8, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->due_code, code.data(), n * 4, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->completed, cw.data(), cw.size() * 4, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->overdue, ow.data(), ow.size() * 4, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->task_id, v->task_id, n * 16, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->created_raw, v->created_raw, n * 8, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->due_raw, v->due_raw, n * 8, cudaMemcpyHostToDevice, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->creator_ref, v->creator_ref, n * 4, cudaMemcpyHostToDevice, t->stream));
    hash_rows_kernel<<<grid_for(n, 256), 256, 0, t->stream>>>(t->creator_ref, d_hash, t->creator_hash, n);
    TT_CUDA(cudaGetLastError());
    TT_CUDA(cudaStreamSynchronize(t->stream));
    cudaFree(d_hash);
  }
  TT_CUDA(cudaStreamSynchronize(t->stream));
  return TT_OK;
}

// ---- ITasksManager operations ---------------------------------------------------------------

int32_t tt_get_yesterdays_due(tt_table* t, int32_t mode, int64_t today_local_ticks, tt_result** out) {
  if (!t || !out) return fail(TT_ERR_ARG, "null argument");
  *out = nullptr;
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  tt_sweep_stats st;
  tt_result* r = nullptr;
  int32_t rc = sweep_owned(t, mode, TT_SWEEP_LIST_ONLY, today_local_ticks, 0, &r, &st, mode != TT_MODE_STORE,
                           /*keys_too=*/mode == TT_MODE_STORE);
  if (rc) return rc;
  if (mode == TT_MODE_STORE && r->n > 0) {
    // TasksStoreManager.cs:136 — OrderBy(TaskCreatedOn), stable; ties keep row order (the sweep left ticks and rows
    // in the sort buffers)
    rc = tt::order_prefilled(t->stream, t->sort, t->task_id, r->d_rows, r->d_ids, r->n, &t->last_order);
    if (rc != 0) {
      result_release(r);
      return fail(TT_ERR_CUDA, "ordering by TaskCreatedOn failed: %s", cudaGetErrorString((cudaError_t)rc));
    }
  }
  *out = r;
  return TT_OK;
}

int32_t tt_sweep(tt_table* t, int32_t mode, int64_t today_local_ticks, int64_t run_at_utc_ticks, tt_result** out,
                 int64_t* n_marked) {
  if (!t || !out) return fail(TT_ERR_ARG, "null argument");
  *out = nullptr;
  if (n_marked) *n_marked = 0;
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  tt_sweep_stats st;
  tt_result* r = nullptr;
  int32_t rc = sweep_owned(t, mode, TT_SWEEP_MARK, today_local_ticks, run_at_utc_ticks, &r, &st);
  if (rc) return rc;
  *out = r;
  if (n_marked) *n_marked = st.n_marked;
  if (mode == TT_MODE_FAKE && r->n > 0 && run_at_utc_ticks / TT_TICKS_PER_DAY > today_local_ticks / TT_TICKS_PER_DAY - 1)
    return fail(TT_ERR_NOT_IMPLEMENTED,
                "FAKE mode: MarkOverdueTasks throws NotImplementedException (FakeTasksManager.cs:104); %lld tasks listed",
                (long long)r->n);
  return TT_OK;
}

int32_t tt_get_due_on_or_before(tt_table* t, int32_t mode, int64_t today_local_ticks, tt_result** out) {
  if (!t || !out) return fail(TT_ERR_ARG, "null argument");
  *out = nullptr;
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  tt_sweep_stats st;
  tt_result* r = nullptr;
  int32_t rc = sweep_owned(t, mode, TT_SWEEP_LIST_ONLY | TT_SWEEP_DUE_ON_OR_BEFORE, today_local_ticks, 0, &r, &st,
                           mode != TT_MODE_STORE, /*keys_too=*/mode == TT_MODE_STORE);
  if (rc) return rc;
  if (mode == TT_MODE_STORE && r->n > 0) {  // the store-backed list is ordered by TaskCreatedOn (TasksStoreManager.cs:136)
    rc = tt::order_prefilled(t->stream, t->sort, t->task_id, r->d_rows, r->d_ids, r->n, &t->last_order);
    if (rc != 0) {
      result_release(r);
      return fail(TT_ERR_CUDA, "ordering by TaskCreatedOn failed: %s", cudaGetErrorString((cudaError_t)rc));
    }
  }
  *out = r;
  return TT_OK;
}

int32_t tt_sweep_backlog(tt_table* t, int32_t mode, int64_t today_local_ticks, int64_t run_at_utc_ticks, tt_result** out,
                         int64_t* n_marked) {
  if (!t || !out) return fail(TT_ERR_ARG, "null argument");
  *out = nullptr;
  if (n_marked) *n_marked = 0;
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  tt_sweep_stats st;
  tt_result* r = nullptr;
  int32_t rc = sweep_owned(t, mode, TT_SWEEP_MARK | TT_SWEEP_DUE_ON_OR_BEFORE, today_local_ticks, run_at_utc_ticks, &r, &st);
  if (rc) return rc;
  *out = r;
  if (n_marked) *n_marked = st.n_marked;
  if (mode == TT_MODE_FAKE && r->n > 0)
    return fail(TT_ERR_NOT_IMPLEMENTED,
                "FAKE mode: MarkOverdueTasks throws NotImplementedException (FakeTasksManager.cs:104); %lld tasks listed",
                (long long)r->n);
  return TT_OK;
}

int32_t tt_sweep_device(tt_table* t, int32_t mode, int32_t flags, int64_t today_local_ticks,
                        int64_t run_at_utc_ticks, int64_t first_row, uint32_t* d_rows, void* d_ids, int64_t cap,
                        tt_sweep_stats* h_stats, void* stream) {
  if (!t || !h_stats) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  return sweep_run(t, mode, flags, today_local_ticks, run_at_utc_ticks, first_row, d_rows, (uint4*)d_ids, cap, h_stats,
                   stream ? (cudaStream_t)stream : t->stream);
}

int32_t tt_sweep_wait(tt_table* t, tt_sweep_stats* h_stats) {
  if (!t || !h_stats) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  if (!t->pending.active) return fail(TT_ERR_STATE, "no tick is pending on this table");
  return sweep_finish(t, h_stats);
}

int32_t tt_sweep_push(tt_table* t, int32_t mode, int32_t flags, int64_t today_local_ticks, int64_t run_at_utc_ticks,
                      uint32_t* d_rows, void* d_ids, int64_t cap, const tt_push_target* target, tt_sweep_stats* h_stats,
                      void* stream) {
  if (!t || !h_stats || !target) return fail(TT_ERR_ARG, "null argument");
  if (!target->off16 || !target->tile_count || !target->count) return fail(TT_ERR_ARG, "push target has a null buffer");
  if ((uintptr_t)target->off16 % 16 != 0) return fail(TT_ERR_ARG, "push target off16 must be 16-byte aligned");
  if (d_ids && !d_rows) return fail(TT_ERR_ARG, "d_ids without d_rows");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  tt::PushTarget push;
  push.off16 = static_cast<uint16_t*>(target->off16);
  push.tile_count = static_cast<uint32_t*>(target->tile_count);
  push.count = static_cast<unsigned long long*>(target->count);
  push.ids = nullptr;
  push.created = nullptr;
  push.tile_first = nullptr;
  push.created_as_key = 0;
  push.shape = 0;
  return sweep_run(t, mode, flags, today_local_ticks, run_at_utc_ticks, 0, d_rows, (uint4*)d_ids, cap, h_stats,
                   stream ? (cudaStream_t)stream : t->stream, &push);
}

// ---- peer-mapped buffers (CUDA IPC) ----------------------------------------------------------

int32_t tt_peer_alloc(int32_t device, uint64_t bytes, void** d_ptr, tt_peer_handle* handle) {
  if (!d_ptr || !handle || bytes == 0) return fail(TT_ERR_ARG, "bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) <= sizeof(tt_peer_handle), "handle too small");
  *d_ptr = nullptr;
  DeviceGuard g(device);
  void* p = nullptr;
  TT_CUDA(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    TT_CUDA(e);
  }
  memset(handle, 0, sizeof *handle);
  memcpy(handle->bytes, &h, sizeof h);
  *d_ptr = p;
  return TT_OK;
}

int32_t tt_peer_open(int32_t device, const tt_peer_handle* handle, void** d_ptr) {
  if (!d_ptr || !handle) return fail(TT_ERR_ARG, "bad argument");
  *d_ptr = nullptr;
  DeviceGuard g(device);
  cudaIpcMemHandle_t h;
  memcpy(&h, handle->bytes, sizeof h);
  void* p = nullptr;
  TT_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *d_ptr = p;
  return TT_OK;
}

int32_t tt_peer_close(int32_t device, void* d_ptr) {
  if (!d_ptr) return TT_OK;
  DeviceGuard g(device);
  TT_CUDA(cudaIpcCloseMemHandle(d_ptr));
  return TT_OK;
}

int32_t tt_peer_free(int32_t device, void* d_ptr) {
  if (!d_ptr) return TT_OK;
  DeviceGuard g(device);
  TT_CUDA(cudaFree(d_ptr));
  return TT_OK;
}

int32_t tt_mark_overdue(tt_table* t, int32_t mode, const uint8_t* ids16, int64_t n, int64_t* n_marked,
                        int64_t* n_missing) {
  if (!t || n < 0 || (n > 0 && !ids16)) return fail(TT_ERR_ARG, "bad argument");
  if (n_marked) *n_marked = 0;
  if (n_missing) *n_missing = 0;
  if (mode == TT_MODE_FAKE)
    return fail(TT_ERR_NOT_IMPLEMENTED, "FAKE mode: MarkOverdueTasks throws NotImplementedException (FakeTasksManager.cs:104)");
  if (mode != TT_MODE_STORE) return fail(TT_ERR_ARG, "mode must be TT_MODE_FAKE or TT_MODE_STORE");
  if (n == 0) return TT_OK;
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  if (t->n == 0) {
    if (n_missing) *n_missing = n;
    return TT_OK;
  }
  // the kernel binary-searches rows in TaskId order: a table whose rows are not (every CreateNewTask
  // with a fresh Guid breaks the order) gets its TaskId index rebuilt first
  int32_t rc = ensure_id_order(t);
  if (rc) return rc;
  uint4* d_ids = nullptr;
  TT_CUDA(cudaMalloc(&d_ids, n * 16));
  cudaError_t e = cudaMemcpyAsync(d_ids, ids16, n * 16, cudaMemcpyHostToDevice, t->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(t->d_small, 0, 16, t->stream);
  if (e == cudaSuccess) {
    tt::mark_by_id_kernel<<<grid_for(n, 128), 128, 0, t->stream>>>(d_ids, n, t->task_id, t->creator_ref,
                                                                   t->ids_sorted ? nullptr : t->id_order, t->n,
                                                                   t->overdue, t->d_small);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(t->h_small, t->d_small, 16, cudaMemcpyDeviceToHost, t->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
  cudaFree(d_ids);
  TT_CUDA(e);
  if (n_marked) *n_marked = (int64_t)t->h_small[0];
  if (n_missing) *n_missing = (int64_t)t->h_small[1];
  return TT_OK;
}

int32_t tt_get_by_creators(tt_table* t, const char* const* utf8, const int32_t* lens, int32_t nq, tt_result** out) {
  if (!t || !out || !utf8 || !lens || nq < 1) return fail(TT_ERR_ARG, "bad argument");
  if (nq > tt::kMaxQueries) return fail(TT_ERR_ARG, "at most %d creators per batch", tt::kMaxQueries);
  for (int32_t i = 0; i < nq; ++i) {
    out[i] = nullptr;
    if (lens[i] < 0 || (lens[i] > 0 && !utf8[i])) return fail(TT_ERR_ARG, "bad query %d", i);
  }
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  t->last_op = tt_op_stats{};
  t->last_op.queries = nq;

  // ---- the query set, built on the host: distinct strings, their hashes and the slot table ----
  std::vector<int32_t> uniq_of(nq);
  std::vector<int32_t> first_of;  // distinct query -> index of its first occurrence
  {
    std::unordered_map<std::string, int32_t> seen;
    for (int32_t i = 0; i < nq; ++i) {
      std::string key(utf8[i] ? utf8[i] : "", (size_t)lens[i]);
      auto it = seen.find(key);
      if (it == seen.end()) {
        it = seen.emplace(std::move(key), (int32_t)first_of.size()).first;
        first_of.push_back(i);
      }
      uniq_of[i] = it->second;
    }
  }
  const uint32_t nu = (uint32_t)first_of.size();
  uint32_t slots = 1024;
  // load factor 1/64 up to 64 queries; 4096 slots (48 KB) keep 4 CTAs per SM, the 8192-slot table
  // (2 CTAs per SM) only pays off for very large batches
  while (slots < 64 * nu && slots < 4096) slots <<= 1;
  if (nu > 512) slots = tt::kMaxSlots;
  std::vector<uint64_t> hash(nu), slot_key(slots, 0);
  std::vector<uint32_t> off(nu + 1, 0), slot_val(slots, tt::kSlotEmpty);
  std::vector<uint8_t> pool;
  for (uint32_t u = 0; u < nu; ++u) {
    const int32_t i = first_of[u];
    hash[u] = tt_hash64((const uint8_t*)utf8[i], (uint32_t)lens[i]);
    pool.insert(pool.end(), (const uint8_t*)utf8[i], (const uint8_t*)utf8[i] + lens[i]);
    off[u + 1] = (uint32_t)pool.size();
    uint32_t s = tt::slot_of(hash[u], slots);
    while (slot_val[s] != tt::kSlotEmpty) s = (s + 1) & (slots - 1);
    slot_key[s] = hash[u];
    slot_val[s] = (tt::tag_of(hash[u]) << 10) | u;
  }
  // one device buffer: [hash nu*8][slot_key slots*8][off (nu+1)*4][slot_val slots*4][pool]
  const size_t o_hash = 0, o_skey = o_hash + (size_t)nu * 8, o_off = o_skey + (size_t)slots * 8,
               o_sval = o_off + ((size_t)nu + 1) * 4, o_pool = o_sval + (size_t)slots * 4;
  const size_t qbytes = o_pool + std::max<size_t>(pool.size(), 1);
  std::vector<uint8_t> host(qbytes, 0);
  memcpy(&host[o_hash], hash.data(), (size_t)nu * 8);
  memcpy(&host[o_skey], slot_key.data(), (size_t)slots * 8);
  memcpy(&host[o_off], off.data(), ((size_t)nu + 1) * 4);
  memcpy(&host[o_sval], slot_val.data(), (size_t)slots * 4);
  if (!pool.empty()) memcpy(&host[o_pool], pool.data(), pool.size());
  if (qbytes > t->d_query_cap) {
    cudaFree(t->d_query);
    t->d_query = nullptr;
    t->d_query_cap = 0;
    const size_t want = std::max<size_t>(262144, qbytes);
    TT_CUDA(cudaMalloc(&t->d_query, want));
    t->d_query_cap = want;
  }
  if ((int64_t)nu > t->d_lists_queries) {
    cudaFree(t->d_lists);
    t->d_lists = nullptr;
    t->d_lists_queries = 0;
    const int64_t want = std::max<int64_t>(64, nu);
    TT_CUDA(cudaMalloc(&t->d_lists, (size_t)want * tt::kListCap * 4));
    t->d_lists_queries = want;
  }
  TT_CUDA(cudaMemcpyAsync(t->d_query, host.data(), qbytes, cudaMemcpyHostToDevice, t->stream));
  tt::QuerySet qs;
  qs.hash = reinterpret_cast<const uint64_t*>(t->d_query + o_hash);
  qs.slot_key = reinterpret_cast<const uint64_t*>(t->d_query + o_skey);
  qs.off = reinterpret_cast<const uint32_t*>(t->d_query + o_off);
  qs.slot_val = reinterpret_cast<const uint32_t*>(t->d_query + o_sval);
  qs.pool = t->d_query + o_pool;
  qs.slots = slots;
  qs.nq = nu;
  const size_t smem = nu == 1 ? 0 : (size_t)slots * 12;
  if (!t->creator_smem_opted) {  // a per-device attribute: kept with the table, which lives on one device
    TT_CUDA(cudaFuncSetAttribute(tt::creator_scan_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)tt::kMaxSlots * 12));
    TT_CUDA(cudaFuncSetAttribute(tt::order_lists_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 tt::kListOrderCap * 12));
    t->creator_smem_opted = true;
  }

  // ---- K4: one pass over the hash column for the whole batch.  First into per-query lists (one
  // CTA per query orders them afterwards); if some creator has more than kListCap tasks, once
  // more into a pair list that the general ordering path takes. ----
  const int64_t n_pairs = t->n_pad / 2;
  const int64_t per_cta = (int64_t)tt::kCreatorThreads * tt::kCreatorUnroll;
  const unsigned grid = (unsigned)std::max<int64_t>(
      1, std::min<int64_t>((n_pairs + per_cta - 1) / per_cta, (int64_t)t->sm_count * 8));
  auto scan = [&](const tt::MatchSink& sink) -> int32_t {
    TT_CUDA(cudaMemsetAsync(t->d_small, 0, 16, t->stream));
    TT_CUDA(cudaMemsetAsync(t->d_qcount, 0, (size_t)nu * 4, t->stream));
    TT_CUDA(cudaEventRecord(t->ev0, t->stream));
    if (nu == 1)
      tt::creator_scan_kernel<true><<<grid, tt::kCreatorThreads, 0, t->stream>>>(
          reinterpret_cast<const ulonglong2*>(t->creator_hash), t->n, n_pairs, qs, t->creator_ref, t->creator_off,
          t->creator_pool, sink);
    else
      tt::creator_scan_kernel<false><<<grid, tt::kCreatorThreads, smem, t->stream>>>(
          reinterpret_cast<const ulonglong2*>(t->creator_hash), t->n, n_pairs, qs, t->creator_ref, t->creator_off,
          t->creator_pool, sink);
    TT_CUDA(cudaGetLastError());
    TT_CUDA(cudaEventRecord(t->ev1, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->h_small, t->d_small, 16, cudaMemcpyDeviceToHost, t->stream));
    TT_CUDA(cudaMemcpyAsync(t->h_qcount, t->d_qcount, (size_t)nu * 4, cudaMemcpyDeviceToHost, t->stream));
    TT_CUDA(cudaStreamSynchronize(t->stream));
    float ms = 0;
    TT_CUDA(cudaEventElapsedTime(&ms, t->ev0, t->ev1));
    t->last_op.scan_ms += ms;
    t->last_op.launches += 1;
    return TT_OK;
  };

  int64_t found = 0;
  bool lists_ok = true;
  std::vector<uint32_t> count(nu, 0);
  if (t->n > 0) {
    tt::MatchSink sink{t->d_lists, t->d_qcount, nullptr, 0, t->d_small};
    int32_t rc = scan(sink);
    if (rc) return rc;
    found = (int64_t)t->h_small[0];
    t->last_op.candidates = (int64_t)t->h_small[1];
    for (uint32_t u = 0; u < nu; ++u) {
      count[u] = t->h_qcount[u];
      lists_ok = lists_ok && count[u] <= tt::kListCap;
    }
    if (!lists_ok) {
      if (found > t->d_pairs_cap) {
        cudaFree(t->d_pairs);
        t->d_pairs = nullptr;
        t->d_pairs_cap = 0;
        TT_CUDA(cudaMalloc(&t->d_pairs, (size_t)found * 8));
        t->d_pairs_cap = found;
      }
      tt::MatchSink psink{nullptr, t->d_qcount, t->d_pairs, t->d_pairs_cap, t->d_small};
      rc = scan(psink);
      if (rc) return rc;
      if ((int64_t)t->h_small[0] != found) return fail(TT_ERR_STATE, "table changed during a creator query");
    }
  }
  t->last_op.matches = found;

  // ---- K5: order every query's matches by TaskCreatedOn descending (stable), gather TaskIds ----
  tt_result* whole = nullptr;
  int32_t rc = result_alloc(t, std::max<int64_t>(found, 1), &whole);
  if (rc) return rc;
  whole->n = found;
  if (found > 0) {
    TT_CUDA(cudaEventRecord(t->ev0, t->stream));
    if (lists_ok) {
      tt::order_lists_kernel<<<nu, tt::kListOrderThreads, tt::kListOrderCap * 12, t->stream>>>(t->d_lists, t->d_qcount, t->created_raw,
                                                                        t->task_id, 1, whole->d_rows, whole->d_ids);
      cudaError_t ce = cudaGetLastError();
      if (ce != cudaSuccess) {
        result_release(whole);
        return fail(TT_ERR_CUDA, "order_lists_kernel: %s", cudaGetErrorString(ce));
      }
    } else {
      cudaMemsetAsync(t->d_qcount, 0, (size_t)nu * 4, t->stream);
      int e = tt::order_pairs(t->stream, t->sort, t->d_pairs, found, t->created_raw, t->task_id, /*descending=*/true,
                              whole->d_rows, whole->d_ids, t->d_qcount);
      if (e != 0) {
        result_release(whole);
        cudaGetLastError();
        return fail(TT_ERR_CUDA, "order_pairs: %s", cudaGetErrorString((cudaError_t)e));
      }
    }
    cudaEventRecord(t->ev1, t->stream);
    cudaError_t ce = cudaStreamSynchronize(t->stream);
    if (ce != cudaSuccess) {
      result_release(whole);
      cudaGetLastError();
      return fail(TT_ERR_CUDA, "creator result ordering: %s", cudaGetErrorString(ce));
    }
    cudaEventElapsedTime(&t->last_op.order_ms, t->ev0, t->ev1);
    t->last_op.launches += 1;
  }
  std::vector<int64_t> first(nu + 1, 0);
  for (uint32_t u = 0; u < nu; ++u) first[u + 1] = first[u] + count[u];
  for (int32_t i = 0; i < nq; ++i) {
    const int32_t u = uniq_of[i];
    out[i] = result_slice(whole, first[u], count[u]);
    if (!out[i]) {
      for (int32_t k = 0; k < i; ++k) {
        result_release(out[k]);
        out[k] = nullptr;
      }
      result_release(whole);
      return fail(TT_ERR_OOM, "out of host memory");
    }
  }
  result_release(whole);  // the slices keep the block alive
  return TT_OK;
}

// ITasksManager.GetTasksByCreator for one request thread.  The reference's manager is a singleton shared by all
// request threads (Api/Program.cs:13), so many of these calls arrive at once: they are combined (see
// tt_table::CreatorRequest) instead of queueing behind each other for a table pass each.
int32_t tt_get_by_creator(tt_table* t, const char* utf8, int32_t len, tt_result** out) {
  if (!t || !out) return fail(TT_ERR_ARG, "bad argument");
  *out = nullptr;
  if (len < 0 || (len > 0 && !utf8)) return fail(TT_ERR_ARG, "bad query");
  tt_table::CreatorRequest me;
  me.utf8 = utf8 ? utf8 : "";
  me.len = len;
  std::unique_lock<std::mutex> lk(t->cq_mu);
  t->cq.push_back(&me);
  while (!me.done) {
    if (t->cq_busy) {
      t->cq_cv.wait(lk);
      continue;
    }
    // nobody is scanning: this thread takes what is queued (its own request included unless the queue is
    // longer than a batch) and answers all of it with one pass
    t->cq_busy = true;
    const size_t take = std::min<size_t>(t->cq.size(), (size_t)tt::kMaxQueries);
    std::vector<tt_table::CreatorRequest*> batch(t->cq.begin(), t->cq.begin() + take);
    t->cq.erase(t->cq.begin(), t->cq.begin() + take);
    lk.unlock();
    std::vector<const char*> q(take);
    std::vector<int32_t> l(take);
    std::vector<tt_result*> res(take, nullptr);
    for (size_t i = 0; i < take; ++i) {
      q[i] = batch[i]->utf8;
      l[i] = batch[i]->len;
    }
    const int32_t rc = tt_get_by_creators(t, q.data(), l.data(), (int32_t)take, res.data());
    const std::string err = rc ? g_err : std::string();
    lk.lock();
    for (size_t i = 0; i < take; ++i) {
      batch[i]->out = res[i];
      batch[i]->rc = rc;
      batch[i]->err = err;
      batch[i]->done = true;
    }
    t->cq_batches += 1;
    t->cq_requests += (int64_t)take;
    t->cq_busy = false;
    t->cq_cv.notify_all();
  }
  lk.unlock();
  if (me.rc) {
    g_err = me.err;
    return me.rc;
  }
  *out = me.out;
  return TT_OK;
}

// How often concurrent GetTasksByCreator calls shared a table pass: *requests answered by *passes passes.
int32_t tt_creator_combining_stats(tt_table* t, int64_t* passes, int64_t* requests) {
  if (!t) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->cq_mu);
  if (passes) *passes = t->cq_batches;
  if (requests) *requests = t->cq_requests;
  return TT_OK;
}

int32_t tt_last_order_info(tt_table* t, int32_t* key_bits, int32_t* passes, int64_t* groups_fixed, int32_t* fell_back) {
  if (!t) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->mu);
  if (key_bits) *key_bits = t->last_order.bits;
  if (passes) *passes = t->last_order.passes;
  if (groups_fixed) *groups_fixed = (int64_t)t->last_order.fixed;
  if (fell_back) *fell_back = t->last_order.fell_back ? 1 : 0;
  return TT_OK;
}

int32_t tt_last_op_stats(tt_table* t, tt_op_stats* out) {
  if (!t || !out) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->mu);
  *out = t->last_op;
  return TT_OK;
}

// ---- results --------------------------------------------------------------------------------

int32_t tt_result_count(const tt_result* r, int64_t* n) {
  if (!r || !n) return fail(TT_ERR_ARG, "null argument");
  *n = r->n;
  return TT_OK;
}

int32_t tt_result_rows(tt_result* r, uint32_t* out) {
  if (!r || (r->n && !out)) return fail(TT_ERR_ARG, "null argument");
  DeviceGuard g(r->core->device);
  if (r->n) TT_CUDA(cudaMemcpy(out, r->d_rows, r->n * 4, cudaMemcpyDeviceToHost));
  return TT_OK;
}

int32_t tt_result_ids(tt_result* r, uint8_t* out16) {
  if (!r || (r->n && !out16)) return fail(TT_ERR_ARG, "null argument");
  DeviceGuard g(r->core->device);
  if (r->n) TT_CUDA(cudaMemcpy(out16, r->d_ids, r->n * 16, cudaMemcpyDeviceToHost));
  return TT_OK;
}

int32_t tt_result_field(tt_result* r, int32_t field, void* out) {
  if (!r || (r->n && !out)) return fail(TT_ERR_ARG, "null argument");
  if (r->n == 0) return TT_OK;
  if (!r->core->alive) return fail(TT_ERR_STATE, "the table of this result was destroyed");
  tt_table* t = r->t;
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  size_t width = (field == TT_FIELD_CREATED_RAW || field == TT_FIELD_DUE_RAW) ? 8 : (field == TT_FIELD_CREATOR_REF ? 4 : 1);
  void* d_out = nullptr;
  TT_CUDA(cudaMalloc(&d_out, r->n * width));
  const unsigned grid = grid_for(r->n, 256);
  switch (field) {
    case TT_FIELD_CREATED_RAW:
      gather_field_kernel<uint64_t><<<grid, 256, 0, t->stream>>>(r->d_rows, t->created_raw, (uint64_t*)d_out, r->n);
      break;
    case TT_FIELD_DUE_RAW:
      gather_field_kernel<uint64_t><<<grid, 256, 0, t->stream>>>(r->d_rows, t->due_raw, (uint64_t*)d_out, r->n);
      break;
    case TT_FIELD_CREATOR_REF:
      gather_field_kernel<uint32_t><<<grid, 256, 0, t->stream>>>(r->d_rows, t->creator_ref, (uint32_t*)d_out, r->n);
      break;
    case TT_FIELD_COMPLETED:
      gather_bit_kernel<<<grid, 256, 0, t->stream>>>(r->d_rows, t->completed, (uint8_t*)d_out, r->n);
      break;
    case TT_FIELD_OVERDUE:
      gather_bit_kernel<<<grid, 256, 0, t->stream>>>(r->d_rows, t->overdue, (uint8_t*)d_out, r->n);
      break;
    default:
      cudaFree(d_out);
      return fail(TT_ERR_ARG, "unknown field %d", field);
  }
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, r->n * width, cudaMemcpyDeviceToHost, t->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
  cudaFree(d_out);
  TT_CUDA(e);
  return TT_OK;
}

int32_t tt_result_free(tt_result* r) {
  if (!r) return TT_OK;
  DeviceGuard g(r->core->device);
  result_release(r);
  return TT_OK;
}

int32_t tt_result_digest(tt_result* r, uint64_t* xor_id_hi, uint64_t* xor_id_lo, uint64_t* sum_rows) {
  if (!r) return fail(TT_ERR_ARG, "null argument");
  if (!r->core->alive) return fail(TT_ERR_STATE, "the table of this result was destroyed");
  tt_table* t = r->t;
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  TT_CUDA(cudaMemsetAsync(t->d_small, 0, 24, t->stream));
  if (r->n) {
    digest_kernel<<<(unsigned)std::min<int64_t>(grid_for(r->n, 256), 4 * t->sm_count), 256, 0, t->stream>>>(
        r->d_rows, r->d_ids, r->n, t->d_small);
    TT_CUDA(cudaGetLastError());
  }
  TT_CUDA(cudaMemcpyAsync(t->h_small, t->d_small, 24, cudaMemcpyDeviceToHost, t->stream));
  TT_CUDA(cudaStreamSynchronize(t->stream));
  if (xor_id_hi) *xor_id_hi = t->h_small[0];
  if (xor_id_lo) *xor_id_lo = t->h_small[1];
  if (sum_rows) *sum_rows = t->h_small[2];
  return TT_OK;
}

// ---- state access -----------------------------------------------------------------------------

int32_t tt_snapshot_flags(tt_table* t, uint8_t* completed_bits, uint8_t* overdue_bits) {
  if (!t) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  size_t bytes = (size_t)(t->n + 7) / 8;
  if (bytes == 0) return TT_OK;
  if (completed_bits) TT_CUDA(cudaMemcpyAsync(completed_bits, t->completed, bytes, cudaMemcpyDefault, t->stream));
  if (overdue_bits) TT_CUDA(cudaMemcpyAsync(overdue_bits, t->overdue, bytes, cudaMemcpyDefault, t->stream));
  TT_CUDA(cudaStreamSynchronize(t->stream));
  return TT_OK;
}

int32_t tt_restore_flags(tt_table* t, const uint8_t* completed_bits, const uint8_t* overdue_bits) {
  if (!t) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  size_t bytes = (size_t)(t->n + 7) / 8;
  if (bytes == 0) return TT_OK;
  if (completed_bits) TT_CUDA(cudaMemcpyAsync(t->completed, completed_bits, bytes, cudaMemcpyDefault, t->stream));
  if (overdue_bits) TT_CUDA(cudaMemcpyAsync(t->overdue, overdue_bits, bytes, cudaMemcpyDefault, t->stream));
  TT_CUDA(cudaStreamSynchronize(t->stream));
  return TT_OK;
}

int32_t tt_count_flags(tt_table* t, int64_t* n_completed, int64_t* n_overdue) {
  if (!t) return fail(TT_ERR_ARG, "null argument");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  TT_CUDA(cudaMemsetAsync(t->d_small, 0, 16, t->stream));
  int64_t words = t->n_pad / 32;
  if (words) {
    count_flags_kernel<<<(unsigned)std::min<int64_t>(grid_for(words, 256), 4 * t->sm_count), 256, 0, t->stream>>>(
        t->completed, t->overdue, words, t->d_small);
    TT_CUDA(cudaGetLastError());
  }
  TT_CUDA(cudaMemcpyAsync(t->h_small, t->d_small, 16, cudaMemcpyDeviceToHost, t->stream));
  TT_CUDA(cudaStreamSynchronize(t->stream));
  if (n_completed) *n_completed = (int64_t)t->h_small[0];
  if (n_overdue) *n_overdue = (int64_t)t->h_small[1];
  return TT_OK;
}

int32_t tt_get_device_view(tt_table* t, tt_device_view* out) {
  if (!t || !out) return fail(TT_ERR_ARG, "null argument");
  out->n = t->n;
  out->n_padded = t->n_pad;
  out->due_code = t->due_code;
  out->completed = t->completed;
  out->overdue = t->overdue;
  out->creator_hash = t->creator_hash;
  out->task_id = t->task_id;
  out->created_raw = t->created_raw;
  out->device = t->device;
  return TT_OK;
}

}  // extern "C"

#include "api_point.inc"

#include "api_snapshot.inc"

// ---- compact hit lists for the cross-shard exchange ------------------------------------------------

extern "C" {

int32_t tt_pack_hits(tt_table* t, const uint32_t* d_rows, int64_t n, uint16_t* d_off16, uint32_t* d_tile_count,
                     int64_t* n_tiles, void* stream_) {
  if (!t || !n_tiles || n < 0 || (n > 0 && (!d_rows || !d_off16)) || !d_tile_count) return fail(TT_ERR_ARG, "bad argument");
  std::lock_guard<std::mutex> lk(t->mu);
  DeviceGuard g(t->device);
  cudaStream_t stream = stream_ ? (cudaStream_t)stream_ : t->stream;
  *n_tiles = t->n_pad / TT_TILE_ROWS;
  if (t->n_pad == 0) return TT_OK;
  // per-tile counts as the last complete sweep recorded them (tile_words[1])
  TT_CUDA(cudaMemcpyAsync(d_tile_count, t->tile_words + t->n_tiles, (size_t)*n_tiles * 4, cudaMemcpyDeviceToDevice, stream));
  if (n) {
    tt::pack_hits16_kernel<<<grid_for(n, 256), 256, 0, stream>>>(d_rows, n, d_off16);
    TT_CUDA(cudaGetLastError());
  }
  return TT_OK;
}

int32_t tt_expand_hits(int32_t device, const uint32_t* d_tile_count, int64_t n_tiles, const uint16_t* d_off16,
                       uint32_t* d_scratch, uint32_t* d_rows_out, void* stream_) {
  if (n_tiles < 0 || (n_tiles > 0 && (!d_tile_count || !d_scratch || !d_rows_out))) return fail(TT_ERR_ARG, "bad argument");
  if (n_tiles == 0) return TT_OK;
  DeviceGuard g(device);
  cudaStream_t stream = (cudaStream_t)stream_;
  const uint32_t nt = (uint32_t)n_tiles;
  const uint32_t groups = (nt + tt::kTilesPerGroup - 1) / tt::kTilesPerGroup;
  uint32_t* tile_first = d_scratch;        // [n_tiles]
  uint32_t* group_sum = d_scratch + nt;    // [groups]
  tt::group_reduce_kernel<<<groups, tt::kTilesPerGroup, 0, stream>>>(d_tile_count, nt, group_sum);
  tt::tile_offsets_kernel<<<groups, tt::kTilesPerGroup, 0, stream>>>(d_tile_count, group_sum, nt, tile_first);
  tt::expand_hits16_kernel<<<(nt + 7) / 8, 256, 0, stream>>>(d_tile_count, tile_first, nt, d_off16, d_rows_out);
  TT_CUDA(cudaGetLastError());
  return TT_OK;
}

}  // extern "C"

#include "api_json.inc"

#include "api_ingest.inc"

#include "api_tick.inc"

#include "api_shards.inc"

// ---- the sweep with its working memory in a fixed device array ---------------------------------
//
// For callers that cannot allocate (see include/tasktable.h).  Everything tt_sweep_device keeps
// in table-owned, growable buffers lives here in one file-scope array:
//   [0, 256)      SweepCtl
//   [256, 512)    group sums (TT_STATIC_MAX_ROWS / 8192 / 1024 = 2 words used)
//   [512, ...)    tile_slot, tile_count, tile_first: 3 x 2048 words
//   [25088, ...)  intermediate hit list: TT_STATIC_MAX_HITS words
// In host code the array's name is only a handle for cudaGetSymbolAddress — it is not a device
// address — so the address is looked up once per device and handed to the kernels as a pointer.
namespace {

constexpr size_t kStaticTiles = TT_STATIC_MAX_ROWS / TT_TILE_ROWS;
constexpr size_t kStaticGroupOff = 256;
constexpr size_t kStaticTileOff = 512;
constexpr size_t kStaticListOff = kStaticTileOff + 3 * kStaticTiles * 4;
constexpr size_t kStaticBytes = kStaticListOff + (size_t)TT_STATIC_MAX_HITS * 4;
static_assert(kStaticListOff % 256 == 0, "hit list must stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// Device code may name the array directly.  Puts the control block and the group sums back
// into their start state; the per-tile words and the list are fully rewritten by every sweep.
__global__ void static_scratch_reset_kernel() {
  tt::SweepCtl* ctl = reinterpret_cast<tt::SweepCtl*>(g_static_scratch);
  uint32_t* groups = reinterpret_cast<uint32_t*>(g_static_scratch + kStaticGroupOff);
  if (threadIdx.x == 0) {
    ctl->tile_counter = 0;
    ctl->pad = 0;
    ctl->cursor = 0;
    ctl->stored = ~0ull;
  }
  if (threadIdx.x < (256 / 4)) groups[threadIdx.x] = 0;
}

constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_addr[kMaxDevices] = {};
int g_static_sms[kMaxDevices] = {};

}  // namespace

extern "C" int32_t tt_sweep_static(const tt_device_view* cols, int32_t mode, int32_t flags, int64_t today_local_ticks,
                                   int64_t run_at_utc_ticks, uint32_t* d_rows, void* d_ids, int64_t cap,
                                   tt_sweep_stats* h_stats, void* stream_) {
  if (!cols || !h_stats) return fail(TT_ERR_ARG, "null argument");
  memset(h_stats, 0, sizeof *h_stats);
  if (mode != TT_MODE_FAKE && mode != TT_MODE_STORE) return fail(TT_ERR_ARG, "mode must be TT_MODE_FAKE or TT_MODE_STORE");
  uint32_t yesterday = 0;
  int32_t rc = check_clock(today_local_ticks, &yesterday);
  if (rc) return rc;
  if (cols->n < 0 || cols->n_padded < cols->n || cols->n_padded % TT_TILE_ROWS != 0)
    return fail(TT_ERR_ARG, "columns must be padded to a multiple of %d rows", TT_TILE_ROWS);
  if (cols->n_padded > TT_STATIC_MAX_ROWS)
    return fail(TT_ERR_ARG, "tt_sweep_static handles at most %d rows (got %lld); use tt_sweep_device", TT_STATIC_MAX_ROWS,
                (long long)cols->n_padded);
  if (cap < 0 || cap > TT_STATIC_MAX_HITS)
    return fail(TT_ERR_ARG, "tt_sweep_static handles at most %d hits (cap %lld); use tt_sweep_device", TT_STATIC_MAX_HITS,
                (long long)cap);
  if (cols->device < 0 || cols->device >= kMaxDevices) return fail(TT_ERR_ARG, "device %d out of range", cols->device);
  if (cols->n > 0 && (!cols->due_code || !cols->completed || !cols->overdue || (d_ids && !cols->task_id)))
    return fail(TT_ERR_ARG, "null column");
  if (cols->n == 0) {
    h_stats->resume_row = 0;
    return TT_OK;
  }
  int n_dev = 0;
  TT_CUDA(cudaGetDeviceCount(&n_dev));
  if (cols->device >= n_dev) return fail(TT_ERR_ARG, "device %d not in [0, %d)", cols->device, n_dev);
  const bool mark = (flags & TT_SWEEP_MARK) != 0;
  const bool range = (flags & TT_SWEEP_DUE_ON_OR_BEFORE) != 0;
  const bool survive = run_at_utc_ticks / TT_TICKS_PER_DAY > (int64_t)yesterday;
  if (range && mark && !survive)  // the same refusal as tt_sweep_device (sweep_run)
    return fail(TT_ERR_ARG, "range sweep: run_at_utc_ticks must lie after yesterday (runAt.Date > TaskDueDate.Date would differ per row)");
  std::lock_guard<std::mutex> lk(g_static_mu);  // one sweep at a time owns the fixed array
  DeviceGuard g(cols->device);
  const int dev = cols->device;
  if (!g_static_addr[dev]) {
    // set-up, once per device, with that device current
    void* addr = nullptr;
    TT_CUDA(cudaGetSymbolAddress(&addr, g_static_scratch));
    int sms = 0;
    TT_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    g_static_sms[dev] = sms;
    g_static_addr[dev] = static_cast<unsigned char*>(addr);
  }
  unsigned char* base = g_static_addr[dev];
  cudaStream_t stream = (cudaStream_t)stream_;
  const bool listing = d_rows != nullptr;
  const uint32_t num_tiles = (uint32_t)(cols->n_padded / TT_TILE_ROWS);

  tt::SweepParams p;
  p.due_code = static_cast<const uint32_t*>(cols->due_code);
  p.completed = static_cast<const uint32_t*>(cols->completed);
  p.overdue = static_cast<uint32_t*>(cols->overdue);
  p.list = listing ? reinterpret_cast<uint32_t*>(base + kStaticListOff) : nullptr;
  p.tile_slot = reinterpret_cast<uint32_t*>(base + kStaticTileOff);
  p.tile_count = p.tile_slot + kStaticTiles;
  uint32_t* tile_first = p.tile_slot + 2 * kStaticTiles;
  p.group_sum = reinterpret_cast<uint32_t*>(base + kStaticGroupOff);
  p.sub_sum = nullptr;
  p.ctl = reinterpret_cast<tt::SweepCtl*>(base);
  p.cap = listing ? (unsigned long long)cap : ~0ull;
  p.first_tile = 0;
  p.num_tiles = num_tiles;
  p.mask = mode == TT_MODE_FAKE ? TT_MASK_FAKE : TT_MASK_STORE;
  p.yesterday = yesterday;
  p.mark_enable = (mark && survive && mode == TT_MODE_STORE) ? 1u : 0u;

  // Once a kernel has been queued the shared array is in use on the caller's stream: every exit
  // below first waits for the stream, so the lock is never released while kernels may still touch it.
  cudaError_t e = cudaSuccess;
  const char* what = "";
  auto step = [&](const char* name) {
    if (e == cudaSuccess) {
      e = cudaGetLastError();
      what = name;
    }
    return e == cudaSuccess;
  };
  static_scratch_reset_kernel<<<1, 64, 0, stream>>>();
  constexpr int kT = tt::kSweepThreads, kC = tt::kSweepCtasPerSm;
  const unsigned grid = (unsigned)std::min<int64_t>(num_tiles, (int64_t)g_static_sms[dev] * kC);
  if (step("reset")) {
    if (range) {
      if (mode == TT_MODE_STORE) {
        if (p.mark_enable) tt::sweep_kernel<true, true, kT, kC, true><<<grid, kT, 0, stream>>>(p);
        else tt::sweep_kernel<true, false, kT, kC, true><<<grid, kT, 0, stream>>>(p);
      } else {
        tt::sweep_kernel<false, false, kT, kC, true><<<grid, kT, 0, stream>>>(p);
      }
    } else if (mode == TT_MODE_STORE) {
      if (p.mark_enable) tt::sweep_kernel<true, true, kT, kC><<<grid, kT, 0, stream>>>(p);
      else tt::sweep_kernel<true, false, kT, kC><<<grid, kT, 0, stream>>>(p);
    } else {
      tt::sweep_kernel<false, false, kT, kC><<<grid, kT, 0, stream>>>(p);
    }
  }
  h_stats->launches = 2;
  if (step("sweep_kernel") && listing) {
    const unsigned n_groups = (num_tiles + tt::kTilesPerGroup - 1) / tt::kTilesPerGroup;
    tt::tile_offsets_kernel<<<n_groups, tt::kTilesPerGroup, 0, stream>>>(p.tile_count, p.group_sum, num_tiles, tile_first);
    if (step("tile_offsets_kernel")) {
      const int64_t per_cta = (int64_t)tt::kPlaceThreads * tt::kPlaceUnroll;
      const unsigned pgrid = (unsigned)std::max<int64_t>(
          1, std::min<int64_t>((cap + per_cta - 1) / per_cta, (int64_t)g_static_sms[dev] * 8));
      tt::place_hits_kernel<TT_TILE_ROWS><<<pgrid, tt::kPlaceThreads, 0, stream>>>(
          p.ctl, p.list, p.tile_slot, tile_first, 0, static_cast<const uint4*>(cols->task_id), d_rows,
          static_cast<uint4*>(d_ids));
      step("place_hits_kernel");
    }
    h_stats->launches = 4;
  }
  tt::SweepCtl done;  // read back into the caller's stack: no pinned buffer is created for it
  done.cursor = 0;
  done.stored = ~0ull;
  if (e == cudaSuccess) {
    e = cudaMemcpyAsync(&done, base, sizeof done, cudaMemcpyDeviceToHost, stream);
    what = "copy of the control block";
  }
  const cudaError_t se = cudaStreamSynchronize(stream);
  if (e == cudaSuccess && se != cudaSuccess) {
    e = se;
    what = "cudaStreamSynchronize";
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(TT_ERR_CUDA, "tt_sweep_static, %s: %s", what, cudaGetErrorString(e));
  }
  h_stats->n_listed = (int64_t)done.cursor;
  const bool complete = done.stored == ~0ull;
  h_stats->n_stored = listing ? (complete ? h_stats->n_listed : (int64_t)done.stored) : 0;
  h_stats->resume_row = complete ? cols->n : 0;
  h_stats->n_marked = p.mark_enable ? (listing ? h_stats->n_stored : h_stats->n_listed) : 0;
  if (!complete)
    return fail(TT_ERR_CAPACITY, "%lld hits do not fit in %lld (%lld stored%s)", (long long)h_stats->n_listed,
                (long long)cap, (long long)h_stats->n_stored, h_stats->n_marked ? " and marked; call again for the rest" : "");
  return TT_OK;
}
