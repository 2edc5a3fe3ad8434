// sweep_lab.cu — development harness, not part of the product: times the library's sweep kernel at
// several grid sizes on synthetic columns (5 % matches, 50 % completed) and checks whether K3 (random
// TaskId gathers) can overlap K1 (streaming) on two streams.  Earlier versions of this file also
// carried the in-order compaction variants (decoupled look-back, published-count scan) whose timings
// are recorded in profiles/r1_summary.md; they were removed together with those kernels.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <algorithm>
#include "../aca-dotnet-workshop_b200/csrc/sweep.cuh"

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
using namespace tt;

__global__ void fill(uint32_t* due, uint32_t* comp, uint32_t* over, int64_t n, uint32_t yday) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t h = tt_mix64(i * 0x9E3779B97F4A7C15ull + 12345);
  due[i] = (h % 100 < 5) ? yday : yday + 1 + (uint32_t)(h >> 40) % 300;
  uint32_t c = __ballot_sync(0xFFFFFFFFu, (h >> 20) & 1);
  if ((threadIdx.x & 31) == 0) { comp[i >> 5] = c; over[i >> 5] = 0; }
}

int main(int argc, char** argv) {
  const int64_t N = argc > 1 ? atoll(argv[1]) : (1ll << 30);
  const int64_t NP = (N + 32767) / 32768 * 32768;
  const uint32_t yday = 739924;
  uint32_t *due, *comp, *over, *rows, *status, *ubase, *tcount, *gsum; SweepCtl* ctl;
  CK(cudaMalloc(&due, NP * 4)); CK(cudaMalloc(&comp, NP / 8)); CK(cudaMalloc(&over, NP / 8));
  CK(cudaMalloc(&rows, NP / 16 * 4)); CK(cudaMalloc(&status, (NP / 8192 + 4) * 4));
  CK(cudaMalloc(&ubase, NP / 8192 * 4)); CK(cudaMalloc(&tcount, NP / 8192 * 4));
  CK(cudaMalloc(&ctl, sizeof(SweepCtl)));
  CK(cudaMemset(due, 0xFF, NP * 4));
  fill<<<(unsigned)((NP + 255) / 256), 256>>>(due, comp, over, N, yday);
  CK(cudaDeviceSynchronize());
  CK(cudaMalloc(&gsum, 4096));
  SweepParams p; p.due_code = due; p.completed = comp; p.overdue = over; p.list = rows; p.tile_slot = ubase; p.tile_count = tcount; p.group_sum = gsum;
  p.ctl = ctl; p.cap = NP / 16; p.first_tile = 0; p.num_tiles = (uint32_t)(NP / 8192); p.mask = TT_MASK_STORE; p.yesterday = yday; p.mark_enable = 1;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  auto run = [&](const char* name, int grid_mult, auto launch) {
    float best = 1e9f; unsigned long long total = 0;
    for (int r = 0; r < 4; ++r) {
      cudaMemset(over, 0, NP / 8);
      SweepCtl init; init.tile_counter = 0; init.pad = 0; init.cursor = 0; init.stored = ~0ull;
      cudaMemset(gsum, 0, 4096);
      cudaMemcpy(ctl, &init, sizeof init, cudaMemcpyHostToDevice);
      cudaMemset(status, 0, (NP / 8192 + 4) * 4);
      cudaEventRecord(a); launch(148 * grid_mult); cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b); best = std::min(best, ms);
      SweepCtl h; cudaMemcpy(&h, ctl, sizeof h, cudaMemcpyDeviceToHost);
      total = h.cursor;
    }
    printf("%-40s %8.3f ms  %7.1f GB/s  hits=%llu  err=%s\n", name, best, N * 4.25 / best / 1e6, total, cudaGetErrorString(cudaGetLastError()));
  };
  p.num_tiles = (uint32_t)(NP / 8192);
  run("library sweep 256thr x4/SM", 4, [&](int g) { sweep_kernel<true, true, 256, 4><<<g, 256>>>(p); });
  run("library sweep 256thr x3/SM", 3, [&](int g) { sweep_kernel<true, true, 256, 4><<<g, 256>>>(p); });
  run("library sweep 256thr x2/SM", 2, [&](int g) { sweep_kernel<true, true, 256, 4><<<g, 256>>>(p); });
  // ---- can K3 (random TaskId gathers) overlap K1 (streaming) on the same DRAM? ----
  {
    uint4* ids; uint4* out_ids; uint32_t *out_rows, *tfirst, *over2;
    CK(cudaMalloc(&ids, NP * 16)); CK(cudaMalloc(&out_ids, NP / 16 * 16)); CK(cudaMalloc(&out_rows, NP / 16 * 4));
    CK(cudaMalloc(&tfirst, NP / 8192 * 4)); CK(cudaMalloc(&over2, NP / 8));
    CK(cudaMemset(ids, 7, NP * 16));
    cudaStream_t s1, s2; cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking); cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking);
    // one K1 run to produce the list + tile words, then K2
    auto reset = [&](uint32_t* ov) {
      cudaMemset(ov, 0, NP / 8);
      SweepCtl init; init.tile_counter = 0; init.pad = 0; init.cursor = 0; init.stored = ~0ull;
      cudaMemcpy(ctl, &init, sizeof init, cudaMemcpyHostToDevice); cudaMemset(gsum, 0, 4096);
    };
    reset(over);
    sweep_kernel<true, true, 256, 4><<<592, 256>>>(p);
    unsigned ng = (p.num_tiles + kTilesPerGroup - 1) / kTilesPerGroup;
    tile_offsets_kernel<<<ng, kTilesPerGroup>>>(p.tile_count, p.group_sum, p.num_tiles, tfirst);
    CK(cudaDeviceSynchronize());
    // a second control block + list for the concurrent K1 so that the K3 inputs stay intact
    SweepCtl* ctl2; uint32_t *rows2, *ubase2, *tcount2, *gsum2;
    CK(cudaMalloc(&ctl2, sizeof(SweepCtl))); CK(cudaMalloc(&rows2, NP / 16 * 4)); CK(cudaMalloc(&ubase2, NP / 8192 * 4));
    CK(cudaMalloc(&tcount2, NP / 8192 * 4)); CK(cudaMalloc(&gsum2, 4096));
    SweepParams p2 = p; p2.ctl = ctl2; p2.list = rows2; p2.tile_slot = ubase2; p2.tile_count = tcount2; p2.group_sum = gsum2; p2.overdue = over2;
    cudaEvent_t e0, e1, e2, e3; cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&e2); cudaEventCreate(&e3);
    for (int mode = 0; mode < 3; ++mode) {
      float best = 1e9f, bk1 = 0, bk3 = 0;
      for (int r = 0; r < 4; ++r) {
        cudaMemset(over2, 0, NP / 8);
        SweepCtl init; init.tile_counter = 0; init.pad = 0; init.cursor = 0; init.stored = ~0ull;
        cudaMemcpy(ctl2, &init, sizeof init, cudaMemcpyHostToDevice); cudaMemset(gsum2, 0, 4096);
        cudaDeviceSynchronize();
        cudaEventRecord(e0, s1); cudaEventRecord(e2, s2);
        if (mode != 1) sweep_kernel<true, true, 256, 4><<<592, 256, 0, s1>>>(p2);
        if (mode != 0) place_hits_kernel<8192><<<148 * 8, 256, 0, s2>>>(ctl, rows, p.tile_slot, tfirst, 0, ids, out_rows, out_ids);
        cudaEventRecord(e1, s1); cudaEventRecord(e3, s2);
        cudaDeviceSynchronize();
        float k1, k3; cudaEventElapsedTime(&k1, e0, e1); cudaEventElapsedTime(&k3, e2, e3);
        float tot = std::max(k1, k3);
        if (tot < best) { best = tot; bk1 = k1; bk3 = k3; }
      }
      printf("%-40s total %8.3f ms (K1 %.3f, K3 %.3f)\n", mode == 0 ? "K1 alone" : mode == 1 ? "K3 alone" : "K1 || K3 on two streams", best, bk1, bk3);
    }
    // ---- the same question with room left on every SM: K1 at 3 or 2 CTAs per SM (it needs the whole
    // register file at 4) next to the tile-ordered K3g, launched in either order ----
    PushTarget none = {nullptr, nullptr, nullptr};
    const unsigned ggrid = (p.num_tiles + kGroupTiles - 1) / kGroupTiles;
    for (int k1ctas = 4; k1ctas >= 2; --k1ctas) {
      for (int order = 0; order < 2; ++order) {
        float best = 1e9f, bk1 = 0, bk3 = 0;
        for (int r = 0; r < 4; ++r) {
          cudaMemset(over2, 0, NP / 8);
          SweepCtl init; init.tile_counter = 0; init.pad = 0; init.cursor = 0; init.stored = ~0ull;
          cudaMemcpy(ctl2, &init, sizeof init, cudaMemcpyHostToDevice); cudaMemset(gsum2, 0, 4096);
          cudaDeviceSynchronize();
          cudaEventRecord(e0, s1); cudaEventRecord(e2, s2);
          if (order == 0) sweep_kernel<true, true, 256, 4><<<148 * k1ctas, 256, 0, s1>>>(p2);
          place_group_kernel<8192><<<ggrid, kGroupThreads, 0, s2>>>(ctl, rows, p.tile_slot, p.tile_count, tfirst, p.num_tiles, ids, out_rows, out_ids, none);
          if (order == 1) sweep_kernel<true, true, 256, 4><<<148 * k1ctas, 256, 0, s1>>>(p2);
          cudaEventRecord(e1, s1); cudaEventRecord(e3, s2);
          cudaDeviceSynchronize();
          float k1, k3; cudaEventElapsedTime(&k1, e0, e1); cudaEventElapsedTime(&k3, e2, e3);
          float tot = std::max(k1, k3);
          if (tot < best) { best = tot; bk1 = k1; bk3 = k3; }
        }
        printf("K1 x%d/SM || K3g, %s first:            total %8.3f ms (K1 %.3f, K3g %.3f)\n", k1ctas, order == 0 ? "K1" : "K3g", best, bk1, bk3);
      }
    }
    {
      float best = 1e9f;
      for (int r = 0; r < 4; ++r) {
        cudaDeviceSynchronize();
        cudaEventRecord(e2, s2);
        place_group_kernel<8192><<<ggrid, kGroupThreads, 0, s2>>>(ctl, rows, p.tile_slot, p.tile_count, tfirst, p.num_tiles, ids, out_rows, out_ids, none);
        cudaEventRecord(e3, s2);
        cudaDeviceSynchronize();
        float k3; cudaEventElapsedTime(&k3, e2, e3);
        best = std::min(best, k3);
      }
      printf("K3g alone                                 total %8.3f ms\n", best);
    }
  }
  return 0;
}
